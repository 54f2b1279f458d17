"""Import stub (see README.md): `mcubes` is imported by the reference's callers at module scope and never used on the per-ray path."""


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    raise RuntimeError("stub module 'mcubes' (oracle/stubs): the reference's per-ray path never calls into it; '%s' was requested" % name)
