"""Import stub (see README.md): `trimesh` is imported by the reference's callers at module scope and never used on the per-ray path."""


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    raise RuntimeError("stub module 'trimesh' (oracle/stubs): the reference's per-ray path never calls into it; '%s' was requested" % name)
