"""Import stub (see ../README.md): `import matplotlib.pyplot as plt` in nerf/utils.py; plotting is off the per-ray path."""


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    raise RuntimeError("stub matplotlib.pyplot: '%s' requested" % name)
