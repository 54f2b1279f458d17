from .grid import GridEncoder, grid_encode, _backend  # noqa: F401
