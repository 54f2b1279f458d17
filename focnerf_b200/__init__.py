"""focnerf_b200 -- B200-native (sm_100a) implementation of FOC-NeRF's per-ray rendering hot path.

Sub-packages mirror the reference's operator packages one to one (`raymarching`, `gridencoder`, `ffmlp`,
`freqencoder`, plus the `shencoder` / `encoding` / `activation` modules its networks import); every operator is a
hand-written CUDA kernel in libfoc_b200.so reached through the C ABI of include/foc_b200.h.  There is no CPU path.
"""
import importlib
import sys

_ALIASES = ("raymarching", "gridencoder", "ffmlp", "freqencoder", "shencoder", "activation", "encoding")


def install_aliases():
    """Make `import raymarching`, `import gridencoder`, ... resolve to this package's modules, so the reference's
    callers (nerf/renderer.py, nerf/network_ff.py, COMBINED.py, editable.py) run unchanged on top of it."""
    for name in _ALIASES:
        sys.modules[name] = importlib.import_module(f"{__name__}.{name}")
    if "tinycudann" not in sys.modules:             # a real tiny-cuda-nn, when installed, is left alone
        sys.modules["tinycudann"] = importlib.import_module(f"{__name__}.tinycudann")
