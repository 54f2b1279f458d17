"""Make the reference's UNMODIFIED Python files (oracle/_ref/py/, copied verbatim by oracle/build_ref.py) importable.
Test / bench infrastructure only -- never imported by the product.

Two ways of standing them up, one per process:

  activate_reference()   the REFERENCE arm: `_raymarching`, `_gridencoder`, `_ffmlp`, `_freqencoder` resolve to the reference's own
                         CUDA extensions (oracle/_ref/*/_ref_*.so, unmodified sources built for sm_100a), `raymarching`, `gridencoder`,
                         `ffmlp`, `freqencoder`, `activation`, `nerf.*` to the reference's own Python, `encoding` / `shencoder` (absent
                         from the reference tree) to oracle/ref_shims (plain torch).  Nothing from focnerf_b200 is imported.
  activate_callers()     the DROP-IN check: focnerf_b200.install_aliases() first, so the reference's callers (`nerf.renderer`,
                         `nerf.network_ff`, `nerf.utils.Trainer`) import THIS repo's operator packages -- "callers run unchanged".

`load_l1_wrapper(pkg, backend)` loads one of the reference's L1 wrapper files (raymarching.py, grid.py, ffmlp.py, freq.py) under a
private module name with `_<pkg>` bound to `backend`: the second drop-in level (reference wrappers over this repo's C-ABI shim).
"""
import importlib
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PY = os.path.join(HERE, "_ref", "py")
STUBS = os.path.join(HERE, "stubs")
SHIMS = os.path.join(HERE, "ref_shims")
_EXT = ("raymarching", "gridencoder", "ffmlp", "freqencoder")
_state = {"mode": None}


def present():
    return os.path.exists(os.path.join(PY, "nerf", "renderer.py"))


def _need(mode):
    if not present():
        raise RuntimeError("oracle/_ref/py is missing: run `python oracle/build_ref.py py` in the build container")
    if _state["mode"] not in (None, mode):
        raise RuntimeError(f"oracle.ref_py already activated as '{_state['mode']}' in this process; '{mode}' needs its own process")
    _state["mode"] = mode


def _load_ext(pkg):
    """oracle/_ref/<pkg>/_ref_<pkg>.so -> module (its PyInit name is _ref_<pkg>)."""
    import torch  # noqa: F401  (the extensions link against libtorch)
    name = f"_ref_{pkg}"
    if name in sys.modules:
        return sys.modules[name]
    path = os.path.join(HERE, "_ref", pkg, name + ".so")
    if not os.path.exists(path):
        raise RuntimeError(f"{path} missing: run `python oracle/build_ref.py {pkg}` in the build container")
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sys.modules[name] = mod
    return mod


def _paths(*dirs):
    for d in reversed(dirs):
        if d in sys.path:
            sys.path.remove(d)
        sys.path.insert(0, d)


def activate_reference():
    _need("reference")
    for pkg in _EXT:
        sys.modules[f"_{pkg}"] = _load_ext(pkg)            # what `import _raymarching as _backend` finds (raymarching.py:9)
    _paths(PY, SHIMS, STUBS)
    return PY


def activate_callers():
    _need("callers")
    import focnerf_b200
    focnerf_b200.install_aliases()
    _paths(PY, STUBS)                                      # the aliases sit in sys.modules, so only `nerf.*` comes from PY
    return PY


def load_l1_wrapper(pkg, backend):
    """-> the reference's L1 wrapper module for `pkg` (e.g. raymarching/raymarching.py), unmodified, with `backend` as its `_backend`."""
    if not present():
        raise RuntimeError("oracle/_ref/py is missing")
    fname = {"raymarching": "raymarching.py", "gridencoder": "grid.py", "ffmlp": "ffmlp.py", "freqencoder": "freq.py"}[pkg]
    if STUBS not in sys.path:
        sys.path.append(STUBS)                             # ffmlp.py:2 imports `turtle`
    key, saved = f"_{pkg}", sys.modules.get(f"_{pkg}")
    sys.modules[key] = backend
    try:
        name = f"ref_l1_{pkg}_{id(backend):x}"
        spec = importlib.util.spec_from_file_location(name, os.path.join(PY, pkg, fname))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        if saved is None:
            sys.modules.pop(key, None)
        else:
            sys.modules[key] = saved
    return mod
