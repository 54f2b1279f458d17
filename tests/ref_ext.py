"""Load the UNMODIFIED reference CUDA extensions from oracle/_ref/ (built by oracle/build_ref.py from /root/reference,
git-ignored but shipped to the GPU box).  Test / bench infrastructure only.  Returns None when a module is absent so
tests can skip; nothing here reads /root/reference at run time."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_REF = os.path.join(ROOT, "oracle", "_ref")
_cache = {}


def load(pkg):
    """pkg in {raymarching, gridencoder, ffmlp, freqencoder} -> pybind module or None."""
    if pkg in _cache:
        return _cache[pkg]
    import torch  # noqa: F401  (the extensions link against libtorch)
    name = f"_ref_{pkg}"
    path = os.path.join(_REF, pkg, name + ".so")
    mod = None
    if os.path.exists(path) and torch.cuda.is_available():
        try:
            spec = importlib.util.spec_from_file_location(name, path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            sys.modules[name] = mod
        except Exception as e:  # pragma: no cover
            print(f"[ref_ext] could not load {path}: {e}")
            mod = None
    _cache[pkg] = mod
    return mod
