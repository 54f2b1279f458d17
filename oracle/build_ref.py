"""Build the UNMODIFIED reference CUDA extensions into oracle/_ref/ (test infrastructure only).

The reference (joejep/FOCNeRF, a torch-ngp fork) has no CPU implementation of its hot path: every
operator is a CUDA extension.  This recipe cross-compiles the reference's own sources, where they lie
under /root/reference, for sm_100a and drops the resulting torch-extension .so files into
oracle/_ref/<name>/ (git-ignored, NOT gpurun-ignored, so they travel to the B200 box).  They are then
used by tests/ (-m gpu), tests/golden/make_golden.py and bench.py's reference legs as the checker and
the timed reference; no product code path imports them.

No reference source is copied into this repository's history: nvcc reads the .cu files in place, and `copy_py()`
drops verbatim copies of the reference's own PYTHON callers / wrappers (PY_FILES below) into oracle/_ref/py/ --
git-ignored like the .so files, shipped to the GPU box with them -- so that
  * bench.py --impl reference drives the reference's unmodified Trainer / NeRFNetwork / run_cuda / L1 wrappers, and
  * tests/test_gpu_dropin.py runs the same unmodified callers on top of focnerf_b200.install_aliases().

usage: python oracle/build_ref.py [raymarching gridencoder ffmlp freqencoder py]
"""
import os
import sys

REF = os.environ.get("FOC_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")

HALF = ['-U__CUDA_NO_HALF_OPERATORS__', '-U__CUDA_NO_HALF_CONVERSIONS__', '-U__CUDA_NO_HALF2_OPERATORS__']
# flags restated from <pkg>/backend.py of the reference (raymarching/backend.py:6-9 etc.)
SPECS = {
    "raymarching": dict(name="_ref_raymarching", srcs=["raymarching.cu", "bindings.cpp"],
                        nvcc=['-O3', '-std=c++17'] + HALF, inc=[]),
    "gridencoder": dict(name="_ref_gridencoder", srcs=["gridencoder.cu", "bindings.cpp"],
                        nvcc=['-O3', '-std=c++17'] + HALF, inc=[]),
    "freqencoder": dict(name="_ref_freqencoder", srcs=["freqencoder.cu", "bindings.cpp"],
                        nvcc=['-O3', '-std=c++17', '-use_fast_math'] + HALF, inc=[]),
    "ffmlp": dict(name="_ref_ffmlp", srcs=["ffmlp.cu", "bindings.cpp"],
                  nvcc=['-O3', '-std=c++17', '--expt-extended-lambda', '--expt-relaxed-constexpr',
                        '-Xcompiler=-mf16c', '-Xcompiler=-Wno-float-conversion',
                        '-Xcompiler=-fno-strict-aliasing'] + HALF,
                  inc=["dependencies/cutlass/include", "dependencies/cutlass/tools/util/include"]),
}


def built_path(pkg):
    s = SPECS[pkg]
    return os.path.join(OUT, pkg, s["name"] + ".so")


def build(pkg, verbose=True):
    if not os.path.isdir(REF):
        raise RuntimeError(f"{REF} not present: the reference can only be compiled in the build container")
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0a"
    os.environ.setdefault("MAX_JOBS", "4")
    from torch.utils.cpp_extension import load
    s = SPECS[pkg]
    bdir = os.path.join(OUT, pkg)
    os.makedirs(bdir, exist_ok=True)
    load(name=s["name"],
         sources=[os.path.join(REF, pkg, "src", f) for f in s["srcs"]],
         extra_cflags=['-O3', '-std=c++17'],
         extra_cuda_cflags=s["nvcc"],
         extra_include_paths=[os.path.join(REF, pkg, i) for i in s["inc"]],
         build_directory=bdir, verbose=verbose, is_python_module=False)
    return built_path(pkg)


# reference python files run unmodified by the reference arm and the drop-in tests: (source under /root/reference, destination under _ref/py)
PY_FILES = [
    ("activation.py", "activation.py"),
    ("raymarching/__init__.py", "raymarching/__init__.py"), ("raymarching/raymarching.py", "raymarching/raymarching.py"),
    ("gridencoder/__init__.py", "gridencoder/__init__.py"), ("gridencoder/grid.py", "gridencoder/grid.py"),
    ("ffmlp/__init__.py", "ffmlp/__init__.py"), ("ffmlp/ffmlp.py", "ffmlp/ffmlp.py"),
    ("freqencoder/__init__.py", "freqencoder/__init__.py"), ("freqencoder/freq.py", "freqencoder/freq.py"),
    # the working cuda_ray callers (legacy/ is the copy whose run_cuda signature is intact)
    ("legacy/nerf/renderer.py", "nerf/renderer.py"), ("legacy/nerf/network_ff.py", "nerf/network_ff.py"),
    ("legacy/nerf/network.py", "nerf/network.py"), ("legacy/nerf/utils.py", "nerf/utils.py"),
    # FOC multi-object callers (script bodies: tests extract the functions they pin, see oracle/foc_extract.py)
    ("COMBINED.py", "foc/COMBINED.py"), ("editable.py", "foc/editable.py"), ("flags.py", "foc/flags.py"),
    ("nerf/renderer.py", "foc/nerf/renderer.py"), ("nerf/network_tcnn.py", "foc/nerf/network_tcnn.py"),
    ("nerf/multiobjectnetwork.py", "foc/nerf/multiobjectnetwork.py"), ("nerf/utils.py", "foc/nerf/utils.py"),
]
PY_OUT = os.path.join(OUT, "py")


def py_present():
    return all(os.path.exists(os.path.join(PY_OUT, d)) for _, d in PY_FILES)


def copy_py():
    """Verbatim copies (byte for byte) of PY_FILES into oracle/_ref/py/; empty __init__.py for the two `nerf` packages."""
    import shutil
    if not os.path.isdir(REF):
        raise RuntimeError(f"{REF} not present: the reference can only be copied in the build container")
    for src, dst in PY_FILES:
        d = os.path.join(PY_OUT, dst)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(os.path.join(REF, src), d)
    for pkg in ("nerf", "foc/nerf"):
        open(os.path.join(PY_OUT, pkg, "__init__.py"), "a").close()
    return PY_OUT


if __name__ == "__main__":
    for p in (sys.argv[1:] or list(SPECS) + ["py"]):
        print("built", copy_py() if p == "py" else build(p))
