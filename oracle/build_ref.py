"""Build the UNMODIFIED reference CUDA extensions into oracle/_ref/ (test infrastructure only).

The reference (joejep/FOCNeRF, a torch-ngp fork) has no CPU implementation of its hot path: every
operator is a CUDA extension.  This recipe cross-compiles the reference's own sources, where they lie
under /root/reference, for sm_100a and drops the resulting torch-extension .so files into
oracle/_ref/<name>/ (git-ignored, NOT gpurun-ignored, so they travel to the B200 box).  They are then
used by tests/ (-m gpu), tests/golden/make_golden.py and bench.py's reference legs as the checker and
the timed reference; no product code path imports them.

No reference source is copied into this repository: nvcc reads them in place.

usage: python oracle/build_ref.py [raymarching gridencoder ffmlp freqencoder]
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


if __name__ == "__main__":
    for p in (sys.argv[1:] or list(SPECS)):
        print("built", build(p))
