"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/foc_b200.h declares."""
import ctypes
import os

import pytest


def test_library_builds_and_exports_every_declared_symbol(built_lib):
    from focnerf_b200 import _lib
    assert os.path.exists(built_lib)
    names = _lib.declared_symbols()
    assert len(names) >= 25
    l = ctypes.CDLL(built_lib)
    missing = [n for n in names if not hasattr(l, n)]
    assert not missing, missing
    assert _lib.lib().foc_abi_version() == 1           # also sets argtypes from the header for every entry point


def test_header_prototypes_parse_to_ctypes(built_lib):
    from focnerf_b200 import _lib
    protos = _lib.parse_header()
    ret, args = protos["foc_march_rays_train"]
    assert ret is ctypes.c_int and len(args) == 20 and args[3] is ctypes.c_float and args[5] is ctypes.c_uint32
    assert protos["foc_ffmlp_backward_scratch_bytes"][0] is ctypes.c_size_t
    assert len(protos["foc_march_rays_train_scratch_bytes"][1]) == 2


def test_sass_contains_blackwell_tensor_path(built_lib):
    """The fused MLP must be tcgen05 / TMEM / TMA code, not mma.sync: SASS mnemonics of B200_PROFILING.md."""
    import shutil
    import subprocess
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", built_lib], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass and "LDTM" in sass and "UTMALDG" in sass
    assert "HMMA." not in sass.replace("UTCHMMA", "")


def test_operators_fail_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from focnerf_b200 import raymarching
    with pytest.raises(RuntimeError):
        raymarching.near_far_from_aabb(torch.zeros(4, 3), torch.ones(4, 3), torch.tensor([-1., -1, -1, 1, 1, 1]))
    from focnerf_b200.trainer import Trainer
    with pytest.raises(RuntimeError):
        Trainer()


def test_product_code_never_imports_the_oracle():
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "focnerf_b200")
    for dp, _, fs in os.walk(root):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "liborc" not in src, os.path.join(dp, f)
