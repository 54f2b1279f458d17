"""Host-side Morton helpers used only to BUILD synthetic occupancy grids (focnerf_b200/synthetic.py).
The operator `raymarching.morton3D` is the CUDA kernel in csrc/raymarch.cu; this is not a fallback for it."""
import torch


def _spread3(v):
    v = v.to(torch.int64)
    v = (v * 0x00010001) & 0xFF0000FF
    v = (v * 0x00000101) & 0x0F00F00F
    v = (v * 0x00000011) & 0xC30C30C3
    v = (v * 0x00000005) & 0x49249249
    return v


def morton3d_torch(coords):
    """coords [n,3] ints in [0,1024) -> int64 Morton codes (x lowest bit), as raymarching.cu:56-71."""
    return _spread3(coords[:, 0]) | (_spread3(coords[:, 1]) << 1) | (_spread3(coords[:, 2]) << 2)
