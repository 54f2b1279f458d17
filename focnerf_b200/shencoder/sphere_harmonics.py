"""`shencoder.SHEncoder` -- the torch-ngp module nerf/network_ff.py:5 reaches through `encoding.get_encoder`.
Its source is ABSENT from the reference tree, so this is the public closed form (real SH, degree <= 4),
forward only: view directions never require gradients on the FOC hot path.  Parity unpinned for this op."""
import torch
import torch.nn as nn

from .._lib import check, lib, ptr, require_cuda, stream


def sh_encode(dirs, degree=4):
    require_cuda(dirs)
    d = dirs.to(torch.float32).contiguous().view(-1, 3)
    out = torch.empty(d.shape[0], degree * degree, dtype=torch.float32, device=d.device)
    check(lib().foc_sh_encode_forward(ptr(d), d.shape[0], degree, ptr(out), stream()), "sh_encode_forward")
    return out


class SHEncoder(nn.Module):
    def __init__(self, input_dim=3, degree=4):
        super().__init__()
        assert input_dim == 3, "SH encoder only support input dim == 3"
        assert 1 <= degree <= 4, "the sm_100a SH encoder implements degree 1..4"
        self.input_dim, self.degree, self.output_dim = input_dim, degree, degree ** 2

    def __repr__(self):
        return f"SHEncoder: input_dim={self.input_dim} degree={self.degree}"

    def forward(self, inputs, size=1):
        prefix = list(inputs.shape[:-1])
        return sh_encode(inputs / size, self.degree).view(prefix + [self.output_dim])
