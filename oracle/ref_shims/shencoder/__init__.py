"""Degree-<=4 real spherical-harmonics direction encoder in plain torch ops, for the REFERENCE arm only (the reference imports
`shencoder` through the absent `encoding` module; torch-ngp's is a one-kernel CUDA extension that is not in the reference tree).
Same closed form and coefficient order as torch-ngp's shencoder.cu (and as focnerf_b200/csrc/freq.cu: k_sh_fwd, which is tested
against oracle.sh4_encode).  No gradient w.r.t. the directions (the reference never asks for one on this path)."""
import torch
import torch.nn as nn


def sh_encode(d, degree=4):
    d = d.float()
    x, y, z = d[..., 0], d[..., 1], d[..., 2]
    out = [torch.full_like(x, 0.28209479177387814)]
    if degree > 1:
        out += [-0.48860251190291987 * y, 0.48860251190291987 * z, -0.48860251190291987 * x]
    if degree > 2:
        xy, xz, yz, x2, y2, z2 = x * y, x * z, y * z, x * x, y * y, z * z
        out += [1.0925484305920792 * xy, -1.0925484305920792 * yz, 0.94617469575755997 * z2 - 0.31539156525251999,
                -1.0925484305920792 * xz, 0.54627421529603959 * x2 - 0.54627421529603959 * y2]
    if degree > 3:
        out += [0.59004358992664352 * y * (-3.0 * x2 + y2), 2.8906114426405538 * xy * z, 0.45704579946446572 * y * (1.0 - 5.0 * z2),
                0.3731763325901154 * z * (5.0 * z2 - 3.0), 0.45704579946446572 * x * (1.0 - 5.0 * z2),
                1.4453057213202769 * z * (x2 - y2), 0.59004358992664352 * x * (-x2 + 3.0 * y2)]
    return torch.stack(out, -1)


class SHEncoder(nn.Module):
    def __init__(self, input_dim=3, degree=4):
        super().__init__()
        assert input_dim == 3 and 1 <= degree <= 4
        self.input_dim, self.degree, self.output_dim = input_dim, degree, degree ** 2

    def __repr__(self):
        return f"SHEncoder: input_dim={self.input_dim} degree={self.degree}"

    def forward(self, inputs, size=1):
        with torch.no_grad():
            return sh_encode(inputs / size, self.degree)
