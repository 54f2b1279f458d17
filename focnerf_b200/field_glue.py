"""Autograd Functions over the fused field-network wiring kernels (csrc/field_glue.cu): the elementwise steps of
nerf/network_ff.py:51-74 between encoder, sigma net, colour net and compositor."""
import torch
from torch.autograd import Function

from ._lib import check, lib, ptr, require_cuda, stream


class _field_mid(Function):
    """(h [B,16] f16, dirs [B,3] f32) -> (sigma [B] f32, cin [B,32] f16)"""

    @staticmethod
    def forward(ctx, h, dirs):
        require_cuda(h, dirs)
        h, dirs = h.contiguous(), dirs.to(torch.float32).contiguous()
        B = h.shape[0]
        sigma = torch.empty(B, dtype=torch.float32, device=h.device)
        cin = torch.empty(B, 32, dtype=torch.float16, device=h.device)
        check(lib().foc_field_mid_forward(ptr(dirs), ptr(h), B, ptr(sigma), ptr(cin), stream()), "field_mid_forward")
        ctx.save_for_backward(h)
        return sigma, cin

    @staticmethod
    def backward(ctx, grad_sigma, grad_cin):
        (h,) = ctx.saved_tensors
        B = h.shape[0]
        grad_h = torch.empty_like(h)
        gs = None if grad_sigma is None else grad_sigma.to(torch.float32).contiguous()
        gc = torch.zeros(B, 32, dtype=torch.float16, device=h.device) if grad_cin is None else grad_cin.to(torch.float16).contiguous()
        check(lib().foc_field_mid_backward(ptr(gs), ptr(gc), ptr(h), B, ptr(grad_h), stream()), "field_mid_backward")
        return grad_h, None


class _rgb_sigmoid(Function):
    """hc [B,16] f16 -> rgb [B,3] f32"""

    @staticmethod
    def forward(ctx, hc):
        require_cuda(hc)
        hc = hc.contiguous()
        B = hc.shape[0]
        rgb = torch.empty(B, 3, dtype=torch.float32, device=hc.device)
        check(lib().foc_rgb_sigmoid_forward(ptr(hc), B, ptr(rgb), stream()), "rgb_sigmoid_forward")
        ctx.save_for_backward(rgb)
        return rgb

    @staticmethod
    def backward(ctx, grad_rgb):
        (rgb,) = ctx.saved_tensors
        B = rgb.shape[0]
        grad_hc = torch.empty(B, 16, dtype=torch.float16, device=rgb.device)
        check(lib().foc_rgb_sigmoid_backward(ptr(grad_rgb.to(torch.float32).contiguous()), ptr(rgb), B, ptr(grad_hc), stream()),
              "rgb_sigmoid_backward")
        return grad_hc


field_mid = _field_mid.apply
rgb_sigmoid = _rgb_sigmoid.apply
