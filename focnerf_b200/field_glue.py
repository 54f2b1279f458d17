"""Autograd Functions over the fused field-network wiring kernels (csrc/field_glue.cu): the elementwise steps of
nerf/network_ff.py:51-74 between encoder, sigma net, colour net and compositor."""
import torch
from torch.autograd import Function

from ._lib import check, lib, ptr, require_cuda, stream, valid_rows


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


class _field_eval(Function):
    """(enc [B,32] f16, dirs [B,3] f32) -> (sigma [B] f32, rgb [B,3] f32): sigma net -> trunc_exp / SH-4 / cat -> colour net -> sigmoid as ONE
    kernel (csrc/field_fwd.cu), and a backward of two kernels whose prologues form the output gradients themselves
    (foc_ffmlp_backward_head): no h / cin / hc / grad_h / grad_hc tensors, no glue launches.  Needs both FFMLPs' fp16 side buffers
    (FFMLP.enable_fp16_side_buffers): the weights are read from the shadows and the weight gradients accumulate into grad16.
    `ws`, `wc` (the fp32 Parameters) only anchor the autograd graph."""

    @staticmethod
    def forward(ctx, enc, dirs, ws, wc, sn, cn, train):
        require_cuda(enc, dirs)
        enc = enc.contiguous()
        if enc.dtype != torch.float16:
            enc = enc.to(torch.float16)
        dirs = dirs.to(torch.float32).contiguous()
        B, dev = enc.shape[0], enc.device
        sigma = torch.empty(B, dtype=torch.float32, device=dev)
        rgb = torch.empty(B, 3, dtype=torch.float32, device=dev)
        ss, sc = sn._current_shadow(), cn._current_shadow()
        if train:
            fb_s = torch.empty(sn.num_layers, B, 64, dtype=torch.float16, device=dev)
            fb_c = torch.empty(cn.num_layers, B, 64, dtype=torch.float16, device=dev)
            cin = torch.empty(B, 32, dtype=torch.float16, device=dev)
        else:
            fb_s = fb_c = cin = None
        rows = valid_rows()
        ctx.rows = rows
        check(lib().foc_field_forward(ptr(enc), ptr(dirs), ptr(ss), ptr(sc), B, sn.num_layers, cn.num_layers, ptr(fb_s), ptr(fb_c), ptr(cin),
                                      ptr(sigma), ptr(rgb), ptr(rows), stream()), "field_forward")
        if train:
            ctx.save_for_backward(enc, cin, fb_s, fb_c, sigma, rgb, ss, sc)
            ctx.nets = (sn, cn)
        return sigma, rgb

    @staticmethod
    def backward(ctx, g_sigma, g_rgb):
        enc, cin, fb_s, fb_c, sigma, rgb, ss, sc = ctx.saved_tensors
        sn, cn = ctx.nets
        B, dev, l, st = enc.shape[0], enc.device, lib(), stream()
        g_rgb = torch.zeros_like(rgb) if g_rgb is None else g_rgb.to(torch.float32).contiguous()
        g_sigma = None if g_sigma is None else g_sigma.to(torch.float32).contiguous()
        g_cin = torch.empty(B, 32, dtype=torch.float16, device=dev)
        scr = torch.empty(l.foc_ffmlp_backward_scratch_bytes(B, 32, 64, cn.num_layers), dtype=torch.uint8, device=dev)
        check(l.foc_ffmlp_backward_head(1, ptr(g_rgb), ptr(rgb), None, ptr(cin), ptr(sc), ptr(fb_c), B, 32, 64, cn.num_layers, 1, ptr(g_cin),
                                        ptr(cn._grad16), 1, ptr(ctx.rows), ptr(scr), st), "ffmlp_backward_head(colour)")
        g_enc = torch.empty(B, 32, dtype=torch.float16, device=dev)
        scr2 = torch.empty(l.foc_ffmlp_backward_scratch_bytes(B, 32, 64, sn.num_layers), dtype=torch.uint8, device=dev)
        check(l.foc_ffmlp_backward_head(2, ptr(g_sigma), ptr(sigma), ptr(g_cin), ptr(enc), ptr(ss), ptr(fb_s), B, 32, 64, sn.num_layers, 1, ptr(g_enc),
                                        ptr(sn._grad16), 1, ptr(ctx.rows), ptr(scr2), st), "ffmlp_backward_head(sigma)")
        return g_enc, None, None, None, None, None, None


def field_eval(enc, dirs, sigma_net, color_net, train):
    return _field_eval.apply(enc, dirs, sigma_net.weights, color_net.weights, sigma_net, color_net, bool(train))
