"""`freqencoder.FreqEncoder` / `freq_encode` of the reference (freqencoder/freq.py:15-76) over libfoc_b200."""
import torch
import torch.nn as nn
from torch.autograd import Function
from torch.amp import custom_bwd, custom_fwd

from .._lib import check, lib, ptr, require_cuda, stream


class _Backend:
    """Reference pybind names (freqencoder/src/bindings.cpp)."""

    @staticmethod
    def freq_encode_forward(inputs, B, D, deg, C, outputs):
        require_cuda(inputs, outputs)
        if inputs.dtype != torch.float32 or not inputs.is_contiguous():
            raise RuntimeError("inputs must be a contiguous float32 tensor")
        check(lib().foc_freq_encode_forward(ptr(inputs), B, D, deg, C, ptr(outputs), stream()), "freq_encode_forward")

    @staticmethod
    def freq_encode_backward(grad, outputs, B, D, deg, C, grad_inputs):
        require_cuda(grad, outputs, grad_inputs)
        if grad.dtype != torch.float32 or not grad.is_contiguous():
            raise RuntimeError("grad must be a contiguous float32 tensor")
        check(lib().foc_freq_encode_backward(ptr(grad), ptr(outputs), B, D, deg, C, ptr(grad_inputs), stream()), "freq_encode_backward")


_backend = _Backend()


class _freq_encoder(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, inputs, degree, output_dim):
        if not inputs.is_cuda:
            inputs = inputs.cuda()
        inputs = inputs.contiguous()
        B, input_dim = inputs.shape
        outputs = torch.empty(B, output_dim, dtype=inputs.dtype, device=inputs.device)
        _backend.freq_encode_forward(inputs, B, input_dim, degree, output_dim, outputs)
        ctx.save_for_backward(inputs, outputs)
        ctx.dims = [B, input_dim, degree, output_dim]
        return outputs

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, grad):
        grad = grad.contiguous()
        inputs, outputs = ctx.saved_tensors
        B, input_dim, degree, output_dim = ctx.dims
        grad_inputs = torch.zeros_like(inputs)
        _backend.freq_encode_backward(grad, outputs, B, input_dim, degree, output_dim, grad_inputs)
        return grad_inputs, None, None


freq_encode = _freq_encoder.apply


class FreqEncoder(nn.Module):
    def __init__(self, input_dim=3, degree=4):
        super().__init__()
        self.input_dim = input_dim
        self.degree = degree
        self.output_dim = input_dim + input_dim * 2 * degree

    def __repr__(self):
        return f"FreqEncoder: input_dim={self.input_dim} degree={self.degree} output_dim={self.output_dim}"

    def forward(self, inputs, **kwargs):
        prefix_shape = list(inputs.shape[:-1])
        inputs = inputs.reshape(-1, self.input_dim)
        outputs = freq_encode(inputs, self.degree, self.output_dim)
        return outputs.reshape(prefix_shape + [self.output_dim])
