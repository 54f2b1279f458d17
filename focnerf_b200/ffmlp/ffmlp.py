"""`ffmlp.FFMLP` and `ffmlp_forward` of the reference (ffmlp/ffmlp.py:15-169) over the tcgen05/TMEM kernels.

Kept: constructor + assertions, the flat fp32 `weights` Parameter and its layout ([hidden,in] + (L-1)x[hidden,hidden]
+ [16,hidden], ffmlp.cu:632), `reset_parameters` (reseeds the global RNG with 42, ffmlp.py:141-144), the batch padding
that always appends 1..128 zero rows (ffmlp.py:157-159), `custom_fwd(cast_inputs=torch.half)`, `__repr__`, and the
`_backend` method names (ffmlp/src/bindings.cpp:5-11).  Not imported: `turtle` (ffmlp.py:2, needs tkinter).

Differences, on purpose: `backward_buffer` is not materialised (the kernel keeps dL/dact in shared memory and reduces
the weight gradients itself); `allocate_splitk/free_splitk` are no-ops.  hidden_dim == 64 with input_dim in {32, 64} (the FOC
field networks) runs on the tcgen05 kernels; every other shape the constructor accepts (hidden 16 / 32 / 128 / 256, input_dim = 16 m)
runs on the CUDA-core kernels of csrc/mlp_generic.cu, which do write `backward_buffer`.
"""
import math

import torch
import torch.nn as nn
from torch.autograd import Function
from torch.amp import custom_bwd, custom_fwd

from .._lib import check, lib, ptr, require_cuda, stream


def _chk_half(t, name):
    if t is None:
        return
    if not t.is_cuda or not t.is_contiguous() or t.dtype != torch.float16:
        raise RuntimeError(f"{name} must be a contiguous CUDA half tensor")


class _Backend:
    @staticmethod
    def ffmlp_forward(inputs, weights, B, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                      forward_buffer, outputs):
        _chk_half(inputs, "inputs"); _chk_half(weights, "weights"); _chk_half(forward_buffer, "forward_buffer"); _chk_half(outputs, "outputs")
        check(lib().foc_ffmlp_forward(ptr(inputs), ptr(weights), B, input_dim, output_dim, hidden_dim, num_layers, activation,
                                      output_activation, ptr(forward_buffer), ptr(outputs), stream()), "ffmlp_forward")

    @staticmethod
    def ffmlp_inference(inputs, weights, B, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                        inference_buffer, outputs):
        # inference_buffer ([B, hidden], ffmlp.py:40) is accepted for signature compatibility and left untouched:
        # activations stay in shared memory / TMEM.
        _chk_half(inputs, "inputs"); _chk_half(weights, "weights"); _chk_half(outputs, "outputs")
        check(lib().foc_ffmlp_forward(ptr(inputs), ptr(weights), B, input_dim, output_dim, hidden_dim, num_layers, activation,
                                      output_activation, None, ptr(outputs), stream()), "ffmlp_inference")

    @staticmethod
    def ffmlp_backward(grad, inputs, weights, forward_buffer, B, input_dim, output_dim, hidden_dim, num_layers, activation,
                       output_activation, calc_grad_inputs, backward_buffer, grad_inputs, grad_weights):
        _chk_half(grad, "grad"); _chk_half(inputs, "inputs"); _chk_half(weights, "weights"); _chk_half(forward_buffer, "forward_buffer")
        _chk_half(backward_buffer, "backward_buffer"); _chk_half(grad_weights, "grad_weights")
        if calc_grad_inputs:
            _chk_half(grad_inputs, "grad_inputs")
        nbytes = lib().foc_ffmlp_backward_scratch_bytes(B, input_dim, hidden_dim, num_layers)
        if nbytes == 0:
            raise RuntimeError("ffmlp_backward: unsupported configuration for the sm_100a kernels")
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=grad.device)
        check(lib().foc_ffmlp_backward(ptr(grad), ptr(inputs), ptr(weights), ptr(forward_buffer), B, input_dim, output_dim,
                                       hidden_dim, num_layers, activation, output_activation, int(bool(calc_grad_inputs)),
                                       ptr(backward_buffer), ptr(grad_inputs) if calc_grad_inputs else None, ptr(grad_weights),
                                       ptr(scratch), stream()), "ffmlp_backward")

    @staticmethod
    def ffmlp_backward_partial(grad, inputs, weights, forward_buffer, B, input_dim, output_dim, hidden_dim, num_layers, activation,
                               output_activation, calc_grad_inputs, grad_inputs, scratch):
        """ffmlp_backward without the final reduction: the per-CTA fp32 partial weight gradients stay in `scratch`."""
        _chk_half(grad, "grad"); _chk_half(inputs, "inputs"); _chk_half(weights, "weights"); _chk_half(forward_buffer, "forward_buffer")
        check(lib().foc_ffmlp_backward(ptr(grad), ptr(inputs), ptr(weights), ptr(forward_buffer), B, input_dim, output_dim,
                                       hidden_dim, num_layers, activation, output_activation, int(bool(calc_grad_inputs)),
                                       None, ptr(grad_inputs) if calc_grad_inputs else None, None, ptr(scratch), stream()), "ffmlp_backward")

    @staticmethod
    def ffmlp_wgrad_finish(scratch, B, input_dim, hidden_dim, num_layers, grad_weights):
        _chk_half(grad_weights, "grad_weights")
        check(lib().foc_ffmlp_wgrad_finish(ptr(scratch), B, input_dim, hidden_dim, num_layers, ptr(grad_weights), stream()), "ffmlp_wgrad_finish")

    @staticmethod
    def ffmlp_wgrad_accumulate(scratch, B, input_dim, hidden_dim, num_layers, grad_weights):
        _chk_half(grad_weights, "grad_weights")
        check(lib().foc_ffmlp_wgrad_accumulate(ptr(scratch), B, input_dim, hidden_dim, num_layers, ptr(grad_weights), stream()), "ffmlp_wgrad_accumulate")

    @staticmethod
    def allocate_splitk(size):
        check(lib().foc_allocate_splitk(size), "allocate_splitk")

    @staticmethod
    def free_splitk():
        check(lib().foc_free_splitk(), "free_splitk")


_backend = _Backend()


class _ffmlp_forward(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.half)
    def forward(ctx, inputs, weights, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                inference=False, calc_grad_inputs=False):
        require_cuda(inputs, weights)
        B = inputs.shape[0]
        inputs, weights = inputs.contiguous(), weights.contiguous()
        outputs = torch.empty(B, output_dim, device=inputs.device, dtype=inputs.dtype)
        if not inference:
            forward_buffer = torch.empty(num_layers, B, hidden_dim, device=inputs.device, dtype=inputs.dtype)
            _backend.ffmlp_forward(inputs, weights, B, input_dim, output_dim, hidden_dim, num_layers, activation,
                                   output_activation, forward_buffer, outputs)
            ctx.save_for_backward(inputs, weights, outputs, forward_buffer)
            ctx.dims = (input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, calc_grad_inputs)
        else:
            _backend.ffmlp_inference(inputs, weights, B, input_dim, output_dim, hidden_dim, num_layers, activation,
                                     output_activation, None, outputs)
        return outputs

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, grad):
        B = grad.shape[0]
        grad = grad.contiguous()
        inputs, weights, outputs, forward_buffer = ctx.saved_tensors
        input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, calc_grad_inputs = ctx.dims
        grad_inputs = torch.empty_like(inputs) if calc_grad_inputs else None
        grad_weights = torch.empty_like(weights)              # fully overwritten by the reduction kernel
        _backend.ffmlp_backward(grad, inputs, weights, forward_buffer, B, input_dim, output_dim, hidden_dim, num_layers,
                                activation, output_activation, calc_grad_inputs, None, grad_inputs, grad_weights)
        return (grad_inputs if calc_grad_inputs else None), grad_weights, None, None, None, None, None, None, None, None


ffmlp_forward = _ffmlp_forward.apply


class _ffmlp_forward_side(Function):
    """Not in the reference: the same kernels with the weights' fp16 side buffers (FFMLP.enable_fp16_side_buffers).  `shadow16`
    is the fp16 copy of the weights the optimiser keeps current (no per-step cast of the fp32 Parameter); the weight gradient is
    written straight into `grad16` (no half -> float copy, no AccumulateGrad).  `weights` (the fp32 Parameter) only anchors the
    autograd graph; its gradient is returned as None."""

    @staticmethod
    def forward(ctx, inputs, weights, shadow16, grad16, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                inference=False, calc_grad_inputs=False, side=None):
        require_cuda(inputs, shadow16)
        ctx.side = side
        if inputs.dtype != torch.float16:
            inputs = inputs.to(torch.float16)
        inputs = inputs.contiguous()
        B = inputs.shape[0]
        outputs = torch.empty(B, output_dim, device=inputs.device, dtype=torch.float16)
        if not inference:
            forward_buffer = torch.empty(num_layers, B, hidden_dim, device=inputs.device, dtype=torch.float16)
            _backend.ffmlp_forward(inputs, shadow16, B, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                                   forward_buffer, outputs)
            ctx.save_for_backward(inputs, shadow16, forward_buffer)
            ctx.grad16 = grad16
            ctx.dims = (input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, calc_grad_inputs)
        else:
            _backend.ffmlp_inference(inputs, shadow16, B, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                                     None, outputs)
        return outputs

    @staticmethod
    def backward(ctx, grad):
        B = grad.shape[0]
        grad = grad.contiguous()
        if grad.dtype != torch.float16:
            grad = grad.to(torch.float16)
        inputs, shadow16, forward_buffer = ctx.saved_tensors
        input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, calc_grad_inputs = ctx.dims
        grad_inputs = torch.empty_like(inputs) if calc_grad_inputs else None
        if ctx.side is None:
            # grad16 ACCUMULATES like `.grad` (the optimiser sweep zeroes it after use): two field evaluations before one step add up
            scratch = torch.empty(lib().foc_ffmlp_backward_scratch_bytes(B, input_dim, hidden_dim, num_layers), dtype=torch.uint8, device=grad.device)
            _backend.ffmlp_backward_partial(grad, inputs, shadow16, forward_buffer, B, input_dim, output_dim, hidden_dim, num_layers,
                                            activation, output_activation, calc_grad_inputs, grad_inputs, scratch)
            _backend.ffmlp_wgrad_accumulate(scratch, B, input_dim, hidden_dim, num_layers, ctx.grad16)
        else:
            # nothing but the optimiser reads the weight gradient: its reduction runs on a side stream, beside the rest of the backward
            # pass (the reference ran its split-K weight-gradient GEMMs on side streams as well, ffmlp.cu:801-812)
            scratch = ctx.side.scratch(id(ctx.grad16), lib().foc_ffmlp_backward_scratch_bytes(B, input_dim, hidden_dim, num_layers), grad.device)
            _backend.ffmlp_backward_partial(grad, inputs, shadow16, forward_buffer, B, input_dim, output_dim, hidden_dim, num_layers,
                                            activation, output_activation, calc_grad_inputs, grad_inputs, scratch)
            with ctx.side.fork():
                _backend.ffmlp_wgrad_accumulate(scratch, B, input_dim, hidden_dim, num_layers, ctx.grad16)
        return (grad_inputs if calc_grad_inputs else None), None, None, None, None, None, None, None, None, None, None, None, None


def convert_activation(act):
    return {'relu': 0, 'exponential': 1, 'sine': 2, 'sigmoid': 3, 'squareplus': 4, 'softplus': 5}.get(act, 6)


class WgradSide:
    """A side stream for work nothing on the step's critical path waits for (the MLP weight-gradient reductions).  fork() orders
    the side stream after everything enqueued so far on the current stream; join() (before the optimiser) orders the current stream
    after the side work.  Scratch buffers are persistent per user, so no allocation made on one stream is recycled under the other.
    Plain stream / event dependencies: capturable in a CUDA graph."""

    def __init__(self):
        self.stream = torch.cuda.Stream()            # created up front: never inside a graph capture
        self._scratch = {}
        self._dirty = False

    def scratch(self, key, nbytes, device):
        t = self._scratch.get(key)
        if t is None or t.numel() < nbytes:
            t = self._scratch[key] = torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)
        return t

    def fork(self):
        self.stream.wait_stream(torch.cuda.current_stream())
        self._dirty = True
        return torch.cuda.stream(self.stream)

    def join(self):
        if self._dirty:
            torch.cuda.current_stream().wait_stream(self.stream)
            self._dirty = False


class FFMLP(nn.Module):
    def __init__(self, input_dim, output_dim, hidden_dim, num_layers, activation='relu'):
        super().__init__()
        self.input_dim = input_dim
        self.output_dim = output_dim
        self.hidden_dim = hidden_dim
        self.num_layers = num_layers
        self.activation = convert_activation(activation)
        self.output_activation = convert_activation('none')     # not supported by the reference either (ffmlp.py:108)
        self.tensorcore_width = 16

        assert hidden_dim in [16, 32, 64, 128, 256], f"FFMLP only support hidden_dim in [16, 32, 64, 128, 256], but got {hidden_dim}"
        assert input_dim > 0 and input_dim % 16 == 0, f"FFMLP input_dim should be 16 * m (m  > 0), but got {input_dim}"
        assert output_dim <= 16, f"FFMLP current only supports output dim <= 16, but got {output_dim}"
        assert num_layers >= 2, f"FFMLP num_layers should be larger than 2 (3 matmuls), but got {num_layers}"

        self.padded_output_dim = int(math.ceil(output_dim / 16)) * 16
        self.num_parameters = hidden_dim * (input_dim + hidden_dim * (num_layers - 1) + self.padded_output_dim)
        self.weights = nn.Parameter(torch.zeros(self.num_parameters))
        self.reset_parameters()
        _backend.allocate_splitk(self.num_layers + 1)

    def cleanup(self):
        _backend.free_splitk()

    def enable_fp16_side_buffers(self, grad16=None, shadow16=None):
        """Not in the reference.  (grad16, shadow16), both flat half tensors of the weights' size: `call_fused` then reads the
        weights from `shadow16` and writes their gradient into `grad16`; `weights.grad` stays None.  The owner (an optimiser such
        as focnerf_b200.optim.ScaledAdam, via attach()) must refresh shadow16 after every update."""
        for t in (grad16, shadow16):
            if t is not None and (t.dtype != torch.half or t.numel() != self.weights.numel() or not t.is_contiguous()):
                raise RuntimeError("fp16 side buffers must be contiguous half tensors of the weights' size")
        if shadow16 is None:
            shadow16 = torch.empty_like(self.weights, dtype=torch.half)
        shadow16.copy_(self.weights.detach())
        if grad16 is None:
            grad16 = torch.empty_like(shadow16)
        grad16.zero_()                                          # the backward accumulates into it
        self._shadow16, self._grad16 = shadow16, grad16
        self._shadow_version = self.weights._version
        return grad16, shadow16

    def _current_shadow(self):
        """See GridEncoder._current_shadow: refresh after an out-of-band write of the fp32 Parameter (load_state_dict, ...)."""
        if self.weights._version != self._shadow_version:
            with torch.no_grad():
                self._shadow16.copy_(self.weights.detach())
            self._shadow_version = self.weights._version
        return self._shadow16

    def call_fused(self, inputs, calc_grad_inputs):
        """The kernels on a batch that is already a multiple of 128 rows (no padding cat), through the side buffers when enabled."""
        if getattr(self, "_shadow16", None) is not None and torch.is_autocast_enabled():
            return _ffmlp_forward_side.apply(inputs, self.weights, self._current_shadow(), self._grad16, self.input_dim, self.padded_output_dim,
                                             self.hidden_dim, self.num_layers, self.activation, self.output_activation, not self.training,
                                             calc_grad_inputs, getattr(self, "_wgrad_side", None))
        return ffmlp_forward(inputs, self.weights, self.input_dim, self.padded_output_dim, self.hidden_dim, self.num_layers, self.activation,
                             self.output_activation, not self.training, calc_grad_inputs)

    def __repr__(self):
        return (f"FFMLP: input_dim={self.input_dim} output_dim={self.output_dim} hidden_dim={self.hidden_dim} "
                f"num_layers={self.num_layers} activation={self.activation}")

    def reset_parameters(self):
        torch.manual_seed(42)
        std = math.sqrt(3 / self.hidden_dim)
        self.weights.data.uniform_(-std, std)
        if getattr(self, "_shadow16", None) is not None:
            self._shadow16.copy_(self.weights.detach())

    def forward(self, inputs):
        B, C = inputs.shape
        pad = 128 - (B % 128)                                   # always pads, by a full 128 when B % 128 == 0 (ffmlp.py:157)
        if pad > 0:
            inputs = torch.cat([inputs, torch.zeros(pad, C, dtype=inputs.dtype, device=inputs.device)], dim=0)
        outputs = self.call_fused(inputs, inputs.requires_grad)       # through the fp16 side buffers when an optimiser enabled them
        if B != outputs.shape[0] or self.padded_output_dim != self.output_dim:
            outputs = outputs[:B, :self.output_dim]
        return outputs
