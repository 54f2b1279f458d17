"""`gridencoder.GridEncoder` and `grid_encode` of the reference (gridencoder/grid.py:24-184) over libfoc_b200.

Same constructor, attribute names (`embeddings` [n_entries, level_dim] fp32 Parameter, `offsets` int32 buffer,
`output_dim`, `n_params`), init, AMP convention (only the TABLE is cast to half under autocast, and only when
level_dim is even; coordinates stay fp32) and `__repr__`, so state_dicts interchange with the reference module.

`_backend.grid_encode_forward/backward/grad_total_variation` keep the reference's positional signatures
(gridencoder/src/gridencoder.h:12-15) and its [L,B,C] kernel layout.  The autograd Function itself asks the
library for the point-major [B,L,C] layout, which IS the [B, L*C] tensor the caller receives: the values are
identical to the reference's `outputs.permute(1,0,2).reshape(B, L*C)` without that copy (grid.py:57), and
backward reads the incoming [B, L*C] gradient in place instead of permuting it to [L,B,C] (grid.py:73).
"""
import math

import numpy as np
import torch
import torch.nn as nn
from torch.autograd import Function
from torch.amp import custom_bwd, custom_fwd

from .._lib import check, lib, ptr, require_cuda, stream, valid_rows

_gridtype_to_id = {'hash': 0, 'tiled': 1}
_interp_to_id = {'linear': 0, 'smoothstep': 1}
FOC_LBC, FOC_BLC = 0, 1


def _dtype_code(t):
    if t.dtype == torch.float32:
        return 0
    if t.dtype == torch.float16:
        return 1
    raise RuntimeError("embeddings must be float32 or float16")


def _chk(t, name, dtype=None):
    if t is None:
        return
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be a contiguous tensor")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"{name} must be a {dtype} tensor")


class _Backend:
    """Reference pybind names (gridencoder/src/bindings.cpp:5-9); optional trailing `layout` is new."""

    @staticmethod
    def grid_encode_forward(inputs, embeddings, offsets, outputs, B, D, C, L, S, H, dy_dx, gridtype, align_corners, interp,
                            layout=FOC_LBC):
        _chk(inputs, "inputs", torch.float32); _chk(embeddings, "embeddings"); _chk(offsets, "offsets", torch.int32)
        _chk(outputs, "outputs", embeddings.dtype); _chk(dy_dx, "dy_dx", embeddings.dtype)
        check(lib().foc_grid_encode_forward(ptr(inputs), ptr(embeddings), ptr(offsets), ptr(outputs), B, D, C, L, float(S), H,
                                            ptr(dy_dx), gridtype, int(align_corners), interp, _dtype_code(embeddings), layout,
                                            stream()), "grid_encode_forward")

    @staticmethod
    def grid_encode_backward(grad, inputs, embeddings, offsets, grad_embeddings, B, D, C, L, S, H, dy_dx, grad_inputs, gridtype,
                             align_corners, interp, layout=FOC_LBC):
        _chk(grad, "grad", embeddings.dtype); _chk(inputs, "inputs", torch.float32); _chk(offsets, "offsets", torch.int32)
        _chk(grad_embeddings, "grad_embeddings", embeddings.dtype); _chk(dy_dx, "dy_dx", embeddings.dtype)
        _chk(grad_inputs, "grad_inputs", embeddings.dtype)
        check(lib().foc_grid_encode_backward(ptr(grad), ptr(inputs), ptr(embeddings), ptr(offsets), ptr(grad_embeddings), B, D, C,
                                             L, float(S), H, ptr(dy_dx), ptr(grad_inputs), gridtype, int(align_corners), interp,
                                             _dtype_code(embeddings), layout, stream()), "grid_encode_backward")

    @staticmethod
    def grad_total_variation(inputs, embeddings, grad, offsets, weight, B, D, C, L, S, H, gridtype, align_corners):
        _chk(inputs, "inputs", torch.float32); _chk(embeddings, "embeddings", torch.float32); _chk(grad, "grad", torch.float32)
        _chk(offsets, "offsets", torch.int32)
        check(lib().foc_grad_total_variation(ptr(inputs), ptr(embeddings), ptr(grad), ptr(offsets), weight, B, D, C, L, float(S),
                                             H, gridtype, int(align_corners), stream()), "grad_total_variation")


def level_scales(per_level_scale, base_resolution, num_levels, device="cuda"):
    """Device-evaluated exp2f(l * log2(per_level_scale)) * H - 1 for l < num_levels (diagnostic / test hook)."""
    out = torch.empty(num_levels, dtype=torch.float32, device=device)
    check(lib().foc_grid_level_scales(float(np.log2(per_level_scale)), base_resolution, num_levels, ptr(out), stream()), "grid_level_scales")
    return out


_backend = _Backend()


class _grid_encode(Function):
    @staticmethod
    @custom_fwd(device_type="cuda")
    def forward(ctx, inputs, embeddings, offsets, per_level_scale, base_resolution, calc_grad_inputs=False, gridtype=0,
                align_corners=False, interpolation=0, shadow16=None, grad16=None):
        require_cuda(inputs, embeddings, offsets)
        inputs = inputs.contiguous()
        B, D = inputs.shape
        L = offsets.shape[0] - 1
        C = embeddings.shape[1]
        S = np.log2(per_level_scale)        # float64 here, narrowed to float at the ABI like the reference's pybind call
        H = base_resolution
        ctx.grad16 = None
        if torch.is_autocast_enabled() and C % 2 == 0:
            # fp16 side buffers (GridEncoder.enable_fp16_side_buffers): read the shadow the optimiser keeps current instead
            # of casting the fp32 table every step, and scatter the gradient into the persistent fp16 buffer.
            embeddings = shadow16 if shadow16 is not None else embeddings.to(torch.half)
            ctx.grad16 = grad16
        outputs = torch.empty(B, L * C, device=inputs.device, dtype=embeddings.dtype)
        dy_dx = torch.empty(B, L * D * C, device=inputs.device, dtype=embeddings.dtype) if calc_grad_inputs else None
        rows = valid_rows() if (dy_dx is None and D == 3) else None
        ctx.rows = rows
        if rows is not None:                                      # graph-replayed step: only the first rows[0] rows are valid
            check(lib().foc_grid_encode_forward_rows(ptr(inputs), ptr(embeddings), ptr(offsets), ptr(outputs), B, D, C, L, float(S), H, gridtype,
                                                     int(align_corners), interpolation, _dtype_code(embeddings), ptr(rows), stream()),
                  "grid_encode_forward_rows")
        else:
            _backend.grid_encode_forward(inputs, embeddings, offsets, outputs, B, D, C, L, S, H, dy_dx, gridtype, align_corners,
                                         interpolation, FOC_BLC)
        ctx.save_for_backward(inputs, embeddings, offsets, dy_dx)
        ctx.dims = [B, D, C, L, S, H, gridtype, interpolation]
        ctx.align_corners = align_corners
        return outputs

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, grad):
        inputs, embeddings, offsets, dy_dx = ctx.saved_tensors
        B, D, C, L, S, H, gridtype, interpolation = ctx.dims
        grad = grad.contiguous()                                   # [B, L*C] read in place (layout FOC_BLC)
        if grad.dtype != embeddings.dtype:
            grad = grad.to(embeddings.dtype)
        side = ctx.grad16 is not None
        grad_embeddings = ctx.grad16 if side else torch.zeros_like(embeddings)      # the side buffer accumulates, like .grad
        grad_inputs = torch.zeros_like(inputs, dtype=embeddings.dtype) if dy_dx is not None else None
        if ctx.rows is not None:
            check(lib().foc_grid_encode_backward_rows(ptr(grad), ptr(inputs), ptr(offsets), ptr(grad_embeddings), B, D, C, L, float(S), H, gridtype,
                                                      int(ctx.align_corners), interpolation, _dtype_code(embeddings), ptr(ctx.rows), stream()),
                  "grid_encode_backward_rows")
        else:
            _backend.grid_encode_backward(grad, inputs, embeddings, offsets, grad_embeddings, B, D, C, L, S, H, dy_dx, grad_inputs,
                                          gridtype, ctx.align_corners, interpolation, FOC_BLC)
        if dy_dx is not None:
            grad_inputs = grad_inputs.to(inputs.dtype)
        return grad_inputs, (None if side else grad_embeddings), None, None, None, None, None, None, None, None, None


grid_encode = _grid_encode.apply


class GridEncoder(nn.Module):
    def __init__(self, input_dim=3, num_levels=16, level_dim=2, per_level_scale=2, base_resolution=16, log2_hashmap_size=19,
                 desired_resolution=None, gridtype='hash', align_corners=False, interpolation='linear'):
        super().__init__()
        # the finest resolution desired at the last level; overrides per_level_scale (grid.py:100-102)
        if desired_resolution is not None:
            per_level_scale = np.exp2(np.log2(desired_resolution / base_resolution) / (num_levels - 1))
        self.input_dim = input_dim
        self.num_levels = num_levels
        self.level_dim = level_dim
        self.per_level_scale = per_level_scale
        self.log2_hashmap_size = log2_hashmap_size
        self.base_resolution = base_resolution
        self.output_dim = num_levels * level_dim
        self.gridtype = gridtype
        self.gridtype_id = _gridtype_to_id[gridtype]
        self.interpolation = interpolation
        self.interp_id = _interp_to_id[interpolation]
        self.align_corners = align_corners

        offsets, offset = [], 0
        self.max_params = 2 ** log2_hashmap_size
        for i in range(num_levels):
            resolution = int(np.ceil(base_resolution * per_level_scale ** i))
            params_in_level = min(self.max_params, (resolution if align_corners else resolution + 1) ** input_dim)
            params_in_level = int(np.ceil(params_in_level / 8) * 8)      # make divisible
            offsets.append(offset)
            offset += params_in_level
        offsets.append(offset)
        self.register_buffer('offsets', torch.from_numpy(np.array(offsets, dtype=np.int32)))
        self.n_params = offsets[-1] * level_dim
        self.embeddings = nn.Parameter(torch.empty(offset, level_dim))
        self._shadow16, self._grad16 = None, None                  # see enable_fp16_side_buffers
        self.reset_parameters()

    def reset_parameters(self):
        std = 1e-4
        self.embeddings.data.uniform_(-std, std)
        if self._shadow16 is not None:
            self._shadow16.copy_(self.embeddings.detach())

    def enable_fp16_side_buffers(self, grad16=None, shadow16=None):
        """Not in the reference.  Allocates (grad16, shadow16), both [offset, C] half: under fp16 autocast the forward then
        gathers from `shadow16` (no per-step `.to(half)` of the 49 MB table) and the backward scatters into `grad16`
        (no per-step zeros_like + half->float copy); `embeddings.grad` stays None.  The owner of the buffers -- an
        optimiser such as focnerf_b200.optim.ScaledAdam -- must refresh shadow16 after every parameter update and zero
        grad16 after consuming it."""
        if self.level_dim % 2 != 0:
            raise RuntimeError("fp16 side buffers need an even level_dim (the fp16 path of the encoder, grid.py:43)")
        for t in (grad16, shadow16):                               # caller-owned storage (e.g. peer-visible memory, optim.PeerAdam)
            if t is not None and (t.dtype != torch.half or tuple(t.shape) != tuple(self.embeddings.shape) or not t.is_contiguous()):
                raise RuntimeError("fp16 side buffers must be contiguous half tensors of the table's shape")
        if shadow16 is None:
            shadow16 = torch.empty_like(self.embeddings, dtype=torch.half)
        shadow16.copy_(self.embeddings.detach())
        if grad16 is None:
            grad16 = torch.empty_like(shadow16)
        grad16.zero_()
        self._shadow16, self._grad16 = shadow16, grad16
        self._shadow_version = self.embeddings._version
        return self._grad16, self._shadow16

    def _current_shadow(self):
        """The fp16 shadow, re-derived first when the fp32 Parameter was written behind the optimiser's back (load_state_dict,
        reset_parameters, any in-place torch op: they bump `_version`; the optimiser's own kernel writes through raw pointers and
        refreshes the shadow itself)."""
        if self._shadow16 is not None and self.embeddings._version != self._shadow_version:
            with torch.no_grad():
                self._shadow16.copy_(self.embeddings.detach())
            self._shadow_version = self.embeddings._version
        return self._shadow16

    def __repr__(self):
        return (f"GridEncoder: input_dim={self.input_dim} num_levels={self.num_levels} level_dim={self.level_dim} "
                f"resolution={self.base_resolution} -> {int(round(self.base_resolution * self.per_level_scale ** (self.num_levels - 1)))} "
                f"per_level_scale={self.per_level_scale:.4f} params={tuple(self.embeddings.shape)} gridtype={self.gridtype} "
                f"align_corners={self.align_corners} interpolation={self.interpolation}")

    def forward(self, inputs, bound=1):
        if math.frexp(float(bound))[0] == 0.5 and inputs.dtype == torch.float32:
            # power-of-two bound: x / (2 bound) is exact, so 0.5 + x / (2 bound) rounds exactly like (x + bound) / (2 bound)
            # (grid.py:149) -- the same bits from one elementwise launch instead of two
            inputs = torch.rsub(inputs, 0.5, alpha=-1.0 / (2.0 * float(bound)))
        else:
            inputs = (inputs + bound) / (2 * bound)                # [-bound, bound] -> [0, 1]
        prefix_shape = list(inputs.shape[:-1])
        inputs = inputs.view(-1, self.input_dim)
        outputs = grid_encode(inputs, self.embeddings, self.offsets, self.per_level_scale, self.base_resolution,
                              inputs.requires_grad, self.gridtype_id, self.align_corners, self.interp_id, self._current_shadow(), self._grad16)
        return outputs.view(prefix_shape + [self.output_dim])

    @torch.amp.autocast("cuda", enabled=False)                     # always fp32 (grid.py:164)
    def grad_total_variation(self, weight=1e-7, inputs=None, bound=1, B=1000000):
        D = self.input_dim
        C = self.embeddings.shape[1]
        L = self.offsets.shape[0] - 1
        S = np.log2(self.per_level_scale)
        H = self.base_resolution
        if inputs is None:
            inputs = torch.rand(B, self.input_dim, device=self.embeddings.device)
        else:
            inputs = ((inputs + bound) / (2 * bound)).view(-1, self.input_dim)
            B = inputs.shape[0]
        if self.embeddings.grad is None:
            raise ValueError('grad is None, should be called after loss.backward() and before optimizer.step()!')
        _backend.grad_total_variation(inputs.contiguous(), self.embeddings, self.embeddings.grad, self.offsets, weight, B, D, C,
                                      L, S, H, self.gridtype_id, self.align_corners)
