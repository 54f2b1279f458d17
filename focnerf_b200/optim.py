"""`ScaledAdam`: torch.optim.Adam(betas, eps) + torch.amp.GradScaler of the reference's train step
(main_nerf.py:211,472; nerf/utils.py:866-868) as ONE device-side sweep (csrc/optim.cu): non-finite check, Adam update,
fp16 shadow of the updated parameters, in-place zeroing of the gradient.  Same arithmetic and the same scale schedule as
the torch pair; lr, scale, step count and the inf flag are device scalars, so the step is CUDA-graph capturable and an
LR scheduler changes `lr` without re-capturing.

    opt = ScaledAdam(model.parameters(), lr=1e-2, betas=(0.9, 0.99), eps=1e-15)
    g16, s16 = model.encoder.enable_fp16_side_buffers(); opt.attach(model.encoder.embeddings, grad16=g16, shadow16=s16)
    opt.zero_grad(); opt.scale_loss(loss).backward(); opt.step()
"""
import ctypes

import torch

from ._lib import check, lib, require_cuda, stream


class FocAdamTensor(ctypes.Structure):
    """include/foc_b200.h: struct FocAdamTensor"""
    _fields_ = [("param", ctypes.c_void_p), ("exp_avg", ctypes.c_void_p), ("exp_avg_sq", ctypes.c_void_p), ("grad", ctypes.c_void_p),
                ("shadow16", ctypes.c_void_p), ("n", ctypes.c_uint64), ("grad_dtype", ctypes.c_int32), ("zero_grad", ctypes.c_int32)]


class ScaledAdam:
    MAX_TENSORS = 8                    # per launch (csrc/optim.cu kMaxTensors)

    def __init__(self, params, lr=1e-2, betas=(0.9, 0.99), eps=1e-15, loss_scaling=True, init_scale=65536.0, growth_factor=2.0,
                 backoff_factor=0.5, growth_interval=2000, persistent_grads=False):
        self.params = [p for p in params]
        # persistent_grads: every gradient tensor (param.grad included) is zeroed by the step kernel after it is consumed and
        # zero_grad() becomes a no-op, so gradients live at fixed addresses -- what a CUDA graph that applies the PREVIOUS
        # step's gradients needs (Trainer's pipelined step)
        self.persistent_grads = bool(persistent_grads)
        if not self.params:
            raise ValueError("ScaledAdam: no parameters")
        for p in self.params:
            require_cuda(p)
            if p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("ScaledAdam: parameters must be contiguous fp32 CUDA tensors")
        dev = self.params[0].device
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.growth_factor, self.backoff_factor, self.growth_interval = float(growth_factor), float(backoff_factor), int(growth_interval)
        self.exp_avg = [torch.zeros_like(p, memory_format=torch.contiguous_format) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p, memory_format=torch.contiguous_format) for p in self.params]
        self.lr = torch.full((1,), float(lr), dtype=torch.float32, device=dev)
        self.scale = torch.full((1,), float(init_scale), dtype=torch.float32, device=dev) if loss_scaling else None
        self.growth_tracker = torch.zeros(1, dtype=torch.int32, device=dev)
        self.found_inf = torch.zeros(1, dtype=torch.float32, device=dev)
        self.step_count = torch.zeros(1, dtype=torch.int32, device=dev)
        self._side = {}                # id(param) -> (grad16, shadow16)

    # ---- wiring ---------------------------------------------------------------------------------------------------
    def attach(self, param, grad16=None, shadow16=None):
        """Route `param`'s gradient through a persistent fp16 accumulation buffer (filled by the operator's backward instead
        of param.grad) and/or keep an fp16 copy of the parameter current for the next forward."""
        if not any(param is p for p in self.params):
            raise ValueError("ScaledAdam.attach: not one of this optimiser's parameters")
        for t in (grad16, shadow16):
            if t is not None and (t.dtype != torch.float16 or t.numel() != param.numel() or not t.is_contiguous() or not t.is_cuda):
                raise RuntimeError("ScaledAdam.attach: side buffers must be contiguous CUDA half tensors of the parameter's size")
        self._side[id(param)] = (grad16, shadow16)
        if shadow16 is not None:
            shadow16.copy_(param.detach())

    def grad_tensors(self):
        """The tensors a data-parallel all-reduce must average before step(): the fp16 side buffer where one is attached,
        else param.grad."""
        out = []
        for p in self.params:
            g16 = self._side.get(id(p), (None, None))[0]
            g = g16 if g16 is not None else p.grad
            if g is not None:
                out.append(g)
        return out

    def scale_loss(self, loss):
        return loss if self.scale is None else loss * self.scale

    def set_lr(self, lr):
        self.lr.fill_(float(lr))

    def zero_grad(self, set_to_none=True):
        if self.persistent_grads:
            return
        for p in self.params:
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                else:
                    p.grad.zero_()
        # attached fp16 gradient buffers are zeroed by the step kernel itself

    # ---- the step -------------------------------------------------------------------------------------------------
    def _descriptors(self):
        rows = []
        for p, m, v in zip(self.params, self.exp_avg, self.exp_avg_sq):
            g16, s16 = self._side.get(id(p), (None, None))
            g = g16 if g16 is not None else p.grad
            if g is None:
                continue
            if g.dtype not in (torch.float32, torch.float16) or not g.is_contiguous():
                if self.persistent_grads:
                    raise RuntimeError("ScaledAdam(persistent_grads=True) needs contiguous fp32 / fp16 gradients")
                g = g.contiguous().float()
            rows.append(FocAdamTensor(p.data_ptr(), m.data_ptr(), v.data_ptr(), g.data_ptr(), 0 if s16 is None else s16.data_ptr(), p.numel(),
                                      1 if g.dtype == torch.float16 else 0, 1 if (g16 is not None or self.persistent_grads) else 0))
        return rows

    @torch.no_grad()
    def step(self):
        rows = self._descriptors()
        if not rows:
            return
        l, st = lib(), stream()
        b1, b2 = self.betas
        sc = None if self.scale is None else self.scale.data_ptr()
        fi = None if self.scale is None else self.found_inf.data_ptr()
        chunks = [rows[i:i + self.MAX_TENSORS] for i in range(0, len(rows), self.MAX_TENSORS)]
        arrays = [(FocAdamTensor * len(c))(*c) for c in chunks]
        if self.scale is not None:
            for a in arrays:
                check(l.foc_adam_nonfinite(ctypes.cast(a, ctypes.c_void_p), len(a), fi, st), "adam_nonfinite")
        for a in arrays:
            check(l.foc_adam_step(ctypes.cast(a, ctypes.c_void_p), len(a), self.lr.data_ptr(), b1, b2, self.eps, sc, fi,
                                  self.step_count.data_ptr(), st), "adam_step")
        check(l.foc_scaler_update(sc, self.growth_tracker.data_ptr(), fi, self.step_count.data_ptr(), self.growth_factor, self.backoff_factor,
                                  self.growth_interval, st), "scaler_update")

    # ---- checkpointing (nerf/utils.py:1434-1556 saves optimizer + scaler state) -------------------------------------
    def state_dict(self):
        return {"exp_avg": [t.clone() for t in self.exp_avg], "exp_avg_sq": [t.clone() for t in self.exp_avg_sq], "lr": float(self.lr.item()),
                "step": int(self.step_count.item()), "scale": None if self.scale is None else float(self.scale.item()),
                "growth_tracker": int(self.growth_tracker.item()), "betas": self.betas, "eps": self.eps}

    def load_state_dict(self, sd):
        for dst, src in zip(self.exp_avg, sd["exp_avg"]):
            dst.copy_(src)
        for dst, src in zip(self.exp_avg_sq, sd["exp_avg_sq"]):
            dst.copy_(src)
        self.lr.fill_(sd["lr"]); self.step_count.fill_(sd["step"]); self.growth_tracker.fill_(sd["growth_tracker"])
        if self.scale is not None and sd.get("scale") is not None:
            self.scale.fill_(sd["scale"])
        self.refresh_shadows()

    def refresh_shadows(self):
        """Re-derive the fp16 shadows from the fp32 parameters (after load_state_dict or any out-of-band parameter write)."""
        for p in self.params:
            s16 = self._side.get(id(p), (None, None))[1]
            if s16 is not None:
                s16.copy_(p.detach())
