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
        self.lr0, self.decay_iters = float(lr), None           # set_exp_decay(): the reference's per-iteration LambdaLR, on the device
        self.sched_iter = torch.zeros(1, dtype=torch.int32, device=dev)
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

    def set_exp_decay(self, iters, lr0=None):
        """main_nerf.py:477 `LambdaLR(optimizer, lambda iter: 0.1 ** min(iter / opt.iters, 1))` stepped after every iteration
        (scheduler_update_every_step=True): from now on step() also advances lr on the device (csrc/optim.cu k_scaler_update)."""
        if lr0 is not None:
            self.lr0 = float(lr0)
        self.decay_iters = None if iters is None else float(iters)
        self.sched_iter.zero_()
        self.lr.fill_(self.lr0)

    def _scaler_update(self, l, sc, fip, st):
        if self.decay_iters is None:
            check(l.foc_scaler_update(sc, self.growth_tracker.data_ptr(), fip, self.step_count.data_ptr(), self.growth_factor, self.backoff_factor,
                                      self.growth_interval, st), "scaler_update")
        else:
            check(l.foc_scaler_update_sched(sc, self.growth_tracker.data_ptr(), fip, self.step_count.data_ptr(), self.growth_factor,
                                            self.backoff_factor, self.growth_interval, self.lr.data_ptr(), self.sched_iter.data_ptr(), self.lr0,
                                            self.decay_iters, st), "scaler_update_sched")

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
            if g16 is not None and p.grad is not None:
                # an operator that does not know about the side buffer delivered this gradient through autograd: applying only
                # grad16 would silently drop it
                raise RuntimeError("ScaledAdam: a parameter with an attached fp16 gradient buffer also has a .grad -- some forward "
                                   "bypassed the side buffers (call the module under fp16 autocast, or detach the buffers)")
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
        self._scaler_update(l, sc, fi, st)

    # ---- checkpointing (nerf/utils.py:1434-1556 saves optimizer + scaler state) -------------------------------------
    def state_dict(self):
        return {"exp_avg": [t.clone() for t in self.exp_avg], "exp_avg_sq": [t.clone() for t in self.exp_avg_sq], "lr": float(self.lr.item()),
                "step": int(self.step_count.item()), "scale": None if self.scale is None else float(self.scale.item()),
                "growth_tracker": int(self.growth_tracker.item()), "betas": self.betas, "eps": self.eps,
                "lr0": self.lr0, "decay_iters": self.decay_iters, "sched_iter": int(self.sched_iter.item())}

    def load_state_dict(self, sd):
        for dst, src in zip(self.exp_avg, sd["exp_avg"]):
            dst.copy_(src)
        for dst, src in zip(self.exp_avg_sq, sd["exp_avg_sq"]):
            dst.copy_(src)
        self.lr.fill_(sd["lr"]); self.step_count.fill_(sd["step"]); self.growth_tracker.fill_(sd["growth_tracker"])
        self.lr0, self.decay_iters = sd.get("lr0", self.lr0), sd.get("decay_iters", self.decay_iters)
        self.sched_iter.fill_(sd.get("sched_iter", 0))
        if self.scale is not None and sd.get("scale") is not None:
            self.scale.fill_(sd["scale"])
        self.refresh_shadows()

    def refresh_shadows(self):
        """Re-derive the fp16 shadows from the fp32 parameters (after load_state_dict or any out-of-band parameter write)."""
        for p in self.params:
            s16 = self._side.get(id(p), (None, None))[1]
            if s16 is not None:
                s16.copy_(p.detach())


class FocDpReduce(ctypes.Structure):
    """include/foc_b200.h: struct FocDpReduce"""
    _fields_ = [("world", ctypes.c_int32), ("rank", ctypes.c_int32), ("grad16", ctypes.c_void_p * 8), ("shard_start", ctypes.c_uint64),
                ("shard_n", ctypes.c_uint64), ("reduced", ctypes.c_void_p), ("tail", ctypes.c_void_p * 8), ("tail_n", ctypes.c_uint64),
                ("reduced_tail", ctypes.c_void_p), ("inf_flag", ctypes.c_void_p), ("shard_blocks", ctypes.c_uint32)]


def shard_range(n, world, rank, block=4096):
    """Rank `rank`'s slice of an n-element table sharded over `world` ranks: equal slices whose length is a multiple of `block` (the
    sweep's elements per CTA), the last ranks' slices shorter or empty.  -> (slice length of a full shard, start, length)."""
    per = -(-n // world)
    per = -(-per // block) * block
    start = min(rank * per, n)
    return per, start, min(start + per, n) - start


class FocDpGrads(ctypes.Structure):
    """include/foc_b200.h: struct FocDpGrads"""
    _fields_ = [("world", ctypes.c_int32), ("rank", ctypes.c_int32), ("grad16", ctypes.c_void_p * 8), ("tail", ctypes.c_void_p * 8)]


class PeerAdam(ScaledAdam):
    """ScaledAdam for data-parallel training on one NVSwitch node, with the gradient exchange fused into the optimiser over
    peer memory instead of an NCCL all-reduce in front of it (csrc/peer.cu has the protocol):

    * `table` (the hash grid, 99.9 % of the parameters) is SHARDED: rank r keeps fp32 master values and Adam moments current
      for elements [r * shard, (r + 1) * shard) only, pulls that slice of every rank's fp16 gradient table, and pushes the
      fp16 copy of the updated slice into every rank's table.  The forward / backward only ever touch the fp16 copies, which
      are complete on every rank.  `sync_master()` (a collective) completes the fp32 tensors, e.g. before a checkpoint.
    * every other parameter (the MLP weights, ~10 k values) is replicated: each rank averages all ranks' gradients in the
      same order and applies the same update -- bit-identical replicas without a broadcast.

    The table's fp16 gradient / parameter copies live in a PeerArena; hand them to the encoder with
    `encoder.enable_fp16_side_buffers(opt.grad16, opt.shadow16)` (Trainer does)."""
    BLOCK = 4096                       # csrc/optim.cu kPerBlock: shard boundaries are multiples of it

    def __init__(self, params, table, group=None, **kw):
        import torch.distributed as dist
        from .peer import PeerArena
        kw.pop("persistent_grads", None)
        super().__init__(params, **kw)
        if not any(table is p for p in self.params):
            raise ValueError("PeerAdam: `table` must be one of the parameters")
        self.group, self.table = group, table
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        n = table.numel()
        if n % 8:
            raise RuntimeError("PeerAdam: the sharded table must hold a multiple of 8 values")
        per, self.shard_start, self.shard_n = shard_range(n, self.world, self.rank, self.BLOCK)
        self.shard_len = per
        self.tail = [p for p in self.params if p is not table]
        self.tail_off, off = [], 0
        for p in self.tail:
            self.tail_off.append(off)
            off += -(-p.numel() // 4) * 4
        self.tail_n = off
        g_bytes = -(-n * 2 // 256) * 256
        t_bytes = -(-max(self.tail_n, 4) * 4 // 256) * 256
        self.arena = PeerArena(g_bytes + t_bytes + g_bytes, group=group, device=table.device)
        self._off_grad, self._off_tail, self._off_shadow = 0, g_bytes, g_bytes + t_bytes
        self.grad16 = self.arena.tensor(self._off_grad, n, torch.float16).view(table.shape)
        self.tail_grad = self.arena.tensor(self._off_tail, max(self.tail_n, 4), torch.float32)
        self._grad_span = self.arena.tensor(self._off_grad, (g_bytes + t_bytes) // 4, torch.float32)     # grad16 + tail, zeroed in one fill
        self.shadow16 = self.arena.tensor(self._off_shadow, n, torch.float16).view(table.shape)
        self.attach(table, grad16=self.grad16, shadow16=self.shadow16)
        dev = table.device
        self.reduced = torch.zeros(max(self.shard_n, 4), dtype=torch.float32, device=dev)
        self.reduced_tail = torch.zeros(max(self.tail_n, 4), dtype=torch.float32, device=dev)
        a = self.arena
        self._reduce_args = FocDpReduce(
            self.world, self.rank, (ctypes.c_void_p * 8)(*[a.address(r, self._off_grad) for r in range(self.world)]), self.shard_start, self.shard_n,
            self.reduced.data_ptr(), (ctypes.c_void_p * 8)(*[a.address(r, self._off_tail) for r in range(self.world)]), self.tail_n,
            self.reduced_tail.data_ptr(), a.state.data_ptr() + 4, 0)
        self._peer_shadow = (ctypes.c_void_p * 8)(*[a.address(r, self._off_shadow + 2 * self.shard_start) for r in range(self.world)])
        # exchange inside the sweep (foc_adam_step_dp): pull -> fp32 mean -> Adam -> push in one pass over the shard, two barriers.
        # FOC_DP_FUSED=0 keeps the three-barrier protocol (foc_dp_reduce + foc_adam_step_peers), bit-identical results.
        import os
        self.fused_exchange = os.environ.get("FOC_DP_FUSED", "1") == "1" and (self.world - 1) * per < n      # same answer on every rank
        self._dp_grads = FocDpGrads(self.world, self.rank,
                                    (ctypes.c_void_p * 8)(*[a.address(r, self._off_grad + 2 * self.shard_start) for r in range(self.world)]),
                                    (ctypes.c_void_p * 8)(*[a.address(r, self._off_tail) for r in range(self.world)]))
        dist.barrier(group=group)

    def grad_tensors(self):
        return []                      # nothing for an NCCL all-reduce: step() exchanges the gradients itself

    def _descriptors(self, pull=False):
        """pull=True (foc_adam_step_dp): the gradients are the ranks' own buffers -- the table row names the local fp16 shard, the
        tail rows their slice of the local staged tail; the kernel reads the same offsets on every rank."""
        rows = []
        s, n = self.shard_start, self.shard_n
        ti = [i for i, p in enumerate(self.params) if p is self.table][0]
        if n > 0:
            flat = lambda t: t.view(-1)[s:s + n]
            g = (self.grad16.view(-1)[s:s + n].data_ptr(), 1) if pull else (self.reduced.data_ptr(), 0)
            rows.append(FocAdamTensor(flat(self.table.data).data_ptr(), flat(self.exp_avg[ti]).data_ptr(), flat(self.exp_avg_sq[ti]).data_ptr(),
                                      g[0], 0, n, g[1], 0))
        for p, off in zip(self.tail, self.tail_off):
            i = [k for k, q in enumerate(self.params) if q is p][0]
            s16 = self._side.get(id(p), (None, None))[1]
            rows.append(FocAdamTensor(p.data_ptr(), self.exp_avg[i].data_ptr(), self.exp_avg_sq[i].data_ptr(),
                                      (self.tail_grad if pull else self.reduced_tail).data_ptr() + 4 * off, 0 if s16 is None else s16.data_ptr(),
                                      p.numel(), 0, 0))
        return rows

    @torch.no_grad()
    def step(self):
        l, st = lib(), stream()
        for p, off in zip(self.tail, self.tail_off):           # stage the small gradients where the peers can read them
            g16 = self._side.get(id(p), (None, None))[0]
            g = g16 if g16 is not None else p.grad
            if g is not None:
                self.tail_grad[off:off + p.numel()].copy_(g.reshape(-1))
                if g16 is not None:
                    g16.zero_()                                # consumed: the side buffers ACCUMULATE between steps (like .grad)
        fi = None if self.scale is None else self.found_inf
        b1, b2 = self.betas
        sc = None if self.scale is None else self.scale.data_ptr()
        fip = None if fi is None else fi.data_ptr()
        if self.fused_exchange:
            if fi is not None:                                 # this rank's own gradients (whole fp16 table + staged tail): one local pass
                own = (FocAdamTensor * 2)(FocAdamTensor(0, 0, 0, self.grad16.data_ptr(), 0, self.grad16.numel(), 1, 0),
                                          FocAdamTensor(0, 0, 0, self.tail_grad.data_ptr(), 0, max(self.tail_n, 4), 0, 0))
                check(l.foc_adam_nonfinite(ctypes.cast(own, ctypes.c_void_p), 2, fip, st), "adam_nonfinite")
            self.arena.barrier(fi)                             # every rank's gradients are complete; non-finite flags exchanged
            rows = self._descriptors(pull=True)
            arr = (FocAdamTensor * len(rows))(*rows)
            check(l.foc_adam_step_dp(ctypes.cast(arr, ctypes.c_void_p), len(rows), self.lr.data_ptr(), b1, b2, self.eps, sc, fip,
                                     self.step_count.data_ptr(), 0, ctypes.byref(self._dp_grads), ctypes.cast(self._peer_shadow, ctypes.c_void_p), st),
                  "adam_step_dp")
            self.arena.barrier()                               # every shard of the fp16 table has arrived, every rank has pulled
            self._grad_span.zero_()
            self._scaler_update(l, sc, fip, st)
            return
        self.arena.barrier()                                   # every rank's gradients are complete
        check(l.foc_dp_reduce(ctypes.byref(self._reduce_args), st), "dp_reduce")
        self.arena.barrier(fi)                                 # every rank has pulled what it needs; non-finite flags exchanged
        self._grad_span.zero_()
        rows = self._descriptors()
        arr = (FocAdamTensor * len(rows))(*rows)
        if self.shard_n > 0:
            check(l.foc_adam_step_peers(ctypes.cast(arr, ctypes.c_void_p), len(rows), self.lr.data_ptr(), b1, b2, self.eps, sc, fip,
                                        self.step_count.data_ptr(), 0, self.world, ctypes.cast(self._peer_shadow, ctypes.c_void_p), st), "adam_step_peers")
        else:
            check(l.foc_adam_step(ctypes.cast(arr, ctypes.c_void_p), len(rows), self.lr.data_ptr(), b1, b2, self.eps, sc, fip,
                                  self.step_count.data_ptr(), st), "adam_step")
        self.arena.barrier()                                   # every shard of the fp16 table has arrived
        self._scaler_update(l, sc, fip, st)

    @torch.no_grad()
    def sync_master(self):
        """Collective: complete the sharded fp32 tensors (table parameters and both Adam moments) on every rank."""
        import torch.distributed as dist
        ti = [i for i, p in enumerate(self.params) if p is self.table][0]
        n, per = self.table.numel(), self.shard_len
        for t in (self.table.data, self.exp_avg[ti], self.exp_avg_sq[ti]):
            flat = t.view(-1)
            mine = flat.new_zeros(per)
            mine[:self.shard_n] = flat[self.shard_start:self.shard_start + self.shard_n]
            parts = [flat.new_empty(per) for _ in range(self.world)]
            dist.all_gather(parts, mine, group=self.group)
            flat.copy_(torch.cat(parts)[:n])

    def state_dict(self):
        self.sync_master()
        return super().state_dict()

    def refresh_shadows(self):
        self.shadow16.copy_(self.table.detach())
        for p in self.tail:
            s16 = self._side.get(id(p), (None, None))[1]
            if s16 is not None:
                s16.copy_(p.detach())
