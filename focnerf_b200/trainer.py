"""Train step and render loop of the cuda_ray path, restating `NeRFRenderer.run_cuda`
(legacy/nerf/renderer.py:256-377 -- the working copy) and the optimisation step of
legacy/nerf/utils.py:468-560, 863-868 (fp16 autocast, MSE, GradScaler, Adam(betas=(0.9,0.99), eps=1e-15)).

This is the repo's public API for the hot path (bench.py's `e2e` goes through `Trainer.train_step_host`); the operators
it calls are the same ones the reference's unchanged renderer reaches through `focnerf_b200.install_aliases()`.
"""
import torch

from . import raymarching
from .network import NeRFNetwork


def _pad128(n):
    return max(128, -(-int(n) // 128) * 128)


class Trainer:
    """Defaults follow the reference's recipe: lr 1e-3 (main_nerf.py:26), Adam(betas=(0.9, 0.99), eps=1e-15) (main_nerf.py:472),
    density_thresh 10, min_near 0.2, max_steps 1024, density-grid refresh every update_extra_interval = 16 steps (main_nerf.py:33),
    `iters` = the horizon of the per-iteration LambdaLR 0.1 ** min(iter / iters, 1) (main_nerf.py:25,477; None = constant lr)."""

    def __init__(self, bound=1.0, density_grid=None, density_thresh=10.0, min_near=0.2, dt_gamma=0.0, max_steps=1024,
                 lr=1e-3, fp16=True, device="cuda", grid_size=128, seed=0, fused=True, use_graph=False, optimizer="fused", pipeline=False,
                 data_parallel=None, group=None, iters=None, update_extra_interval=16):
        if not torch.cuda.is_available():
            raise RuntimeError("focnerf_b200.Trainer needs a CUDA device (there is no CPU path)")
        import math
        import os
        self._wgrad_side = None
        torch.manual_seed(seed)
        self.device = torch.device(device)
        self.bound, self.min_near, self.dt_gamma, self.max_steps = float(bound), min_near, dt_gamma, max_steps
        self.cascade = 1 + math.ceil(math.log2(bound))
        self.grid_size, self.fp16 = grid_size, fp16
        self.model = NeRFNetwork(bound=bound).to(self.device)
        self.aabb = torch.tensor([-bound, -bound, -bound, bound, bound, bound], dtype=torch.float32, device=self.device)
        self.density_bitfield = torch.zeros(self.cascade * grid_size ** 3 // 8, dtype=torch.uint8, device=self.device)
        self.density_grid = torch.zeros(self.cascade, grid_size ** 3, dtype=torch.float32, device=self.device)   # renderer.py:88-89
        self.mean_density = torch.zeros(1, dtype=torch.float32, device=self.device)      # stays on the device
        self.density_thresh, self.density_scale, self.iter_density = float(density_thresh), 1.0, 0
        if density_grid is not None:
            self.set_density_grid(density_grid, density_thresh)
        self.step_counter = torch.zeros(16, 2, dtype=torch.int32, device=self.device)     # renderer.py:96-98
        self._fuse_rgb = os.environ.get("FOC_FUSE_RGB") == "1"
        self.mean_count, self.local_step = 0, 0
        self.global_step, self.update_extra_interval, self.iters, self.lr = 0, int(update_extra_interval), iters, float(lr)
        self.lr_scheduler = None
        self.grad_sync = None
        # fused: field-network wiring through field_glue.cu (same function, ~25 fewer launches per step)
        # use_graph: replay the whole optimisation step as ONE CUDA graph once the sample-buffer size has settled
        self.fused, self.use_graph = fused, use_graph
        self._graph = None
        self._refresh_graph, self._refresh_buf = None, None
        self._render_cap = {}                           # rays per frame -> sample-buffer capacity learnt from the previous frame
        self.render_overflows = 0                       # frames the speculative single pass had to render again with exact sizes
        self.capture_count = 0                          # how many times the step was (re-)captured (bench.py reports it for the timed region)
        # pipeline=True (fused optimiser + graph only, opt-in): the optimiser sweep of step i-1 runs on a side stream next to
        # the march of step i (the march reads rays + bitfield, never the parameters).  `_pending` = a backward has left
        # gradients that no sweep has consumed yet; flush() applies them.  Measured on B200: 0.441 ms vs 0.443 ms per step --
        # the sweep saturates HBM/L2 and the latency-bound march slows down beside it by as much as the overlap saves, so the
        # default keeps the plain order (and parameters that are always current).
        self.pipeline = bool(pipeline) and optimizer == "fused" and data_parallel != "peer"
        # data_parallel (needs an initialised torch.distributed group, one process per GPU): every rank marches its own rays,
        # the replicas start from rank 0's parameters and see the mean gradient every step.
        #   "peer": the exchange is fused into the optimiser over NVLink peer memory (optim.PeerAdam, csrc/peer.cu) --
        #           sharded table sweep, no NCCL on the data path;   "nccl": all-reduce of the gradient tensors, then the
        #           replicated sweep (the baseline the peer path is measured against)
        if data_parallel not in (None, "peer", "nccl"):
            raise ValueError("data_parallel must be None, 'peer' or 'nccl'")
        if data_parallel == "peer" and not (optimizer == "fused" and fp16):
            data_parallel = "nccl"
        self.data_parallel = data_parallel
        self._dp_group = group
        if data_parallel is not None:
            from . import dist as fdist
            fdist.broadcast_parameters(list(self.model.parameters()), group=group)
        self._pending = False
        self._side_stream, self._capture_stream = None, None
        # optimizer="fused": ScaledAdam (csrc/optim.cu) with the encoder's fp16 side buffers; "torch": the reference's
        # torch.optim.Adam + torch.amp.GradScaler pair, kept for A/B checks
        if optimizer == "fused":
            from .optim import PeerAdam, ScaledAdam
            if data_parallel == "peer":
                self.optimizer = PeerAdam(self.model.parameters(), table=self.model.encoder.embeddings, group=group, lr=lr, betas=(0.9, 0.99),
                                          eps=1e-15, loss_scaling=True)
                self.model.encoder.enable_fp16_side_buffers(self.optimizer.grad16, self.optimizer.shadow16)
            else:
                self.optimizer = ScaledAdam(self.model.parameters(), lr=lr, betas=(0.9, 0.99), eps=1e-15, loss_scaling=fp16,
                                            persistent_grads=self.pipeline)
                if fp16:
                    g16, s16 = self.model.encoder.enable_fp16_side_buffers()
                    self.optimizer.attach(self.model.encoder.embeddings, grad16=g16, shadow16=s16)
            if fp16 and fused:                        # the MLP weights too: no per-step casts, gradients written in place
                from .ffmlp.ffmlp import WgradSide
                self._wgrad_side = WgradSide() if os.environ.get("FOC_WGRAD_SIDE", "0") == "1" else None
                for net in (self.model.sigma_net, self.model.color_net):
                    g16, s16 = net.enable_fp16_side_buffers()
                    self.optimizer.attach(net.weights, grad16=g16, shadow16=s16)
                    net._wgrad_side = self._wgrad_side    # weight-gradient reductions beside the rest of the backward pass
            self.scaler = None
            if iters is not None:
                self.optimizer.set_exp_decay(iters, lr)
        else:
            self.optimizer = torch.optim.Adam(self.model.get_params(lr), betas=(0.9, 0.99), eps=1e-15, fused=True, capturable=use_graph)
            self.scaler = torch.amp.GradScaler("cuda", enabled=fp16)
            if iters is not None:
                self.lr_scheduler = torch.optim.lr_scheduler.LambdaLR(self.optimizer, lambda it: 0.1 ** min(it / iters, 1))
        if data_parallel == "nccl":
            from . import dist as fdist
            self.grad_sync = lambda: fdist.all_reduce_mean(self.grad_tensors(), group=group)

    def set_density_grid(self, density_grid, density_thresh=10.0):
        """density_grid [cascade, H^3] (Morton order) -> bitfield, as update_extra_state does (renderer.py:529-531)."""
        g = density_grid.to(self.device, torch.float32).contiguous()
        self.density_grid.copy_(g.view_as(self.density_grid))
        thresh = min(float(g.clamp(min=0).mean()), density_thresh)
        raymarching.packbits(g, thresh, self.density_bitfield)

    @torch.no_grad()
    def mark_untrained_grid(self, poses, intrinsic):
        """NeRFRenderer.mark_untrained_grid (legacy/nerf/renderer.py:380-443): poses [B,4,4] c2w (numpy or tensor), intrinsic = (fx, fy, cx,
        cy); grid cells outside every camera frustum get density -1 (never refreshed, never marched).  Returns how many were marked."""
        from ._lib import check, lib, ptr, stream
        poses = torch.as_tensor(poses, dtype=torch.float32).to(self.device).contiguous().view(-1, 4, 4)
        fx, fy, cx, cy = (float(v) for v in intrinsic)
        n_marked = torch.zeros(1, dtype=torch.int32, device=self.device)
        check(lib().foc_mark_untrained_grid(ptr(poses), poses.shape[0], fx, fy, cx, cy, self.grid_size, self.cascade, self.bound,
                                            ptr(self.density_grid), ptr(n_marked), stream()), "mark_untrained_grid")
        return int(n_marked.item())

    def _density_at(self, xyzs):
        """sigma [m] f32 at xyzs [m,3]: `self.density(x)['sigma'] * density_scale` (renderer.py:480-481), through the fused operators when
        m is a multiple of 128 (no zero-row padding cat)."""
        with torch.autocast("cuda", dtype=torch.float16, enabled=self.fp16):
            m_ = self.model
            if self.fp16 and self.fused and xyzs.shape[0] % 128 == 0 and m_.in_dim == 32 and m_.sigma_net.hidden_dim == 64:
                h = m_.sigma_net.call_fused(m_.encoder(xyzs, bound=m_.bound), False)
                sig = torch.exp(h[:, 0].float())                       # trunc_exp forward (activation.py:9-11)
            else:
                sig = m_.density(xyzs)["sigma"].reshape(-1).float()
        return sig * self.density_scale

    def _refresh_partial_body(self, decay):
        """The steady-state update (renderer.py:487-516) with STATIC shapes, so that it can be captured: per cascade H^3/4 uniformly
        random cells + H^3/4 cells drawn uniformly from the occupied ones.  The reference draws the latter as
        `nonzero(grid > 0)[randint(0, n_occ, [m])]` (a host sync for the shape); here the k-th occupied cell, k = floor(u * n_occ) with
        u ~ U[0,1), is found by a binary search in the running count of occupied cells: the same distribution, no host read."""
        from ._lib import check, lib, ptr, stream
        H, n, dev = self.grid_size, self.grid_size ** 3, self.device
        m = n // 4
        tmp = self._refresh_buf["tmp"]
        tmp.fill_(-1.0)
        for cas in range(self.cascade):
            bound_c = float(min(2 ** cas, self.bound))
            coords = torch.randint(0, H, (m, 3), device=dev)
            idx_rand = raymarching.morton3D(coords)
            running = torch.cumsum((self.density_grid[cas] > 0).to(torch.int32), 0, dtype=torch.int32)
            n_occ = running[-1:]                                                  # device scalar
            k = torch.minimum((torch.rand(m, device=dev) * n_occ.float()).to(torch.int32), (n_occ - 1).clamp(min=0))
            occ = torch.searchsorted(running, k + 1, out_int32=True).clamp_(max=n - 1)
            indices = torch.cat([idx_rand, occ], 0).contiguous()
            noise = torch.rand(2 * m, 3, device=dev)
            xyzs = self._refresh_buf["xyzs"]
            check(lib().foc_density_cell_positions(ptr(indices), 2 * m, H, bound_c, ptr(noise), ptr(xyzs), stream()), "density_cell_positions")
            # a cell drawn twice keeps the LARGER of its two samples: the reference's `tmp_grid[cas, indices] = sigmas` keeps an arbitrary
            # one (renderer.py:516); the maximum is one of them and makes the result deterministic, so data-parallel replicas that
            # share the seed hold bit-identical grids and bitfields without a broadcast
            tmp[cas].scatter_reduce_(0, indices.long(), self._density_at(xyzs), "amax", include_self=True)
        total = self.cascade * n
        check(lib().foc_density_grid_update(ptr(self.density_grid), ptr(tmp), total, decay, self.density_thresh, ptr(self.mean_density),
                                            ptr(self.density_bitfield), ptr(self._refresh_buf["scratch"]), stream()), "density_grid_update")

    @torch.no_grad()
    def update_extra_state(self, decay=0.95):
        """NeRFRenderer.update_extra_state (legacy/nerf/renderer.py:445-536): re-sample the field's density at jittered cell
        positions (every cell for the first 16 calls, then H^3/4 random + H^3/4 occupied cells per cascade), EMA-max into
        the density grid, threshold at min(mean, density_thresh) into the bitfield the marcher reads.  Cells are generated in
        Morton order on the device (no meshgrid loops) and the mean / threshold never visit the host (csrc/density_grid.cu).
        The steady-state partial update has static shapes and, with use_graph, is replayed as ONE CUDA graph; the only host read
        left is the reference's own `mean_count` (renderer.py:533-535)."""
        from ._lib import check, lib, ptr, stream
        self.flush()
        H, n, dev = self.grid_size, self.grid_size ** 3, self.device
        was_training = self.model.training
        self.model.eval()
        if self.iter_density >= 16:
            if self._refresh_buf is None:
                total = self.cascade * n
                self._refresh_buf = {"tmp": torch.empty_like(self.density_grid), "xyzs": torch.empty(n // 2, 3, dtype=torch.float32, device=dev),
                                     "scratch": torch.empty(max(4, lib().foc_density_grid_update_scratch_bytes(total)), dtype=torch.uint8, device=dev)}
            if self.use_graph and self.fp16 and self.fused:
                if self._refresh_graph is None or self._refresh_graph[0] != decay:
                    self._refresh_partial_body(decay)              # warm-up outside the capture (allocator, lazy module loads)
                    torch.cuda.synchronize()
                    g = torch.cuda.CUDAGraph()
                    if self._capture_stream is None:
                        self._capture_stream = torch.cuda.Stream(priority=-1)
                    with torch.cuda.graph(g, stream=self._capture_stream):
                        self._refresh_partial_body(decay)
                    self._refresh_graph = (decay, g)
                self._refresh_graph[1].replay()
            else:
                self._refresh_partial_body(decay)
        else:
            tmp = torch.full_like(self.density_grid, -1.0)
            for cas in range(self.cascade):
                bound_c = float(min(2 ** cas, self.bound))
                noise = torch.rand(n, 3, device=dev)
                xyzs = torch.empty(n, 3, dtype=torch.float32, device=dev)
                check(lib().foc_density_cell_positions(None, n, H, bound_c, ptr(noise), ptr(xyzs), stream()), "density_cell_positions")
                tmp[cas] = self._density_at(xyzs)
            total = self.cascade * n
            scratch = torch.empty(max(4, lib().foc_density_grid_update_scratch_bytes(total)), dtype=torch.uint8, device=dev)
            check(lib().foc_density_grid_update(ptr(self.density_grid), ptr(tmp), total, decay, self.density_thresh, ptr(self.mean_density),
                                                ptr(self.density_bitfield), ptr(scratch), stream()), "density_grid_update")
        self.iter_density += 1
        self.model.train(was_training)
        self.update_mean_count()
        self.local_step = 0                              # renderer.py:536

    # ------------------------------------------------------------------------------------------------ training
    def render_train(self, rays_o, rays_d, bg_color=1.0, perturb=True, force_all_rays=False, counter=None, mean_count=None):
        """run_cuda, training branch (legacy/nerf/renderer.py:268-319)."""
        nears, fars = raymarching.near_far_from_aabb(rays_o, rays_d, self.aabb, self.min_near)
        if counter is None:
            counter = self.step_counter[self.local_step % 16]
            self.local_step += 1
        counter.zero_()
        xyzs, dirs, deltas, rays = raymarching.march_rays_train(
            rays_o, rays_d, self.bound, self.density_bitfield, self.cascade, self.grid_size, nears, fars, counter,
            self.mean_count if mean_count is None else mean_count, perturb, 128, force_all_rays, self.dt_gamma, self.max_steps)
        sigmas, rgbs = (self.model.forward_fused if self.fused else self.model)(xyzs, dirs)
        weights_sum, depth, image = raymarching.composite_rays_train(sigmas, rgbs, deltas, rays)
        image = image + (1 - weights_sum).unsqueeze(-1) * bg_color
        depth = torch.clamp(depth - nears, min=0) / (fars - nears)
        return image, depth, weights_sum

    def render_train_loss(self, rays_o, rays_d, gt_rgb, bg_color=1.0, perturb=True, counter=None, mean_count=None, after_march=None):
        """render_train + `mse_loss(image, gt)` with the background mix and the loss fused into the compositor
        (raymarching.composite_rays_train_mse): same values, ~14 elementwise / reduction launches fewer per step."""
        nears, fars = raymarching.near_far_from_aabb(rays_o, rays_d, self.aabb, self.min_near)
        if counter is None:
            counter = self.step_counter[self.local_step % 16]
            self.local_step += 1
        counter.zero_()
        xyzs, dirs, deltas, rays = raymarching.march_rays_train(
            rays_o, rays_d, self.bound, self.density_bitfield, self.cascade, self.grid_size, nears, fars, counter,
            self.mean_count if mean_count is None else mean_count, perturb, 128, False, self.dt_gamma, self.max_steps)
        if after_march is not None:                    # the pipelined step joins the optimiser's side stream here
            after_march()
        # FOC_FUSE_RGB=1: the sigmoid of the colour net and its backward inside the compositor (composite_rays_train_mse_hc).  Same
        # bits, three launches fewer -- and measured 4 us SLOWER per step on B200 (0.402 vs 0.398 ms): the warp-per-ray compositors
        # are latency-bound, the three expf + divisions per sample lengthen their serial chains by more than the two bandwidth-bound
        # sigmoid kernels cost.  Off by default.
        if self._fuse_rgb and xyzs.shape[0] % 128 == 0:
            sigmas, hc = self.model.forward_fused(xyzs, dirs, raw_colour=True)
            loss, _, _, _ = raymarching.composite_rays_train_mse_hc(sigmas, hc, deltas, rays, gt_rgb, bg_color)
        else:
            sigmas, rgbs = self.model.forward_fused(xyzs, dirs)
            loss, _, _, _ = raymarching.composite_rays_train_mse(sigmas, rgbs, deltas, rays, gt_rgb, bg_color)
        return loss

    def _forward_backward(self, rays_o, rays_d, gt_rgb, bg_color=1.0, counter=None, mean_count=None, after_march=None):
        """march -> field -> composite -> loss -> backward (+ gradient all-reduce); no optimiser step (fused optimiser path)."""
        with torch.autocast("cuda", dtype=torch.float16, enabled=self.fp16):
            if self.fused and self.fp16:
                loss = self.render_train_loss(rays_o, rays_d, gt_rgb, bg_color, counter=counter, mean_count=mean_count,
                                              after_march=after_march)
            else:
                if after_march is not None:
                    after_march()
                image, _, _ = self.render_train(rays_o, rays_d, bg_color, counter=counter, mean_count=mean_count)
                loss = torch.nn.functional.mse_loss(image, gt_rgb)
        scale = getattr(self.optimizer, "scale", None)
        if scale is not None and loss.dim() == 0 and loss.dtype == scale.dtype:
            loss.backward(gradient=scale.view(()))     # d(scale * loss): the device scalar itself is the seed gradient (no mul / fill launches)
        else:
            self.optimizer.scale_loss(loss).backward()
        if self._wgrad_side is not None:
            self._wgrad_side.join()                    # the optimiser (or a gradient exchange) reads the weight gradients next
        if self.grad_sync is not None:                 # data-parallel: all-reduce the (scaled) gradients over NCCL
            self.grad_sync()
        return loss

    def reset_graph(self):
        """Drop the captured step: the next train_step captures it again with buffers sized from the CURRENT mean sample count (a
        caller that is about to time a steady state does this once, so that a slow drift since the last capture cannot trigger a
        re-capture in the middle of its measurement)."""
        self.flush()
        self._graph = None

    def flush(self):
        """Apply the gradients a pipelined step left pending.  Called before anything reads the parameters."""
        if self._pending:
            self.optimizer.step()
            self._pending = False

    def _step_body(self, rays_o, rays_d, gt_rgb, bg_color=1.0, counter=None, mean_count=None, rows=None):
        """rows (int32 [1] device tensor, graph-replayed step only): the march publishes this step's valid row count into it and the
        grid / field / compositor operators process that many rows of the fixed-size sample buffers (_lib.set_valid_rows)."""
        from ._lib import set_valid_rows
        prev = set_valid_rows(rows)
        try:
            if self.scaler is None:
                self.flush()
                self.optimizer.zero_grad(set_to_none=True)      # no-op with persistent gradients (pipeline=True)
                loss = self._forward_backward(rays_o, rays_d, gt_rgb, bg_color, counter, mean_count)
                self.optimizer.step()
                return loss
            self.optimizer.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.float16, enabled=self.fp16):
                image, _, _ = self.render_train(rays_o, rays_d, bg_color, counter=counter, mean_count=mean_count)
                loss = torch.nn.functional.mse_loss(image, gt_rgb)
            self.scaler.scale(loss).backward()
            if self.grad_sync is not None:
                self.grad_sync()
            self.scaler.step(self.optimizer)
            self.scaler.update()
            if self.lr_scheduler is not None:
                self.lr_scheduler.step()
            return loss
        finally:
            set_valid_rows(prev)

    def _rows_mode(self):
        """Device-side row counts need every consumer of the sample buffers to honour them: the one-kernel field path with fp16 side
        buffers (the unfused FFMLP kernels would accumulate weight gradients over stale rows beyond the count)."""
        m = self.model
        return (self.scaler is None and self.fused and self.fp16 and not self.pipeline and not self._fuse_rgb and self._wgrad_side is None
                and getattr(m, "_field_fused", False) and getattr(m.sigma_net, "_shadow16", None) is not None
                and getattr(m.color_net, "_shadow16", None) is not None and getattr(m.encoder, "_shadow16", None) is not None
                and __import__("os").environ.get("FOC_ROWS", "1") == "1")

    def _pipelined_body(self, rays_o, rays_d, gt_rgb, bg_color, counter, mean_count):
        """[ optimiser sweep of the PREVIOUS step || near/far + march of this step ] -> field -> composite -> backward."""
        main = torch.cuda.current_stream()
        if self._side_stream is None:
            self._side_stream = torch.cuda.Stream(priority=0)
        side = self._side_stream
        side.wait_stream(main)
        with torch.cuda.stream(side):
            self.optimizer.step()
        return self._forward_backward(rays_o, rays_d, gt_rgb, bg_color, counter, mean_count, after_march=lambda: main.wait_stream(side))

    def grad_tensors(self):
        """Gradient tensors a data-parallel all-reduce must average between backward and the optimiser step."""
        if self.scaler is None:
            return self.optimizer.grad_tensors()
        return [p.grad for p in self.model.parameters() if p.grad is not None]

    def train_step(self, rays_o, rays_d, gt_rgb, bg_color=1.0):
        """One optimisation step on device-resident rays; returns the loss tensor (no host sync)."""
        self.model.train()
        if self.use_graph and self.mean_count > 0:
            return self._graph_step(rays_o, rays_d, gt_rgb, bg_color)
        # detached: a caller holding on to the loss must not keep this step's autograd graph (and its AccumulateGrad nodes,
        # bound to the eager stream) alive into a later graph capture
        return self._step_body(rays_o, rays_d, gt_rgb, bg_color).detach()

    # ---- CUDA-graph replay of the step -------------------------------------------------------------------------------
    GRAPH_QUANTUM = 16384          # sample-buffer sizes are rounded up to this, so the captured shapes rarely change

    def _graph_step(self, rays_o, rays_d, gt_rgb, bg_color):
        """The step is ~90 small launches for 4096 rays; replaying it as one graph removes the launch gaps.  Shapes must
        be static: the sample buffers are sized by the reference's own `mean_count` rule (raymarching.py:198-203), here
        rounded up to GRAPH_QUANTUM so that the 16-step refresh of mean_count only re-captures when it moves by > 16k."""
        q = self.GRAPH_QUANTUM
        rows_mode = self._rows_mode()
        if rows_mode:
            # fixed allocation with 50 % headroom over the running mean: the kernels read the step's real row count from device memory, so
            # the size only matters when a step would not fit (its trailing rays are then dropped, as the reference drops them against its
            # mean-sized buffers, raymarching.py:198-203).  Re-capture only when the mean has grown to within 10 % of the allocation or
            # shrunk below a third of it -- a drifting sample count no longer re-captures the step every few refreshes.
            m_q = ((int(self.mean_count * 1.5) + q - 1) // q) * q
            if self._graph is not None and self.mean_count * 1.1 <= self._graph[0][0] <= max(self.mean_count, 1) * 3 + q:
                m_q = self._graph[0][0]
        else:
            m_q = ((self.mean_count + q - 1) // q) * q
            if self._graph is not None and m_q <= self._graph[0][0] <= m_q + q:       # hysteresis: keep a slightly larger buffer
                m_q = self._graph[0][0]
        pipelined = self.pipeline
        key = (m_q, tuple(rays_o.shape), float(bg_color), pipelined)
        if pipelined and not self._pending:
            # the captured body starts with the sweep of the previous step's gradients: make sure there are some
            loss = self._forward_backward(rays_o, rays_d, gt_rgb, bg_color)
            self._pending = True
            return loss
        if self._graph is None or self._graph[0] != key:
            self._graph = None                                                     # frees the old graph's pool first
            g_in = [torch.empty_like(rays_o), torch.empty_like(rays_d), torch.empty_like(gt_rgb)]
            for dst, src in zip(g_in, (rays_o, rays_d, gt_rgb)):
                dst.copy_(src)
            g_counter = torch.zeros(2, dtype=torch.int32, device=rays_o.device)
            g_rows = torch.full((1,), m_q, dtype=torch.int32, device=rays_o.device) if rows_mode else None
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            if not pipelined:
                self.optimizer.zero_grad(set_to_none=True)
            if self._capture_stream is None:           # the step's own kernels outrank the deferred optimiser sweep
                self._capture_stream = torch.cuda.Stream(priority=-1)
            self.capture_count += 1
            with torch.cuda.graph(graph, stream=self._capture_stream):
                if pipelined:
                    g_loss = self._pipelined_body(g_in[0], g_in[1], g_in[2], bg_color, g_counter, m_q)
                else:
                    g_loss = self._step_body(g_in[0], g_in[1], g_in[2], bg_color, counter=g_counter, mean_count=m_q, rows=g_rows)
            self._graph = (key, graph, g_in, g_counter, g_loss, g_rows)
        _, graph, g_in, g_counter, g_loss = self._graph[:5]
        for dst, src in zip(g_in, (rays_o, rays_d, gt_rgb)):
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        graph.replay()
        self.step_counter[self.local_step % 16].copy_(g_counter)
        self.local_step += 1
        return g_loss

    def train_step_host(self, rays_o_h, rays_d_h, gt_rgb_h, stage):
        """End-to-end step from PINNED HOST buffers: H2D copies of the batch, the step, D2H read of the loss."""
        if self.use_graph and self._graph is not None and self.mean_count > 0 and tuple(self._graph[2][0].shape) == tuple(rays_o_h.shape):
            stage = self._graph[2]                     # straight into the captured step's input buffers: no device-to-device staging copies
        stage[0].copy_(rays_o_h, non_blocking=True)
        stage[1].copy_(rays_d_h, non_blocking=True)
        stage[2].copy_(gt_rgb_h, non_blocking=True)
        loss = self.train_step(stage[0], stage[1], stage[2])
        return float(loss.item())

    def train_step_pixels(self, dataset, pix_h, pose_idx_h, stage):
        """End-to-end step from PIXEL INDICES: the host uploads (pixel, camera) int32 pairs -- 8 bytes per ray instead of 36 -- and
        rays.PixelDataset builds rays and ground truth on the device (csrc/rays.cu), straight into the captured step's input buffers.
        stage = (pix, pose_idx) int32 device buffers.  Returns the loss as a Python float."""
        stage[0].copy_(pix_h, non_blocking=True)
        stage[1].copy_(pose_idx_h, non_blocking=True)
        n = int(pix_h.shape[0])
        out = None
        if self.use_graph and self._graph is not None and self.mean_count > 0 and self._graph[2][0].shape[0] == n:
            out = tuple(self._graph[2][:3])
        o, d, gt = dataset.batch(stage[0], stage[1], out=out)
        return float(self.train_step(o, d, gt).item())

    def fit_step_pixels(self, dataset, pix_h, pose_idx_h, stage):
        if self.global_step % self.update_extra_interval == 0:
            self.update_extra_state()
        self.global_step += 1
        return self.train_step_pixels(dataset, pix_h, pose_idx_h, stage)

    # ---- the reference's training loop body (nerf/utils.py:851-870, train_one_epoch) ---------------------------------------
    def fit_step(self, rays_o, rays_d, gt_rgb, bg_color=1.0):
        """One iteration of train_one_epoch: refresh the density grid when `global_step % update_extra_interval == 0`
        (nerf/utils.py:853-856, under autocast), then the optimisation step; the learning-rate schedule advances inside it."""
        if self.global_step % self.update_extra_interval == 0:
            self.update_extra_state()
        self.global_step += 1
        return self.train_step(rays_o, rays_d, gt_rgb, bg_color)

    def fit_step_host(self, rays_o_h, rays_d_h, gt_rgb_h, stage):
        """fit_step from PINNED HOST buffers, returning the loss as a Python float (the reference's loop reads `loss.item()` every
        step, nerf/utils.py:872)."""
        if self.global_step % self.update_extra_interval == 0:
            self.update_extra_state()
        self.global_step += 1
        return self.train_step_host(rays_o_h, rays_d_h, gt_rgb_h, stage)

    # ---- checkpoint interchange with the reference (nerf/utils.py:1015-1136) ------------------------------------------------
    _RENDERER_KEYS = ("aabb_train", "aabb_infer", "density_grid", "density_bitfield", "step_counter")

    def state_dict(self):
        """The `model` entry of a reference checkpoint: NeRFNetwork IS its renderer there (network_ff.py:10), so its state_dict
        holds the renderer buffers (renderer.py:79-99) next to `encoder.embeddings`, `encoder.offsets`, `sigma_net.weights`,
        `color_net.weights`."""
        self.flush()
        if hasattr(self.optimizer, "sync_master"):
            self.optimizer.sync_master()                 # data-parallel: complete the sharded fp32 table first (a collective)
        sd = {"aabb_train": self.aabb.detach().clone(), "aabb_infer": self.aabb.detach().clone(), "density_grid": self.density_grid.clone(),
              "density_bitfield": self.density_bitfield.clone(), "step_counter": self.step_counter.clone()}
        sd.update({k: v.detach().clone() for k, v in self.model.state_dict().items()})
        return sd

    def save_checkpoint(self, path, full=False):
        """Trainer.save_checkpoint's file format (nerf/utils.py:1020-1053): epoch / global_step / stats / mean_count / mean_density /
        model (+ optimizer when full)."""
        state = {"epoch": 0, "global_step": self.global_step, "stats": {}, "mean_count": self.mean_count,
                 "mean_density": float(self.mean_density.item()), "model": self.state_dict()}
        if full and hasattr(self.optimizer, "state_dict"):
            state["optimizer"] = self.optimizer.state_dict()
        torch.save(state, path)

    def load_checkpoint(self, checkpoint, model_only=True):
        """Load a checkpoint written by the reference's Trainer.save_checkpoint (or by save_checkpoint above; or a bare model
        state_dict, nerf/utils.py:1090-1093), as Trainer.load_checkpoint does (nerf/utils.py:1077-1110): parameters with
        strict=False, the renderer buffers, mean_count / mean_density.  Returns (missing_keys, unexpected_keys)."""
        self.flush()
        if not isinstance(checkpoint, dict):
            checkpoint = torch.load(checkpoint, map_location=self.device, weights_only=False)
        sd = checkpoint["model"] if "model" in checkpoint else checkpoint
        own = {k: v for k, v in sd.items() if k not in self._RENDERER_KEYS}
        missing, unexpected = self.model.load_state_dict(own, strict=False)
        with torch.no_grad():
            if "density_grid" in sd:                      # dropped from `best` checkpoints (nerf/utils.py:1065-1066)
                self.density_grid.copy_(sd["density_grid"].to(self.device).view_as(self.density_grid))
            if "density_bitfield" in sd:
                self.density_bitfield.copy_(sd["density_bitfield"].to(self.device))
            if "step_counter" in sd:
                self.step_counter.copy_(sd["step_counter"].to(self.device))
            if "aabb_train" in sd:
                self.aabb.copy_(sd["aabb_train"].to(self.device))
            if "mean_density" in checkpoint:
                self.mean_density.fill_(float(checkpoint["mean_density"]))
        if "mean_count" in checkpoint:
            self.mean_count = int(checkpoint["mean_count"])
        if "global_step" in checkpoint and not model_only:
            self.global_step = int(checkpoint["global_step"])
        if hasattr(self.optimizer, "refresh_shadows"):
            self.optimizer.refresh_shadows()             # fp16 copies the kernels read (the modules also notice the version bump)
        self._graph = None                                # captured shapes may no longer fit mean_count
        return list(missing), list(unexpected)

    def update_mean_count(self):
        """renderer.py:533-536: mean samples of the last 16 steps (one D2H sync, every 16 steps)."""
        total = min(self.local_step, 16)
        if total > 0:
            mean = self.step_counter[:total, 0].sum().float() / total
            if self.data_parallel is not None:
                # data-parallel: every rank sizes its sample buffers (and decides about a re-capture of the step) from the SAME number, the
                # largest rank's mean -- otherwise the ranks re-capture at different steps and each stall is paid by all of them at the barrier
                import torch.distributed as dist
                dist.all_reduce(mean, op=dist.ReduceOp.MAX, group=self._dp_group)
            self.mean_count = int(mean.item())

    # ------------------------------------------------------------------------------------------------ inference
    @torch.no_grad()
    def _render_rows_ok(self):
        """The speculative single pass needs every consumer of the sample buffers to honour the device-side row count."""
        import os
        m = self.model
        return (self.fused and self.fp16 and getattr(m, "_field_fused", False) and getattr(m.sigma_net, "_shadow16", None) is not None
                and getattr(m.color_net, "_shadow16", None) is not None and getattr(m.encoder, "_shadow16", None) is not None
                and os.environ.get("FOC_RENDER_SPECULATIVE", "1") == "1")

    def _render_size(self, N, total):
        """Capacity for the next N-ray frame: this frame's total + 12.5 %, in 16 384-row quanta; kept while the total stays inside
        [cap / 1.5, cap / 1.03] so that a slowly moving camera does not change the allocation every frame."""
        cap = self._render_cap.get(N)
        if cap is None or total * 1.03 > cap or total * 1.5 < cap:
            self._render_cap[N] = max(16384, -(-int(total * 1.125 + 128) // 16384) * 16384)

    def _render_size_windows(self, key, totals):
        """Per-window capacities of the speculative windowed render: the totals of this frame + 12.5 % (+ 16 384 rows: the late windows
        hold few samples and move more from frame to frame), kept while every total stays inside its band (see _render_size)."""
        caps = self._render_cap.get(key)
        if caps is None or any(t * 1.03 > c or (t + 16384) * 1.5 < c for t, c in zip(totals, caps)):
            self._render_cap[key] = [max(16384, -(-int(t * 1.125 + 16384) // 16384) * 16384) for t in totals]

    def render_frames(self, frames, bg_color=1.0, T_thresh=1e-4, exchange=None, mode="single_pass", windows=(32, 64, 928)):
        """A camera path: `frames` = [(rays_o, rays_d), ...] -> [(image, depth, weights_sum), ...] with NO host read between frames (the
        reference's test loop, legacy/nerf/utils.py test(), pays ~45 reads per frame).  Every frame is enqueued into buffers sized from
        the previous one; the exact totals of all frames come back in one read at the end and any frame whose samples did not fit is
        rendered again with exact sizes.  With an `exchange` the frames are copies (its two buffers alternate)."""
        outs, checks = [], []
        for o, d in frames:
            pend = []
            out = self.render(o, d, bg_color, T_thresh, mode=mode, exchange=exchange, windows=windows, _defer=pend)
            if exchange is not None:
                out = tuple(t.clone() for t in out)
            outs.append(out)
            checks.append(pend[0] if pend else None)
        live = [i for i, c in enumerate(checks) if c is not None]
        if live:
            totals = torch.cat([checks[i][0].reshape(-1) for i in live]).tolist()  # the one host read
            pos = 0
            for i in live:
                caps = checks[i][1] if isinstance(checks[i][1], list) else [checks[i][1]]
                t = totals[pos:pos + len(caps)]
                pos += len(caps)
                N = frames[i][0].reshape(-1, 3).shape[0]
                key = checks[i][2] if len(checks[i]) > 2 else N
                if any(_pad128(a) > c for a, c in zip(t, caps)):
                    self.render_overflows += 1
                    self._render_cap.pop(key, None)
                    out = self.render(frames[i][0], frames[i][1], bg_color, T_thresh, mode=mode, exchange=exchange, windows=windows)
                    outs[i] = tuple(x.clone() for x in out) if exchange is not None else out
                elif i == live[-1]:
                    if isinstance(checks[i][1], list):
                        self._render_size_windows(key, t)
                    else:
                        self._render_size(N, t[0])
        return outs

    def render(self, rays_o, rays_d, bg_color=1.0, T_thresh=1e-4, device_compaction=True, mode="single_pass", exchange=None,
               windows=(48, 976), _defer=None):
        """run_cuda, inference branch (legacy/nerf/renderer.py:331-372).

        mode="loop": the reference's caller-owned loop (march_rays -> field -> composite_rays, n_step = min(N // n_alive, 8)).
          One host read of the alive count per iteration (it sizes the next launches): ~45 iterations of ~15
          launches for an 800x800 frame, host-bound on B200.  device_compaction=True swaps the boolean-mask indexing for
          raymarching.compact_alive.
        mode="single_pass" (default): the same samples and the same front-to-back recurrence, but marched for all rays at once
          into exactly-sized ray-ordered buffers (march_rays_all: one 4-byte host read per frame), one field evaluation, one
          compositor launch with the TRAINING kernel's termination rule (K7: stop once T after a sample < T_thresh; the loop's K10 tests T
          before the sample -- at most one trailing sample per ray, <= T_thresh per pixel).  Samples behind the termination point are evaluated and
          discarded -- the tensor pipe and HBM are far from saturated, the launch/sync chain was the bound.
        mode="windowed": the loop's early termination without its host-driven iterations -- the march is recorded once, then each
          depth window (`windows` = samples per ray: the first 48, then the rest) is expanded for the rays still alive, evaluated and
          composited with K10's recurrence (raymarching.WindowedMarch).  One 4-byte host read per window.
        The epilogue (background mix, depth normalisation) is one kernel writing [r, g, b, depth, weights_sum] rows; the
        returned image / depth / weights_sum are views of that buffer.
        exchange (peer.FrameExchange, multi-GPU): rays_o / rays_d are this rank's share (`exchange.my_rays()`) of the frame; the
          epilogue stores its rows into every rank's frame buffer over NVLink and the views returned cover the WHOLE frame."""
        from . import peer
        self.flush()
        self.model.eval()
        rays_o, rays_d = rays_o.contiguous().view(-1, 3), rays_d.contiguous().view(-1, 3)
        N, dev = rays_o.shape[0], rays_o.device
        nears, fars = raymarching.near_far_from_aabb(rays_o, rays_d, self.aabb, self.min_near)

        def finish(image, depth, weights_sum, from_first):
            if exchange is not None:
                frame = exchange.finish(image, depth, weights_sum, nears, fars, bg_color, from_first)
            else:
                frame, out = peer.local_frame(N, dev)
                peer.render_finish(image, depth, weights_sum, nears, fars, bg_color, from_first, out)
            return frame[:, :3], frame[:, 3], frame[:, 4]

        if mode == "windowed":
            field = self.model.forward_fused if self.fused and self.fp16 else self.model
            widths, start = [], 0
            for width in windows:
                width = min(int(width), self.max_steps - start)
                if width > 0:
                    widths.append((start, width)); start += width
            key = (N, tuple(widths))
            caps = self._render_cap.get(key) if self._render_rows_ok() else None
            if caps is not None:
                # speculative sizing per window (see single_pass below): every window's samples go into buffers of the previous frame's
                # size, the row counts stay on the device, and the exact totals of ALL windows are read once, after the last launch
                from ._lib import set_valid_rows
                wm = raymarching.WindowedMarch(rays_o, rays_d, self.bound, self.density_bitfield, self.cascade, self.grid_size, nears, fars,
                                               self.dt_gamma, self.max_steps)
                counters = torch.zeros(len(widths), 2, dtype=torch.int32, device=dev)
                rows = torch.zeros(len(widths), 1, dtype=torch.int32, device=dev)
                for wi, (st, width) in enumerate(widths):
                    w = wm.window_spec(st, width, caps[wi], counters[wi], rows[wi])
                    prev = set_valid_rows(rows[wi])
                    try:
                        with torch.autocast("cuda", dtype=torch.float16, enabled=self.fp16):
                            sigmas, rgbs = field(w[0], w[1])
                    finally:
                        set_valid_rows(prev)
                    wm.composite(sigmas, rgbs, w[2], T_thresh)
                out = finish(wm.image, wm.depth, wm.weights_sum, False)
                totals = counters[:, 0].contiguous()
                if exchange is not None and exchange.world > 1:           # every rank must take the same decision
                    import torch.distributed as dist
                    dist.all_reduce(totals, op=dist.ReduceOp.MAX, group=exchange.arena.group)
                if _defer is not None:
                    _defer.append((totals, list(caps), key))
                    return out
                tl = totals.tolist()
                fits = all(_pad128(t) <= c for t, c in zip(tl, caps))
                self._render_size_windows(key, tl)
                if fits:
                    return out
                self.render_overflows += 1                              # fall through: exact sizes, one read per window
            wm = raymarching.WindowedMarch(rays_o, rays_d, self.bound, self.density_bitfield, self.cascade, self.grid_size, nears, fars,
                                           self.dt_gamma, self.max_steps)
            tl = []
            for st, width in widths:
                w = wm.window(st, width)
                tl.append(0 if w is None else w[0].shape[0])
                if w is None:
                    break
                with torch.autocast("cuda", dtype=torch.float16, enabled=self.fp16):
                    sigmas, rgbs = field(w[0], w[1])
                wm.composite(sigmas, rgbs, w[2], T_thresh)
            tl += [0] * (len(widths) - len(tl))
            if self._render_rows_ok():
                if exchange is not None and exchange.world > 1:
                    import torch.distributed as dist
                    tt = torch.tensor(tl, dtype=torch.int32, device=dev)
                    dist.all_reduce(tt, op=dist.ReduceOp.MAX, group=exchange.arena.group)
                    tl = tt.tolist()
                self._render_size_windows(key, tl)
            return finish(wm.image, wm.depth, wm.weights_sum, False)
        if mode == "single_pass":
            field = self.model.forward_fused if self.fused and self.fp16 else self.model
            cap = self._render_cap.get(N) if self._render_rows_ok() else None
            if cap is not None:
                # Speculative sizing: the whole frame is enqueued into buffers of the PREVIOUS frame's size (+12.5 %) with the row count
                # published on the device (march_rays_train_rows -> _lib.valid_rows consumers), and the one 4-byte read of the exact total
                # happens AFTER the last launch -- the GPU never waits for the host inside a frame.  A total beyond the capacity (rays
                # dropped exactly as raymarching.cu:415-416 drops them) re-renders the frame below with exact sizes, so the result never
                # depends on the guess.  defer=True hands the check to the caller (render_frames: one read for a whole camera path).
                from ._lib import set_valid_rows
                counter = torch.zeros(2, dtype=torch.int32, device=dev)
                rows = torch.zeros(1, dtype=torch.int32, device=dev)
                prev = set_valid_rows(rows)
                try:
                    xyzs, dirs, deltas, rays = raymarching.march_rays_train(
                        rays_o, rays_d, self.bound, self.density_bitfield, self.cascade, self.grid_size, nears, fars, counter, cap,
                        False, 128, False, self.dt_gamma, self.max_steps)
                    with torch.autocast("cuda", dtype=torch.float16, enabled=self.fp16):
                        sigmas, rgbs = field(xyzs, dirs)
                finally:
                    set_valid_rows(prev)
                with torch.autocast("cuda", dtype=torch.float16, enabled=self.fp16):
                    weights_sum, depth, image = raymarching.composite_rays_train(sigmas, rgbs, deltas, rays, T_thresh)
                out = finish(image.float().contiguous(), depth.float().contiguous(), weights_sum.float().contiguous(), True)
                total = counter[:1]
                if exchange is not None and exchange.world > 1:           # every rank must take the same decision: the frame barrier is collective
                    import torch.distributed as dist
                    total = total.clone()
                    dist.all_reduce(total, op=dist.ReduceOp.MAX, group=exchange.arena.group)
                if _defer is not None:
                    _defer.append((total, cap))
                    return out
                t = int(total.item())
                self._render_size(N, t)
                if _pad128(t) <= cap:
                    return out
                self.render_overflows += 1                              # fall through: exact sizes, one read in the middle
            xyzs, dirs, deltas, rays = raymarching.march_rays_all(rays_o, rays_d, self.bound, self.density_bitfield, self.cascade,
                                                                  self.grid_size, nears, fars, 128, self.dt_gamma, self.max_steps)
            if exchange is not None and exchange.world > 1 and self._render_rows_ok():
                import torch.distributed as dist
                m = torch.tensor([xyzs.shape[0]], dtype=torch.int32, device=dev)
                dist.all_reduce(m, op=dist.ReduceOp.MAX, group=exchange.arena.group)
                self._render_size(N, int(m.item()))
            else:
                self._render_size(N, xyzs.shape[0])
            with torch.autocast("cuda", dtype=torch.float16, enabled=self.fp16):
                sigmas, rgbs = field(xyzs, dirs)
                weights_sum, depth, image = raymarching.composite_rays_train(sigmas, rgbs, deltas, rays, T_thresh)
            # the training compositor measures depth from the first sample's start (raymarching.cu:545 starts t at 0), the
            # inference one from the ray origin (raymarching.cu:858 starts at rays_t): the epilogue adds sum_i w_i * near
            return finish(image.float().contiguous(), depth.float().contiguous(), weights_sum.float().contiguous(), True)
        weights_sum = torch.zeros(N, dtype=torch.float32, device=dev)
        depth = torch.zeros(N, dtype=torch.float32, device=dev)
        image = torch.zeros(N, 3, dtype=torch.float32, device=dev)
        n_alive = N
        rays_alive = torch.arange(n_alive, dtype=torch.int32, device=dev)
        rays_t = nears.clone()
        step = 0
        with torch.autocast("cuda", dtype=torch.float16, enabled=self.fp16):
            while step < self.max_steps:
                n_alive = rays_alive.shape[0]
                if n_alive <= 0:
                    break
                n_step = max(min(N // n_alive, 8), 1)
                xyzs, dirs, deltas = raymarching.march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, self.bound,
                                                            self.density_bitfield, self.cascade, self.grid_size, nears, fars,
                                                            128, False, self.dt_gamma, self.max_steps)
                sigmas, rgbs = self.model(xyzs, dirs)
                raymarching.composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image,
                                           T_thresh)
                if device_compaction:
                    out, n_out = raymarching.compact_alive(rays_alive, n_alive)
                    rays_alive = out[:int(n_out.item())]
                else:
                    rays_alive = rays_alive[rays_alive >= 0]
                step += n_step
        return finish(image, depth, weights_sum, False)
