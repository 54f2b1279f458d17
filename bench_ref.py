"""Reference arm of bench.py: the REFERENCE's own CUDA extensions (oracle/_ref/*.so, compiled unmodified from
/root/reference by oracle/build_ref.py) timed on the same B200, same workload, same train step.

The reference has no CPU implementation of this path (every operator `.cuda()`s its inputs), and its Python packages
cannot travel to the GPU box (/root/reference does not exist there; ffmlp/ffmlp.py:2 imports tkinter's `turtle`).  So the
L0 kernels are the reference's, unmodified, and the thin L1 autograd wrappers are restated here with the reference's
allocation / copy behaviour, which is part of what it costs per step:
  * gridencoder/grid.py:24-89   -- [L,B,C] kernel output + permute/reshape copy; backward permutes grad to [L,B,C], zero-fills
                                   grad_embeddings;
  * ffmlp/ffmlp.py:15-83,147-169 -- always-pad by 128 (torch.cat), forward_buffer, zero-filled backward_buffer / grad_weights,
                                   split-K side streams (allocate_splitk);
  * raymarching/raymarching.py:161-291 -- zero-filled xyzs/dirs/deltas, composite fwd/bwd with zero-filled grads.
The SH-4 direction encoder is absent from the reference tree; both arms use the same kernel for it.
None of this file is imported by the product; bench.py --impl reference is its only caller.
"""
import math
import os
import sys

import numpy as np
import torch
import torch.nn as nn
from torch.autograd import Function
from torch.amp import custom_bwd, custom_fwd

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import ref_ext  # noqa: E402

_rm = _ge = _ff = None


def _load():
    global _rm, _ge, _ff
    _rm, _ge, _ff = ref_ext.load("raymarching"), ref_ext.load("gridencoder"), ref_ext.load("ffmlp")
    if _rm is None or _ge is None or _ff is None:
        raise RuntimeError("oracle/_ref extensions not present or no CUDA device")


class _grid_encode(Function):
    @staticmethod
    @custom_fwd(device_type="cuda")
    def forward(ctx, inputs, embeddings, offsets, per_level_scale, base_resolution):
        inputs = inputs.contiguous()
        B, D = inputs.shape
        L, C = offsets.shape[0] - 1, embeddings.shape[1]
        S, H = np.log2(per_level_scale), base_resolution
        if torch.is_autocast_enabled() and C % 2 == 0:
            embeddings = embeddings.to(torch.half)
        outputs = torch.empty(L, B, C, device=inputs.device, dtype=embeddings.dtype)
        _ge.grid_encode_forward(inputs, embeddings, offsets, outputs, B, D, C, L, S, H, None, 0, False, 0)
        outputs = outputs.permute(1, 0, 2).reshape(B, L * C)
        ctx.save_for_backward(inputs, embeddings, offsets)
        ctx.dims = [B, D, C, L, S, H]
        return outputs

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, grad):
        inputs, embeddings, offsets = ctx.saved_tensors
        B, D, C, L, S, H = ctx.dims
        grad = grad.view(B, L, C).permute(1, 0, 2).contiguous()
        grad_embeddings = torch.zeros_like(embeddings)
        _ge.grid_encode_backward(grad, inputs, embeddings, offsets, grad_embeddings, B, D, C, L, S, H, None, None, 0, False, 0)
        return None, grad_embeddings, None, None, None


class _ffmlp(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.half)
    def forward(ctx, inputs, weights, input_dim, output_dim, hidden_dim, num_layers, calc_grad_inputs):
        B = inputs.shape[0]
        inputs, weights = inputs.contiguous(), weights.contiguous()
        outputs = torch.empty(B, output_dim, device=inputs.device, dtype=inputs.dtype)
        forward_buffer = torch.empty(num_layers, B, hidden_dim, device=inputs.device, dtype=inputs.dtype)
        _ff.ffmlp_forward(inputs, weights, B, input_dim, output_dim, hidden_dim, num_layers, 0, 6, forward_buffer, outputs)
        ctx.save_for_backward(inputs, weights, outputs, forward_buffer)
        ctx.dims = (input_dim, output_dim, hidden_dim, num_layers, calc_grad_inputs)
        return outputs

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, grad):
        B = grad.shape[0]
        grad = grad.contiguous()
        inputs, weights, outputs, forward_buffer = ctx.saved_tensors
        input_dim, output_dim, hidden_dim, num_layers, calc_grad_inputs = ctx.dims
        grad_inputs = torch.zeros_like(inputs) if calc_grad_inputs else torch.zeros(1, device=grad.device, dtype=grad.dtype)
        grad_weights = torch.zeros_like(weights)
        backward_buffer = torch.zeros(num_layers, B, hidden_dim, device=grad.device, dtype=grad.dtype)
        _ff.ffmlp_backward(grad, inputs, weights, forward_buffer, B, input_dim, output_dim, hidden_dim, num_layers, 0, 6,
                           calc_grad_inputs, backward_buffer, grad_inputs, grad_weights)
        return (grad_inputs if calc_grad_inputs else None), grad_weights, None, None, None, None, None


class _composite(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, sigmas, rgbs, deltas, rays, T_thresh=1e-4):
        sigmas, rgbs = sigmas.contiguous(), rgbs.contiguous()
        M, N = sigmas.shape[0], rays.shape[0]
        ws = torch.empty(N, dtype=sigmas.dtype, device=sigmas.device)
        depth = torch.empty(N, dtype=sigmas.dtype, device=sigmas.device)
        image = torch.empty(N, 3, dtype=sigmas.dtype, device=sigmas.device)
        _rm.composite_rays_train_forward(sigmas, rgbs, deltas, rays, M, N, T_thresh, ws, depth, image)
        ctx.save_for_backward(sigmas, rgbs, deltas, rays, ws, depth, image)
        ctx.dims = [M, N, T_thresh]
        return ws, depth, image

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, g_ws, g_depth, g_image):
        g_ws, g_image = g_ws.contiguous(), g_image.contiguous()
        sigmas, rgbs, deltas, rays, ws, depth, image = ctx.saved_tensors
        M, N, T_thresh = ctx.dims
        gs, gc = torch.zeros_like(sigmas), torch.zeros_like(rgbs)
        _rm.composite_rays_train_backward(g_ws, g_image, sigmas, rgbs, deltas, rays, ws, image, M, N, T_thresh, gs, gc)
        return gs, gc, None, None, None


class RefFFMLP(nn.Module):
    def __init__(self, input_dim, output_dim, hidden_dim, num_layers):
        super().__init__()
        self.input_dim, self.output_dim, self.hidden_dim, self.num_layers = input_dim, output_dim, hidden_dim, num_layers
        self.padded_output_dim = int(math.ceil(output_dim / 16)) * 16
        n = hidden_dim * (input_dim + hidden_dim * (num_layers - 1) + self.padded_output_dim)
        self.weights = nn.Parameter(torch.zeros(n))
        torch.manual_seed(42)
        self.weights.data.uniform_(-math.sqrt(3 / hidden_dim), math.sqrt(3 / hidden_dim))
        _ff.allocate_splitk(num_layers + 1)

    def forward(self, inputs):
        B, C = inputs.shape
        pad = 128 - (B % 128)
        inputs = torch.cat([inputs, torch.zeros(pad, C, dtype=inputs.dtype, device=inputs.device)], dim=0)
        out = _ffmlp.apply(inputs, self.weights, self.input_dim, self.padded_output_dim, self.hidden_dim, self.num_layers,
                           inputs.requires_grad)
        return out[:B, :self.output_dim]


class RefNetwork(nn.Module):
    """nerf/network_ff.py:10-74 on the reference extensions."""

    def __init__(self, bound=1.0):
        super().__init__()
        from focnerf_b200.gridencoder import GridEncoder      # used only for its offsets / per_level_scale / init (no kernels)
        g = GridEncoder(desired_resolution=2048 * bound)
        self.bound, self.pls, self.H = bound, g.per_level_scale, g.base_resolution
        self.register_buffer("offsets", g.offsets.clone())
        self.embeddings = nn.Parameter(g.embeddings.data.clone())
        self.sigma_net = RefFFMLP(32, 16, 64, 2)
        self.color_net = RefFFMLP(32, 3, 64, 3)

    def forward(self, x, d):
        from focnerf_b200.activation import trunc_exp
        from focnerf_b200.shencoder import sh_encode
        x = (x + self.bound) / (2 * self.bound)
        h = self.sigma_net(_grid_encode.apply(x.view(-1, 3), self.embeddings, self.offsets, self.pls, self.H))
        sigma = trunc_exp(h[..., 0])
        geo = h[..., 1:]
        dd = sh_encode(d)
        h = torch.cat([dd.to(geo.dtype), geo, torch.zeros_like(geo[..., :1])], dim=-1)
        return sigma, torch.sigmoid(self.color_net(h))


def _ref_density(self, x):
    """nerf/network_ff.py:76-90"""
    from focnerf_b200.activation import trunc_exp
    x = (x + self.bound) / (2 * self.bound)
    h = self.sigma_net(_grid_encode.apply(x.view(-1, 3), self.embeddings, self.offsets, self.pls, self.H))
    return {'sigma': trunc_exp(h[..., 0]), 'geo_feat': h[..., 1:]}


def _ref_color(self, x, d, mask=None, geo_feat=None, **kwargs):
    """nerf/network_ff.py:92-123: colour net on the masked subset (boolean gather / scatter)"""
    from focnerf_b200.shencoder import sh_encode
    if mask is not None:
        rgbs = torch.zeros(mask.shape[0], 3, dtype=x.dtype, device=x.device)
        if not mask.any():
            return rgbs
        x, d, geo_feat = x[mask], d[mask], geo_feat[mask]
    dd = sh_encode(d)
    h = torch.cat([dd.to(geo_feat.dtype), geo_feat, torch.zeros_like(geo_feat[..., :1])], dim=-1)
    h = torch.sigmoid(self.color_net(h))
    if mask is not None:
        rgbs[mask] = h.to(rgbs.dtype)
        return rgbs
    return h


RefNetwork.density = _ref_density
RefNetwork.color = _ref_color


def multi_bench(dev, chunks=3, K=4, T=512, batch=4096):
    """configs[3] on the reference stack: COMBINED.py's per-image loop body (tests/foc_eager.py restates it op for op) over
    RefNetwork objects, in the reference's own 4096-ray batches; a bounded sample of the 640x480 frame (75 batches)."""
    import bench
    import foc_eager

    def near_far(o, d, aabb, min_near):
        n = o.shape[0]
        nears, fars = torch.empty(n, device=o.device), torch.empty(n, device=o.device)
        _rm.near_far_from_aabb(o.contiguous(), d.contiguous(), aabb, n, min_near, nears, fars)
        return nears, fars

    models = bench.multi_objects(lambda: RefNetwork(1.0), K, dev)
    aabb = torch.tensor([-1.0] * 3 + [1.0] * 3, device=dev)
    o, d = bench.multi_rays(device=dev)
    n = o.shape[0]
    starts = [(n // 2 // batch) * batch + i * batch for i in range(chunks + 1)]      # batches from the middle of the frame
    foc_eager.render_chunk(models, near_far, o[starts[0]:starts[0] + batch], d[starts[0]:starts[0] + batch], aabb, T)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s0 in starts[1:]:
        out = foc_eager.render_chunk(models, near_far, o[s0:s0 + batch], d[s0:s0 + batch], aabb, T)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / chunks
    return {"metric": "640x480 multi-object composite Mrays/s", "value": batch / (ms * 1e-3) / 1e6, "unit": "Mrays/s",
            "ms_per_frame": ms * n / batch, "rays": n, "objects": K, "samples_per_ray": T,
            "field_evaluations_per_s": batch * K * T / (ms * 1e-3), "coverage": float((out[2] > 0.5).float().mean().item()),
            "sample": f"{chunks} of the frame's {n // batch} batches of {batch} rays (the reference's max_ray_batch), objects resident "
                      "(the reference also reloads every checkpoint from disk per image: not timed)",
            "workload": "configs[3] on the reference extensions: eager COMBINED.py loop body (linspace / cumprod / masked colour / running max)"}


class RefTrainer:
    """Same step as focnerf_b200.trainer.Trainer.train_step (legacy/nerf/renderer.py:268-319 + utils.py:510-560, 863-868)."""

    def __init__(self, density_grid, bound=1.0, lr=1e-2):
        self.bound, self.cascade = bound, 1 + math.ceil(math.log2(bound))
        dev = torch.device("cuda")
        torch.manual_seed(0)
        self.model = RefNetwork(bound).to(dev)
        self.aabb = torch.tensor([-bound] * 3 + [bound] * 3, dtype=torch.float32, device=dev)
        g = density_grid.to(dev, torch.float32).contiguous()
        self.bitfield = torch.empty(g.numel() // 8, dtype=torch.uint8, device=dev)
        _rm.packbits(g.view(-1), g.numel() // 8, min(float(g.clamp(min=0).mean()), 10.0), self.bitfield)
        self.step_counter = torch.zeros(16, 2, dtype=torch.int32, device=dev)
        self.mean_count, self.local_step = 0, 0
        self.optimizer = torch.optim.Adam(self.model.parameters(), lr=lr, betas=(0.9, 0.99), eps=1e-15, fused=True)
        self.scaler = torch.amp.GradScaler("cuda")

    def train_step(self, rays_o, rays_d, gt):
        N, dev = rays_o.shape[0], rays_o.device
        self.optimizer.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.float16):
            nears, fars = torch.empty(N, device=dev), torch.empty(N, device=dev)
            _rm.near_far_from_aabb(rays_o, rays_d, self.aabb, N, 0.2, nears, fars)
            counter = self.step_counter[self.local_step % 16]
            counter.zero_()
            self.local_step += 1
            M = N * 1024
            if self.mean_count > 0:
                M = self.mean_count + (128 - self.mean_count % 128)
            xyzs, dirs = torch.zeros(M, 3, device=dev), torch.zeros(M, 3, device=dev)
            deltas = torch.zeros(M, 2, device=dev)
            rays = torch.empty(N, 3, dtype=torch.int32, device=dev)
            noises = torch.rand(N, device=dev)
            _rm.march_rays_train(rays_o, rays_d, self.bitfield, self.bound, 0.0, 1024, N, self.cascade, 128, M, nears, fars, xyzs, dirs,
                                 deltas, rays, counter, noises)
            if self.mean_count <= 0:
                m = int(counter[0].item())
                m += 128 - m % 128
                xyzs, dirs, deltas = xyzs[:m], dirs[:m], deltas[:m]
            sigmas, rgbs = self.model(xyzs, dirs)
            ws, depth, image = _composite.apply(sigmas, rgbs, deltas, rays)
            image = image + (1 - ws).unsqueeze(-1)
            loss = torch.nn.functional.mse_loss(image, gt)
        self.scaler.scale(loss).backward()
        self.scaler.step(self.optimizer)
        self.scaler.update()
        return loss

    def update_mean_count(self):
        total = min(self.local_step, 16)
        self.mean_count = int(self.step_counter[:total, 0].sum().item() / total)

    @torch.no_grad()
    def render(self, rays_o, rays_d, T_thresh=1e-4):
        """run_cuda, inference branch (legacy/nerf/renderer.py:331-372) on the reference extensions."""
        N, dev = rays_o.shape[0], rays_o.device
        nears, fars = torch.empty(N, device=dev), torch.empty(N, device=dev)
        _rm.near_far_from_aabb(rays_o, rays_d, self.aabb, N, 0.2, nears, fars)
        ws, depth, image = torch.zeros(N, device=dev), torch.zeros(N, device=dev), torch.zeros(N, 3, device=dev)
        n_alive = N
        rays_alive = torch.arange(n_alive, dtype=torch.int32, device=dev)
        rays_t = nears.clone()
        step = 0
        with torch.autocast("cuda", dtype=torch.float16):
            while step < 1024:
                n_alive = rays_alive.shape[0]
                if n_alive <= 0:
                    break
                n_step = max(min(N // n_alive, 8), 1)
                M = n_alive * n_step
                M += 128 - M % 128 if M % 128 else 0
                xyzs, dirs, deltas = torch.zeros(M, 3, device=dev), torch.zeros(M, 3, device=dev), torch.zeros(M, 2, device=dev)
                noises = torch.zeros(n_alive, device=dev)
                _rm.march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, self.bound, 0.0, 1024, self.cascade, 128, self.bitfield,
                               nears, fars, xyzs, dirs, deltas, noises)
                sigmas, rgbs = self.model(xyzs, dirs)
                _rm.composite_rays(n_alive, n_step, T_thresh, rays_alive, rays_t, sigmas.float().contiguous(), rgbs.float().contiguous(),
                                   deltas, ws, depth, image)
                rays_alive = rays_alive[rays_alive >= 0]
                step += n_step
        return image + (1 - ws).unsqueeze(-1), depth, ws


def run(args):
    """-> the JSON dict for `bench.py --impl reference` (rank 0, one GPU: the reference has no multi-GPU path)."""
    import json  # noqa: F401
    import cases
    import bench
    _load()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    tr = RefTrainer(torch.from_numpy(grid))
    assert np.array_equal(tr.bitfield.cpu().numpy(), bits), "reference packbits disagrees with the oracle"
    pool_dev, pool_host = bench.make_pool(scene, bench.POOL, bench.RAYS_PER_GPU, 1000, dev)
    stage = [torch.empty_like(t) for t in pool_dev[0]]

    def loop(n, host, first):
        for i in range(n):
            j = (first + i) % bench.POOL
            if host:
                for s, h in zip(stage, pool_host[j]):
                    s.copy_(h, non_blocking=True)
                float(tr.train_step(*stage).item())
            else:
                tr.train_step(*pool_dev[j])
            if tr.local_step % 16 == 0:
                tr.update_mean_count()

    loop(16, False, 0)
    warm = max(3, args.warmup)
    loop(warm, False, 16)
    clocks = bench.ClockSampler(0)
    clocks.start()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loop(args.steps, False, 100)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clk = clocks.stop()
    samples = float(tr.step_counter[:, 0].float().mean().item())
    loop(3, True, 200)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    loop(args.steps, True, 300)
    t1.record()
    torch.cuda.synchronize()
    ms_e2e = t0.elapsed_time(t1)
    rays = bench.RAYS_PER_GPU * args.steps
    val, val_e2e = rays / (ms * 1e-3), rays / (ms_e2e * 1e-3)
    render = bench.render_bench(tr.render, dev, 0, 1, frames=args.frames) if args.frames > 0 else None
    multi = multi_bench(dev) if getattr(args, "multi_frames", 0) > 0 else None
    quality = bench.psnr_run(lambda: RefTrainer(torch.from_numpy(grid)), scene, dev, args.psnr_steps, pool_dev) if args.psnr_steps > 0 else None
    return {
        "impl": "reference", "metric": "train rays/s", "value": val, "unit": "rays/s", "n_gpus": 1, "steps": args.steps, "warmup": warm,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {"workload": "configs[1]: single-object hash-grid NeRF train step (16 levels, 2^19 table, F=2, 64-wide FFMLPs, "
                               "4096 rays/batch per GPU, cuda_ray, fp16, Adam)", "rays_per_gpu": bench.RAYS_PER_GPU, "samples_per_step": samples,
                   "note": "the reference's own CUDA extensions (raymarching, gridencoder, ffmlp: unmodified sources compiled for sm_100a) on the "
                           "same B200; the reference has no CPU implementation of this path and no multi-GPU path, so this arm always runs on "
                           "one GPU"},
        "clocks": clk, "render": render, "multi_object": multi, "quality": quality,
        "cpu_baseline": {"value": val, "unit": "rays/s", "cores": 0, "kind": "reference",
                         "sample": f"{args.steps} full train steps of 4096 rays on the reference CUDA extensions (GPU; the reference has no CPU path)"},
        "e2e": {"value": val_e2e, "unit": "rays/s", "ms_per_step": ms_e2e / args.steps, "h2d_bytes_per_step": bench.RAYS_PER_GPU * 36,
                "d2h_bytes_per_step": 4},
    }
