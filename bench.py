"""bench.py -- train rays/s of the FOC-NeRF per-ray hot path (BASELINE.json configs[1]).

Workload: single-object hash-grid NeRF train step -- 16 levels, 2^19 table, F=2, 64-wide fused MLPs, 4096 rays/batch per
GPU, cuda_ray occupancy marching, fp16 autocast, MSE, GradScaler, Adam -- on the synthetic table-top scene (analytic
ground truth; there is no dataset in the reference tree).  A "step" = march -> encode -> MLPs -> composite -> loss ->
backward -> optimizer on one fresh batch of rays.

  python bench.py --gpus N --steps K --warmup W              (N>1 under torchrun, one rank per GPU, NCCL grad all-reduce)
  python bench.py --impl reference ...                       the reference's own CUDA extensions (oracle/_ref) on the
                                                             same GPU, same config (falls back to the CPU port)
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

RAYS_PER_GPU = 4096
POOL = 32                      # distinct pre-generated ray batches cycled through (fresh rays every step)
HBM_FALLBACK_GBS, TC_FALLBACK_TFLOPS = 6650.0, 1590.0


def peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index),
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def make_pool(scene, n_batches, rays, seed, device):
    """Pre-generated batches: device copies for `value`, pinned host copies for `e2e`."""
    import cases
    import torch
    dev, host = [], []
    for i in range(n_batches):
        o, d = cases.camera_rays(rays, seed + i)
        o, d = torch.from_numpy(o).to(device), torch.from_numpy(d).to(device)
        gt = scene.render_gt(o, d)
        dev.append((o, d, gt))
        host.append(tuple(t.cpu().pin_memory() for t in (o, d, gt)))
    return dev, host


# algorithmic traffic per SAMPLE of each timed entry point (SURVEY.md 8d / DESIGN.md "algorithmic bytes")
ALG_BYTES = {
    "foc_grid_encode_forward": 12 + 512 + 64,          # coords + 16 levels x 8 corners x 4 B + fp16 [B,32] out
    "foc_grid_encode_backward": 12 + 64 + 512,         # coords + fp16 grad + 16 x 8 x 4 B atomic targets
    "foc_composite_rays_train_forward": 24,
    "foc_composite_rays_train_backward": 24 + 16,
    "foc_composite_mse_forward": 24,
    "foc_composite_mse_backward": 24 + 16,
    "foc_march_rays_train": 32,
}
TIMED = list(ALG_BYTES) + ["foc_ffmlp_forward", "foc_ffmlp_backward"]
# DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per CALL of each entry point, from the round-1 `ncu --set full`
# capture of this same workload (profiles/r1_kernels.md; two-kernel entry points summed, two-net entry points averaged)
NCU_TRAFFIC = {
    "foc_ffmlp_backward": 78.8e6, "foc_ffmlp_forward": 16.8e6, "foc_grid_encode_backward": 42.9e6, "foc_grid_encode_forward": 21.5e6,
    "foc_march_rays_train": 1.4e6, "foc_composite_rays_train_forward": 3.9e6, "foc_composite_rays_train_backward": 3.9e6,
}


def cpu_baseline(scene, bits, cascade, rays=1024):
    """The oracle port of the same train step on the host cores (OpenMP), bounded sample: `rays` rays of one batch."""
    import numpy as np
    import torch
    import cases
    from oracle import cpu_step
    st = cpu_step.CpuState()
    o, d = cases.camera_rays(rays, 999)
    gt = scene.render_gt(torch.from_numpy(o), torch.from_numpy(d)).numpy()
    nz = np.random.RandomState(0).rand(rays).astype(np.float32)
    cpu_step.train_step(st, o[:128], d[:128], gt[:128], bits, cascade, nz[:128])          # warm-up (builds liborc, page-in)
    t = time.perf_counter()
    cpu_step.train_step(st, o, d, gt, bits, cascade, nz)
    dt = time.perf_counter() - t
    return {"value": rays / dt, "unit": "rays/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"1 train step (fwd+bwd+Adam) on {rays} rays of the same scene, oracle C port with OpenMP, {dt:.1f} s"}


def psnr_run(make_trainer, scene, dev, steps, pool_dev):
    """Fixed-iteration quality check (north_star: PSNR within 0.1 dB of the reference): a FRESH trainer, `steps` train steps
    over the same batch pool, then PSNR of the inference render against the analytic ground truth on 32768 held-out rays."""
    import cases
    import torch
    tr = make_trainer()
    for i in range(steps):
        tr.train_step(*pool_dev[i % len(pool_dev)])
        if tr.local_step % 16 == 0:
            tr.update_mean_count()
    o, d = cases.camera_rays(32768, 424242)
    o, d = torch.from_numpy(o).to(dev), torch.from_numpy(d).to(dev)
    gt = scene.render_gt(o, d)
    image = tr.render(o, d)[0]
    mse = float(torch.mean((image.float() - gt) ** 2).item())
    return {"psnr_db": -10.0 * math.log10(max(mse, 1e-12)), "steps": steps, "eval_rays": 32768,
            "note": "fresh model, fixed number of steps on the benchmark's batch pool, held-out rays of the synthetic scene"}


def render_bench(render_fn, dev, rank, world, frames=3, H=800, W=800):
    """configs[2]: 800x800 inference render, single pass with early termination.  N > 1: the frame's rays are dealt to the ranks
    in 128-ray blocks round-robin and every rank's finished rows are stored straight into every rank's frame buffer over NVLink
    peer memory by the render epilogue (peer.FrameExchange; FOC_RENDER_GATHER=nccl selects all_gather + un-deal copy instead).
    -> {"value": Mrays/s whole job, ...} (same dict on every rank)."""
    import torch
    import torch.distributed as dist
    from focnerf_b200 import synthetic
    from focnerf_b200.dist import cyclic_tile, gather_cyclic
    pose = synthetic.ring_poses(8, radius=2.2, seed=123)[3]
    o, d = synthetic.get_rays(pose, synthetic.intrinsics(H, W), H, W)
    n = H * W
    via_peer = world > 1 and os.environ.get("FOC_RENDER_GATHER", "peer") == "peer"
    ex = None
    if via_peer:
        from focnerf_b200.peer import FrameExchange
        ex = FrameExchange(n, device=dev)
        idx = ex.my_rays()
        o, d = o[idx], d[idx]
    elif world > 1:                                  # 128-ray blocks dealt round-robin: balances the samples, not just the rays
        idx = cyclic_tile(n, rank, world)
        o, d = o[idx], d[idx]
    o, d = o.contiguous().to(dev), d.contiguous().to(dev)

    buf = {"out": None}
    brk = os.environ.get("FOC_RENDER_BREAKDOWN") == "1"

    def frame():
        """-> (image [n,3], depth [n], weights_sum [n]) of the WHOLE frame on this rank"""
        t0 = time.perf_counter()
        if via_peer:
            return render_fn(o, d, exchange=ex)
        image, depth, ws = render_fn(o, d)
        if world == 1:
            return image, depth, ws
        out = torch.cat([image, depth.unsqueeze(-1), ws.unsqueeze(-1)], -1)
        if brk:
            torch.cuda.synchronize(); t1 = time.perf_counter()
        out, buf["out"] = gather_cyclic(out, n, out=buf["out"])
        if brk:
            torch.cuda.synchronize(); t2 = time.perf_counter()
            if rank == 0:
                print(f"[render breakdown] render {1e3 * (t1 - t0):.2f} ms, gather {1e3 * (t2 - t1):.2f} ms", file=sys.stderr)
        return out[:, :3], out[:, 3], out[:, 4]

    frame()                                          # warm-up (allocator, autotune-free: just first-touch)
    if brk:                                          # one profiled frame: device time per kernel, to stderr (rank 0 prints)
        from torch.profiler import ProfilerActivity, profile
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            frame()
            torch.cuda.synchronize()
        if rank == 0:
            print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=60), file=sys.stderr)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(frames):
        out = frame()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item()) / frames
    return {"metric": "800x800 render Mrays/s", "value": n / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_frame": ms, "frames": frames,
            "rays": n, "scaling": "strong", "coverage": float((out[2] > 0.5).float().mean().item()),
            "exchange": "peer-memory stores in the render epilogue" if via_peer else ("all_gather" if world > 1 else None),
            "workload": "configs[2]: one 800x800 test pose of the synthetic scene, T_thresh 1e-4, max_steps 1024, model as trained by the timed steps"}


def multi_rays(H=480, W=640, device="cpu"):
    """configs[3]: the 640x480 table-top view (fx = fy = 554, radius 1.8) -> rays_o, rays_d [H*W, 3]"""
    import torch
    from focnerf_b200 import synthetic
    pose = synthetic.ring_poses(8, radius=1.8, seed=321)[2]
    intr = (554.0, 554.0, W / 2.0, H / 2.0)
    o, d = synthetic.get_rays(pose, intr, H, W)
    return o.contiguous().to(device), d.contiguous().to(device)


def multi_objects(make_model, K, dev):
    """K independently seeded object fields (the workload does not depend on the weights: the reference composites with no early
    termination); tables drawn wide so densities differ from place to place."""
    import torch
    models = []
    for k in range(K):
        m = make_model().to(dev)
        torch.manual_seed(4000 + k)                  # after construction: FFMLP.reset_parameters re-seeds the global generator (ffmlp.py:139)
        with torch.no_grad():
            emb = m.encoder.embeddings if hasattr(m, "encoder") else m.embeddings
            emb.copy_(torch.randn_like(emb) * 1.5)
        models.append(m.eval())
    return models


def multi_bench(dev, rank, world, frames=2, K=4, T=512, H=480, W=640, chunk=16384):
    """configs[3]: FOC multi-object composition (COMBINED.py), K = 4 resident object hash grids, 640x480 rays, T = 512 uniform
    samples per ray and object, arg-max-density selection + per-object colour masks + 4-channel composite.  N > 1: 128-ray
    blocks dealt round-robin, no data-path collective, one gather per frame."""
    import torch
    import torch.distributed as dist
    from focnerf_b200.multiobject import MultiObjectRenderer
    from focnerf_b200.network import NeRFNetwork
    from focnerf_b200.dist import cyclic_tile, gather_cyclic
    o, d = multi_rays(H, W)
    n = H * W
    if world > 1:
        idx = cyclic_tile(n, rank, world)
        o, d = o[idx], d[idx]
    o, d = o.contiguous().to(dev), d.contiguous().to(dev)
    aabb = torch.tensor([-1.0] * 3 + [1.0] * 3, device=dev)
    r = MultiObjectRenderer(multi_objects(lambda: NeRFNetwork(bound=1.0), K, dev), aabb, num_steps=T)
    buf = {"out": None}

    def frame():
        img, dep, ws = r.render(o, d, bg=1.0, chunk=chunk)
        if world == 1:
            return img, dep, ws
        out = torch.cat([img, dep.unsqueeze(-1), ws.unsqueeze(-1)], -1)
        out, buf["out"] = gather_cyclic(out, n, out=buf["out"])
        return out[:, :4], out[:, 4], out[:, 5]

    frame()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(frames):
        out = frame()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item()) / frames
    return {"metric": "640x480 multi-object composite Mrays/s", "value": n / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_frame": ms,
            "frames": frames, "rays": n, "objects": K, "samples_per_ray": T, "field_evaluations_per_s": n * K * T / (ms * 1e-3),
            "scaling": "strong", "coverage": float((out[2] > 0.5).float().mean().item()),
            "workload": "configs[3]: 4 resident object hash grids (16 levels, 2^19, F=2, 64-wide FFMLPs), 640x480 rays, 512 uniform samples "
                        "per ray and object, per-sample arg-max density + per-object colour masks + 4-channel composite, no early termination"}


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import cases
    from focnerf_b200 import _lib
    from focnerf_b200.trainer import Trainer

    world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        if os.environ.get("FOC_KEEP_NCCL_DEBUG") != "1":
            os.environ.pop("NCCL_DEBUG", None)      # NCCL_DEBUG=VERSION/WARN prints a banner on stdout; stdout carries ONE JSON line
        dist.init_process_group("nccl")
    dev = torch.device("cuda", local)
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    use_graph = (not args.no_graph) and (world == 1 or os.environ.get("FOC_GRAPH_DP", "1") == "1")
    # N > 1: data parallel.  Default: gradient exchange fused into the optimiser over NVLink peer memory (csrc/peer.cu);
    # FOC_DP=nccl selects the all-reduce + replicated sweep it replaces (A/B)
    dp = None if world == 1 else os.environ.get("FOC_DP", "peer")
    tr = Trainer(bound=1.0, density_grid=torch.from_numpy(grid), device=dev, seed=0, fused=not args.no_fused, use_graph=use_graph,
                 optimizer="torch" if args.torch_optim else "fused", pipeline=args.pipeline, data_parallel=dp)
    pool_dev, pool_host = make_pool(scene, POOL, RAYS_PER_GPU, 1000 + 7919 * rank, dev)
    stage = [torch.empty_like(t) for t in pool_dev[0]]

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def loop(n, host, first):
        for i in range(n):
            j = (first + i) % POOL
            if host:
                tr.train_step_host(*pool_host[j], stage)
            else:
                tr.train_step(*pool_dev[j])
            if tr.local_step % 16 == 0:
                tr.update_mean_count()              # renderer.py:533-536, every update_extra_interval = 16 steps

    loop(16, False, 0)                              # first "epoch": sizes the sample buffers (mean_count), untimed
    loop(max(3, args.warmup), False, 16)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    sync_all()
    prof = os.environ.get("FOC_PROFILE") == "1"     # ncu --profile-from-start off: capture the timed region only
    if prof:
        torch.cuda.profiler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loop(args.steps, False, 100)
    tr.flush()                                      # the pipelined step defers each optimiser sweep into the next step: apply the last one
    e1.record()
    sync_all()
    if prof:
        torch.cuda.profiler.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    clk = clocks.stop() if rank == 0 else None
    samples = float(tr.step_counter[:, 0].float().mean().item())       # mean samples/step over the last 16 steps

    # end to end: pinned host buffers -> H2D -> step -> D2H loss, through the public API
    loop(3, True, 200)
    sync_all()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    loop(args.steps, True, 300)
    tr.flush()
    t1.record()
    sync_all()
    ms_e2e = torch.tensor([t0.elapsed_time(t1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_e2e, op=dist.ReduceOp.MAX)
    ms_e2e = float(ms_e2e.item())

    # per-entry-point device times: the SAME step run eagerly (a graph replay cannot carry per-kernel events), with CUDA
    # events around every C-ABI call on the launching stream.  Not part of `value`; feeds roofline + gpu_launches.
    tr.use_graph = False
    n_prof = min(args.steps, 32)
    loop(3, False, 400)
    with _lib.KernelTimer(TIMED) as kt:
        sync_all()
        loop(n_prof, False, 500)
        sync_all()
    tr.use_graph = use_graph

    render = render_bench(tr.render, dev, rank, world, frames=args.frames) if args.frames > 0 else None
    multi = multi_bench(dev, rank, world, frames=args.multi_frames) if args.multi_frames > 0 else None

    # density-grid refresh (update_extra_state, every 16 steps in the reference's loop): timed on a copy of the state, after
    # everything else, because it replaces the analytic bitfield the timed steps marched through
    def time_refresh(full):
        tr.iter_density = 0 if full else 16
        tr.update_extra_state()
        torch.cuda.synchronize()
        t = time.perf_counter()
        for _ in range(3):
            tr.iter_density = 0 if full else 16
            tr.update_extra_state()
        torch.cuda.synchronize()
        return (time.perf_counter() - t) / 3 * 1e3
    refresh = {"full_ms": time_refresh(True), "partial_ms": time_refresh(False),
               "note": "one update_extra_state call (128^3 cells x cascade full; 2 x 128^3/4 cells partial), host wall clock; not part of `value`"}

    quality = None
    if args.psnr_steps > 0 and world == 1:
        quality = psnr_run(lambda: Trainer(bound=1.0, density_grid=torch.from_numpy(grid), device=dev, seed=0, fused=not args.no_fused,
                                           use_graph=use_graph, optimizer="torch" if args.torch_optim else "fused"),
                           scene, dev, args.psnr_steps, pool_dev)

    def shutdown():
        """Captured graphs hold NCCL work: drop them before the communicator goes, and never let teardown hang the job."""
        if world == 1:
            return
        sys.stdout.flush()
        threading.Timer(30.0, lambda: os._exit(0)).start()
        tr._graph = None
        import gc
        gc.collect()
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
        os._exit(0)

    if rank != 0:
        shutdown()
        return
    total_rays = world * RAYS_PER_GPU * args.steps
    hbm, how = peaks()
    # dominant kernel among OUR entry points, by mean device time per call inside the timed region
    per = {k: kt.mean_ms(k) * len(kt.events[k]) / n_prof for k in TIMED if kt.events.get(k)}
    top = max(per, key=per.get)
    if top in ALG_BYTES:
        achieved = ALG_BYTES[top] * samples / (kt.mean_ms(top) * 1e-3) / 1e9
        roof = {"kernel": top, "bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm, "traffic": None,
                "peak_source": how, "ms_per_launch": kt.mean_ms(top)}
    else:                                            # fused MLPs: the only dense contraction -> tensor pipe
        calls = len(kt.events[top]) / n_prof         # sigma + colour net per step
        flops = (36864 if top == "foc_ffmlp_forward" else 2 * 36864) * samples / calls
        try:
            tc = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
        except Exception:
            tc, how = TC_FALLBACK_TFLOPS, "fallback"
        achieved = flops / (kt.mean_ms(top) * 1e-3) / 1e12
        roof = {"kernel": top, "bound": "tensor", "achieved": achieved, "peak": tc, "unit": "TFLOP/s", "frac": achieved / tc, "traffic": None,
                "peak_source": how, "ms_per_launch": kt.mean_ms(top)}
    roof["traffic"] = NCU_TRAFFIC.get(top)
    roof["traffic_source"] = "profiles/r1_kernels.md (ncu --set full, DRAM bytes per call)"
    roof["ms_per_step_by_entry_point"] = {k: round(v, 4) for k, v in sorted(per.items(), key=lambda kv: -kv[1])}
    out = {
        "metric": "train rays/s", "value": total_rays / (ms * 1e-3), "unit": "rays/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {"workload": "configs[1]: single-object hash-grid NeRF train step (16 levels, 2^19 table, F=2, 64-wide FFMLPs, "
                               "4096 rays/batch per GPU, cuda_ray, fp16, Adam)", "rays_per_gpu": RAYS_PER_GPU,
                   "samples_per_step": samples, "parallelism": f"dp{world}" if world > 1 else "single",
                   "cuda_graph": use_graph, "fused_glue": not args.no_fused, "optimizer": "torch" if args.torch_optim else "fused ScaledAdam",
                   "l2": "fresh ray batch every step; step working set (fp32 table + Adam moments + fp16 copy + grads ~ 260 MB) exceeds the 126 MB L2"},
        "clocks": clk,
        "e2e": {"value": total_rays / (ms_e2e * 1e-3), "unit": "rays/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": world * RAYS_PER_GPU * 9 * 4, "d2h_bytes_per_step": world * 4},
        "gpu_launches": int(round(kt.kernel_launches() / n_prof * args.steps)),
        "render": render, "multi_object": multi, "density_refresh": refresh, "quality": quality,
        "roofline": roof,
    }
    if world == 1:
        out["cpu_baseline"] = cpu_baseline(scene, bits, cascade)
        # configs[0] (north_star: "the reference's PyTorch-only renderer path is timed on the box's own host cores in the same
        # run, with core count stated, as a reported baseline"): NeRFRenderer.run restated on torch CPU ops, 4096 rays x 128 samples
        from oracle import torch_renderer
        ro, rd = cases.camera_rays(RAYS_PER_GPU, 5)
        out["cpu_torch_renderer"] = torch_renderer.time_cpu(ro, rd, num_steps=128, runs=5, warmup=2)
    print(json.dumps(out))
    shutdown()


def run_reference(args):
    """Reference arm.  The reference has no CPU implementation of this path; its own CUDA extensions (oracle/_ref) are
    timed by bench_ref.py when they load, else the oracle port on the host cores."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    try:
        import bench_ref
        print(json.dumps(bench_ref.run(args)))
        return
    except Exception as e:
        note = f"reference extensions unavailable ({type(e).__name__}: {e}); CPU port timed instead"
    import cases
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    cb = cpu_baseline(scene, bits, cascade)
    print(json.dumps({"impl": "reference", "metric": "train rays/s", "value": cb["value"], "unit": "rays/s", "n_gpus": int(os.environ.get("WORLD_SIZE", 1)),
                      "steps": 1, "warmup": 1, "ms_per_step": 1e3 * 1024 / cb["value"], "higher_is_better": True, "scaling": "weak",
                      "vs_baseline": None, "dtype": "f16", "data": "synthetic", "config": {"workload": "configs[1] (bounded sample)", "note": note},
                      "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1024)
    ap.add_argument("--psnr-steps", type=int, default=512, help="steps of the fixed-iteration PSNR run (0 = skip)")
    ap.add_argument("--frames", type=int, default=3, help="800x800 frames of the render workload (0 = skip)")
    ap.add_argument("--multi-frames", type=int, default=2, help="640x480 frames of the 4-object composite workload (0 = skip)")
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="launch the step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--pipeline", action="store_true", help="defer each optimiser sweep into the next step's graph, beside the march")
    ap.add_argument("--torch-optim", action="store_true", help="torch.optim.Adam + GradScaler instead of the fused ScaledAdam")
    ap.add_argument("--no-fused", action="store_true", help="field network through the unfused operator modules")
    a = ap.parse_args()
    (run_reference if a.impl == "reference" else run_ours)(a)
