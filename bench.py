"""bench.py -- train rays/s of the FOC-NeRF per-ray hot path (BASELINE.json configs[1]) + the render / multi-object legs.

Workload and protocol: bench_common.py (the reference's own recipe: 4096 rays/step, Adam lr 1e-3 with its per-step LambdaLR, fp16,
the density grid learnt from an empty start and refreshed every 16 steps INSIDE the timed region).  A "step" = refresh when due ->
march -> encode -> MLPs -> composite -> loss -> backward -> optimizer on one fresh batch of rays.

  python bench.py --gpus N --steps K --warmup W              (N>1 under torchrun, one rank per GPU; gradient exchange over NVLink peer memory)
  python bench.py --impl reference ...                       the reference's own unmodified stack on the same GPU (bench_ref.py)
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import bench_common as C  # noqa: E402

HBM_FALLBACK_GBS, TC_FALLBACK_TFLOPS = 6650.0, 1590.0        # B200_PROFILING.md fallbacks when MEASURED_PEAKS.json is absent

# algorithmic traffic per SAMPLE of each timed entry point (SURVEY.md 8d / DESIGN.md "algorithmic bytes")
ALG_BYTES = {
    "foc_grid_encode_forward": 12 + 512 + 64,          # coords + 16 levels x 8 corners x 4 B + fp16 [B,32] out
    "foc_grid_encode_backward": 12 + 64 + 512,         # coords + fp16 grad + 16 x 8 x 4 B atomic targets
    "foc_composite_rays_train_forward": 24,
    "foc_composite_rays_train_backward": 24 + 16,
    "foc_composite_mse_forward": 24,
    "foc_composite_mse_backward": 24 + 16,
    "foc_march_rays_train": 32,
    # the same kernels reading their row count from device memory (graph-replayed step)
    "foc_march_rays_train_rows": 32, "foc_grid_encode_forward_rows": 12 + 512 + 64, "foc_grid_encode_backward_rows": 12 + 64 + 512,
    "foc_composite_mse_backward_rows": 24 + 16,
    "foc_field_mid_forward": 108, "foc_field_mid_backward": 96, "foc_rgb_sigmoid_forward": 44, "foc_rgb_sigmoid_backward": 44,
}
ALG_BYTES_PER_PARAM = {"foc_adam_step": 30, "foc_adam_nonfinite": 2}
# data-parallel exchange (csrc/peer.cu): timed, no single-GPU roofline (NVLink pulls / pushes of 1/N of the table + barrier latency)
PEER_TIMED = ["foc_dp_reduce", "foc_adam_step_peers", "foc_adam_step_dp", "foc_peer_barrier"]
N_PARAMS = 6119864 * 2 + 7168 + 11264
# MLP entry points: FLOPs per sample and call (sigma net 32-64-64-16, colour net 32-64-64-64-16), forward; backward = 2x
MLP_FLOPS_FWD = 36864
# per sample and call: foc_ffmlp_* mean of the two nets; foc_field_forward (training) = enc 64 + dirs 12 + stored activations 5 x 128 + cin 64
# + sigma 4 + rgb 12; foc_ffmlp_backward_head = mean of colour (24 + 64 + 384 + 64) and sigma (8 + 32 + 64 + 256 + 64)
MLP_BYTES = {"foc_ffmlp_forward": (352 + 480) / 2, "foc_ffmlp_backward": (416 + 544) / 2, "foc_field_forward": 796, "foc_ffmlp_backward_head": (536 + 424) / 2}
MLP_FLOPS = {"foc_ffmlp_forward": 36864, "foc_ffmlp_backward": 2 * 36864, "foc_field_forward": 36864, "foc_ffmlp_backward_head": 2 * 36864}   # per sample, all calls of a step
TIMED = list(ALG_BYTES) + list(ALG_BYTES_PER_PARAM) + list(MLP_BYTES) + PEER_TIMED


def peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), float(d["bf16_tflops"]), float(d["bf16_tflops_sustained"]), "measured"
    except Exception:
        return HBM_FALLBACK_GBS, TC_FALLBACK_TFLOPS, TC_FALLBACK_TFLOPS, "fallback"


def ncu_traffic():
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per CALL of each entry point, extracted from the committed
    `ncu --set full` capture of this workload by tools/ncu_summary.py into profiles/r2_traffic.json; {} when absent."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
    except Exception:
        return {}


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
            "sample": f"1 train step (fwd+bwd+Adam) on {rays} rays of the same scene through the analytic occupancy grid, oracle C port with OpenMP, {dt:.1f} s"}


def render_bench(tr, dev, rank, world, frames=3, H=800, W=800, tag="own", mode=None):
    """configs[2]: 800x800 inference render, single pass with early termination.  N > 1: the frame's rays are dealt to the ranks
    in 128-ray blocks round-robin and every rank's finished rows are stored straight into every rank's frame buffer over NVLink
    peer memory by the render epilogue (peer.FrameExchange; FOC_RENDER_GATHER=nccl selects all_gather + un-deal copy instead)."""
    import torch
    import torch.distributed as dist
    from focnerf_b200.dist import cyclic_tile, gather_cyclic
    o, d = C.frame_rays(H, W)
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

    mode = mode or os.environ.get("FOC_RENDER_MODE", "single_pass")      # "windowed": depth windows (32, 64, 928), speculative sizing per window
    wins = (32, 64, 928)

    def frame():
        """-> (image [n,3], depth [n], weights_sum [n]) of the WHOLE frame on this rank"""
        if via_peer:
            return tr.render(o, d, exchange=ex, mode=mode, windows=wins)
        image, depth, ws = tr.render(o, d, mode=mode, windows=wins)
        if world == 1:
            return image, depth, ws
        out = torch.cat([image, depth.unsqueeze(-1), ws.unsqueeze(-1)], -1)
        out, buf["out"] = gather_cyclic(out, n, out=buf["out"])
        return out[:, :3], out[:, 3], out[:, 4]

    frame()
    frame()                                          # second warm-up: the sample buffers are sized from the previous frame
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if via_peer or world == 1:                       # a camera path: no host read between frames, one at the end (Trainer.render_frames)
        out = tr.render_frames([(o, d)] * frames, exchange=ex, mode=mode, windows=wins)[-1]
    else:
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
    res = {"metric": "800x800 render Mrays/s", "value": n / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_frame": ms, "frames": frames,
            "rays": n, "scaling": "strong", "mode": mode, "host_reads": "one 4-byte read per %d-frame path (Trainer.render_frames)" % frames if (via_peer or world == 1) else "one per frame",
            "frames_rendered_twice": tr.render_overflows, "coverage": C.coverage_of(out[0]), "mean_image": float(out[0].float().mean().item()), "checkpoint": tag,
            "exchange": "peer-memory stores in the render epilogue" if via_peer else ("all_gather" if world > 1 else None),
            "workload": "configs[2]: one 800x800 test pose of the synthetic scene, T_thresh 1e-4, max_steps 1024"}
    if ex is not None:                               # the frame buffers are peer-mapped: every rank is done with them before anyone unmaps
        del out
        torch.cuda.synchronize()
        dist.barrier()
        ex.close()
    return res


def multi_bench(dev, rank, world, frames=2, K=4, T=512, H=480, W=640, chunk=16384, offsets=None):
    """configs[3] (K = 4) / configs[4]-edit (K = 8, object 0 moved by editable.py's default offset): FOC multi-object composition,
    K resident object hash grids, 640x480 rays, T = 512 uniform samples per ray and object, arg-max-density selection + per-object
    colour masks + 4-channel composite.  N > 1: 128-ray blocks dealt round-robin, no data-path collective, one gather per frame."""
    import torch
    import torch.distributed as dist
    from focnerf_b200.multiobject import MultiObjectRenderer
    from focnerf_b200.network import NeRFNetwork
    from focnerf_b200.dist import cyclic_tile, gather_cyclic
    o, d = C.multi_rays(H, W)
    n = H * W
    if world > 1:
        idx = cyclic_tile(n, rank, world)
        o, d = o[idx], d[idx]
    o, d = o.contiguous().to(dev), d.contiguous().to(dev)
    aabb = torch.tensor([-1.0] * 3 + [1.0] * 3, device=dev)
    models = C.multi_object_tables([NeRFNetwork(bound=1.0).to(dev).eval() for _ in range(K)], lambda m: m.encoder.embeddings)
    r = MultiObjectRenderer(models, aabb, offsets=offsets, num_steps=T)
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
            "frames": frames, "rays": n, "objects": K, "samples_per_ray": T, "offsets": offsets is not None,
            "field_evaluations_per_s": n * K * T / (ms * 1e-3), "scaling": "strong",
            "mean_image": float(out[0].float().mean().item()), "mean_depth": float(out[1].float().mean().item()),
            "workload": f"{K} resident object hash grids (16 levels, 2^19, F=2, 64-wide FFMLPs), 640x480 rays, 512 uniform samples per ray and "
                        "object, per-sample arg-max density + per-object colour masks + 4-channel composite, no early termination"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    import cases
    from focnerf_b200 import _lib
    from focnerf_b200.trainer import Trainer

    world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    out_fd = None
    if world > 1:
        # stdout carries ONE JSON line, but NCCL writes its banner / INFO lines (NCCL_DEBUG set by a launcher) to fd 1 from C: point fd 1 at
        # stderr for the whole run -- NCCL's log stays visible there, e.g. for counting ranks -- and keep the real stdout for the JSON line
        sys.stdout.flush()
        out_fd = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl")
    dev = torch.device("cuda", local)
    scene, cascade, grid, bits, thresh = C.scene_and_grid()
    use_graph = (not args.no_graph) and (world == 1 or os.environ.get("FOC_GRAPH_DP", "1") == "1")
    # N > 1: data parallel.  Default: gradient exchange fused into the optimiser over NVLink peer memory (csrc/peer.cu);
    # FOC_DP=nccl selects the all-reduce + replicated sweep it replaces (A/B)
    dp = None if world == 1 else os.environ.get("FOC_DP", "peer")

    def make_trainer(seed, data_parallel=None, **kw):
        return Trainer(bound=C.BOUND, density_grid=None, density_thresh=C.DENSITY_THRESH, min_near=C.MIN_NEAR, dt_gamma=C.DT_GAMMA,
                       max_steps=C.MAX_STEPS, lr=C.LR, iters=C.ITERS, update_extra_interval=C.UPDATE_EXTRA_INTERVAL, device=dev, seed=seed,
                       fused=not args.no_fused, use_graph=use_graph, optimizer="torch" if args.torch_optim else "fused",
                       data_parallel=data_parallel, **kw)

    t_pool = time.time()
    pool_dev, pool_host = C.make_pool(scene, C.POOL, C.RAYS_PER_GPU, 1000 + 7919 * rank, dev)
    t_pool = time.time() - t_pool
    eo, ed, egt = C.heldout(scene, dev)
    stage = [torch.empty_like(t) for t in pool_dev[0]]

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def loop(tr, n, host, first, pool=None):
        pool = pool or (pool_host if host else pool_dev)
        for i in range(n):
            b = pool[(first + i) % len(pool)]
            if host:
                tr.fit_step_host(*b, stage)
            else:
                tr.fit_step(*b)

    clocks = C.ClockSampler(local)
    if rank == 0:
        clocks.start()
    # ---- fixed-iteration run = pretraining of the timed model (seed 0), the protocol's loop from an empty density grid ---------
    tr = make_trainer(C.QUALITY_SEEDS[0], data_parallel=dp)
    t0 = time.time()
    loop(tr, C.PRETRAIN, False, 0)
    torch.cuda.synchronize()
    pretrain_s = time.time() - t0
    psnrs = [C.psnr_of(tr.render(eo, ed)[0], egt)]            # N > 1: the data-parallel model (N x 4096 rays per step), reported as quality_dp

    # ---- timed region ---------------------------------------------------------------------------------------------------------------
    warm = max(3, args.warmup)
    if use_graph:
        tr.reset_graph()                              # steady state: the step is captured once more, sized from the current sample count
    loop(tr, warm, False, C.PRETRAIN)
    sync_all()
    prof = os.environ.get("FOC_PROFILE") == "1"     # ncu --profile-from-start off: capture the timed region only
    if prof:
        torch.cuda.profiler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    captures0 = tr.capture_count
    e0.record()
    loop(tr, args.steps, False, C.PRETRAIN + warm)
    tr.flush()
    e1.record()
    captures_timed = tr.capture_count - captures0
    sync_all()
    if prof:
        torch.cuda.profiler.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    samples = float(tr.step_counter[:, 0].float().mean().item())       # mean samples/step over the last <= 16 steps

    # end to end: pinned host buffers -> H2D -> step -> D2H loss, through the public API
    loop(tr, 3, True, 7)
    sync_all()
    t0e, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0e.record()
    loop(tr, args.steps, True, 10)
    tr.flush()
    t1e.record()
    sync_all()
    ms_e2e = torch.tensor([t0e.elapsed_time(t1e)], device=dev)
    if world > 1:
        dist.all_reduce(ms_e2e, op=dist.ReduceOp.MAX)
    ms_e2e = float(ms_e2e.item())
    clk = clocks.stop() if rank == 0 else None

    # data-parallel replicas must hold the same bits after the timed steps (the driver's GPU test box has one GPU: checked here)
    dp_parity = None
    if world > 1:
        def digest(t):
            return t.detach().contiguous().view(torch.uint8).to(torch.int64).sum()
        enc = tr.model.encoder
        table16 = enc._shadow16 if getattr(enc, "_shadow16", None) is not None else enc.embeddings
        d_ = torch.stack([digest(table16), digest(tr.model.sigma_net.weights), digest(tr.model.color_net.weights), digest(tr.density_bitfield)])
        all_d = [torch.empty_like(d_) for _ in range(world)]
        dist.all_gather(all_d, d_)
        dp_parity = bool(all(torch.equal(a, all_d[0]) for a in all_d))

    # configs[4] train leg: 32768 rays per GPU (262144 per step on 8 GPUs), same model, same loop
    big = None
    if args.big_steps > 0:
        per = args.big_rays // C.RAYS_PER_GPU
        big_pool = [tuple(torch.cat([pool_dev[(j * per + k) % C.POOL][t] for k in range(per)], 0) for t in range(3)) for j in range(C.POOL // per)]
        loop(tr, 32, False, 0, big_pool)             # two refreshes: mean_count follows the new batch size, the step is re-captured
        sync_all()
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b0.record()
        loop(tr, args.big_steps, False, 32, big_pool)
        tr.flush()
        b1.record()
        sync_all()
        bms = torch.tensor([b0.elapsed_time(b1)], device=dev)
        if world > 1:
            dist.all_reduce(bms, op=dist.ReduceOp.MAX)
        bms = float(bms.item())
        big = {"metric": "train rays/s", "value": world * args.big_rays * args.big_steps / (bms * 1e-3), "unit": "rays/s", "rays_per_gpu": args.big_rays,
               "global_rays_per_step": world * args.big_rays, "ms_per_step": bms / args.big_steps, "steps": args.big_steps,
               "samples_per_step_per_gpu": float(tr.step_counter[:, 0].float().mean().item()), "scaling": "weak",
               "workload": "configs[4] train leg: the same loop at 32768 rays per GPU (262144 rays per step on 8 GPUs)"}
        del big_pool
        loop(tr, 16, False, 0)                       # back to the 4096-ray batches (mean_count settles again)

    # per-entry-point device times.  With the CUDA graph: the step is re-captured with two EXTERNAL event-record nodes around every
    # C-ABI call (_lib.GraphKernelTimer), so the times are those of the kernels as they run inside the replayed step that `value` times
    # (an eager pass lets host launch gaps leak into the intervals: measured 61 us for a 13 us kernel).  Launch counts come from a short
    # eager pass (a replay makes no Python calls to count).  Neither pass is part of `value`.
    n_prof = min(args.steps, 32)
    tr.use_graph = False
    loop(tr, 2, False, 400)
    with _lib.KernelTimer(TIMED if not use_graph else ()) as kc:
        sync_all()
        loop(tr, n_prof if not use_graph else 4, False, 500)
        sync_all()
    launches_per_step = kc.kernel_launches() / (n_prof if not use_graph else 4)
    tr.use_graph = use_graph
    if use_graph:
        tr._graph = None                               # re-capture with the timer's wrappers in place
        with _lib.GraphKernelTimer(TIMED) as kt:
            for i in range(n_prof + 4):
                b_ = pool_dev[(600 + i) % C.POOL]
                kt.begin_step()
                tr.fit_step(*b_)
                torch.cuda.synchronize()
                kt.collect()
            tr._graph = None                           # the captured graph refers to the timer's events: drop it before they go
        timing_mode = "external event-record nodes inside the graph-replayed step"
    else:
        kt = kc
        kt.calls_per_step = lambda k: len(kt.events.get(k, ())) / n_prof
        kt.timed_names = lambda: [k for k in TIMED if kt.events.get(k)]
        timing_mode = "CUDA events around each call of an eager step"
    loop(tr, 2, False, 700)                            # clean re-capture for the legs below
    samples_prof = float(tr.step_counter[:min(16, max(1, tr.local_step)), 0].float().mean().item()) if tr.local_step > 0 else samples

    # ---- render / multi-object legs -------------------------------------------------------------------------------------------------
    render = render_own = render_w = None
    if args.frames > 0:
        render_own = render_bench(tr, dev, rank, world, frames=args.frames, tag="own (trained by this arm)")
        render = render_own
        if os.path.exists(C.SHARED_CKPT) and os.environ.get("FOC_NO_SHARED_CKPT") != "1":      # every N: the 1 -> 8 render curve is of ONE model
            # the reference arm (run first) left ITS trained model in the reference's checkpoint format: render the same model
            try:
                tr_s = make_trainer(0)
                missing, unexpected = tr_s.load_checkpoint(C.SHARED_CKPT)
                render = render_bench(tr_s, dev, rank, world, frames=args.frames, tag="shared: written by the reference arm's Trainer.save_checkpoint")
                render["checkpoint_keys"] = {"missing": missing, "unexpected": unexpected}
                render["psnr_db_heldout"] = C.psnr_of(tr_s.render(eo, ed)[0], egt)
                render_w = render_bench(tr_s, dev, rank, world, frames=args.frames, tag=render["checkpoint"], mode="windowed")
                del tr_s
            except Exception as e:
                render["shared_checkpoint_error"] = f"{type(e).__name__}: {e}"
    if args.frames > 0 and render_w is None:           # N > 1 (or no shared checkpoint): the depth-windowed mode on this arm's own model
        render_w = render_bench(tr, dev, rank, world, frames=args.frames, tag="own (trained by this arm)", mode="windowed")
    if render_w is not None:
        render_w["note"] = ("the same frames in depth windows (32, 64, 928 samples per ray) with the reference loop's exact termination rule and "
                            "speculative sizing per window: faster than `render` in absolute terms at every N, but it gains less on a 1/N share "
                            "(more launches per frame), so `render` (single pass) stays the headline whose 1 -> 8 scaling is quoted")
    multi = multi_bench(dev, rank, world, frames=args.multi_frames) if args.multi_frames > 0 else None
    multi8 = multi_bench(dev, rank, world, frames=1, K=8, offsets=[C.EDIT_OFFSET] + [None] * 7, chunk=8192) if args.multi_frames > 0 else None

    # one refresh on its own (it is already inside every timed region; this only says what it costs)
    def time_refresh():
        torch.cuda.synchronize()
        t = time.perf_counter()
        for _ in range(4):
            tr.update_extra_state()
        torch.cuda.synchronize()
        return (time.perf_counter() - t) / 4 * 1e3
    refresh = {"partial_ms": time_refresh(), "note": "one update_extra_state call in the steady state (2 x 128^3/4 cells), host wall clock; "
                                                     "already inside `value` every 16 steps"}

    # ---- quality: the other seeds of the fixed-iteration run ---------------------------------------------------------------------------
    quality = None
    quality_dp = None if world == 1 else {"psnr_db": psnrs[0], "steps": C.PRETRAIN, "rays_per_step": world * C.RAYS_PER_GPU, "eval_rays": C.EVAL_RAYS,
                                          "note": "rank 0's replica after the protocol's PRETRAIN data-parallel steps, held-out rays"}
    if world == 1:
        if args.psnr_steps > 0:
            for s in C.QUALITY_SEEDS[1:args.psnr_seeds]:
                tq = make_trainer(s)
                loop(tq, C.PRETRAIN, False, 0)
                psnrs.append(C.psnr_of(tq.render(eo, ed)[0], egt))
                del tq
        m, sd = C.mean_sd(psnrs)
        quality = {"psnr_db": m, "psnr_sd_db": sd, "psnr_per_seed": psnrs, "seeds": list(C.QUALITY_SEEDS[:len(psnrs)]), "steps": C.PRETRAIN,
                   "eval_rays": C.EVAL_RAYS, "note": "fresh model per seed, the protocol's first PRETRAIN steps from an empty density grid, held-out rays"}

    # end to end from PIXEL INDICES: a resident 8-view dataset of the same scene (nerf/provider.py --preload keeps
    # poses and images on the GPU), the host uploads (pixel, camera) int32 pairs and reads the loss back; rays and ground truth are built on the
    # device by csrc/rays.cu.  Same model, continuing from the steady state; timing only.
    e2e_pixels = None
    if world == 1 and args.steps > 0 and not args.no_pixels:
        import numpy as np
        import synthetic
        from focnerf_b200 import rays as R
        Hp = Wp = 800
        poses = torch.from_numpy(np.asarray(synthetic.ring_poses(8, radius=2.2, seed=777), np.float32)).to(dev)
        intr = synthetic.intrinsics(Hp, Wp)
        imgs = torch.stack([scene.render_gt(*R.pixel_batch(poses[k:k + 1], intr, Hp, Wp)[:2]) for k in range(8)])
        ds = R.PixelDataset(poses, intr, Hp, Wp, imgs, bg=1.0)
        g = np.random.RandomState(99)
        idx_pool = [(torch.from_numpy(g.randint(0, Hp * Wp, C.RAYS_PER_GPU).astype(np.int32)).pin_memory(),
                     torch.from_numpy(g.randint(0, 8, C.RAYS_PER_GPU).astype(np.int32)).pin_memory()) for _ in range(64)]
        st_idx = (torch.empty(C.RAYS_PER_GPU, dtype=torch.int32, device=dev), torch.empty(C.RAYS_PER_GPU, dtype=torch.int32, device=dev))
        for i in range(48):
            tr.fit_step_pixels(ds, *idx_pool[i % 64], st_idx)
        torch.cuda.synchronize()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        for i in range(args.steps):
            tr.fit_step_pixels(ds, *idx_pool[(48 + i) % 64], st_idx)
        p1.record()
        torch.cuda.synchronize()
        pms = p0.elapsed_time(p1) / args.steps
        e2e_pixels = {"value": C.RAYS_PER_GPU / (pms * 1e-3), "unit": "rays/s", "ms_per_step": pms, "h2d_bytes_per_step": C.RAYS_PER_GPU * 8, "d2h_bytes_per_step": 4,
                      "samples_per_step": float(tr.step_counter[:, 0].float().mean().item()),
                      "note": "Trainer.fit_step_pixels: the host uploads 4096 (pixel, camera) int32 pairs per step, foc_pixel_batch builds rays + ground truth from a "
                              "resident 8-view 800x800 dataset of the same scene; density refresh every 16 steps inside, loss read back every step"}
        del ds, imgs

    # round-1 workload for continuity (frozen analytic occupancy grid, lr 1e-2, no refresh): NOT the headline
    frozen = None
    if world == 1 and args.frozen_steps > 0:
        tf = Trainer(bound=C.BOUND, density_grid=torch.from_numpy(grid), device=dev, seed=0, lr=1e-2, use_graph=use_graph,
                     optimizer="torch" if args.torch_optim else "fused", fused=not args.no_fused)
        for i in range(32 + 8):
            tf.train_step(*pool_dev[i % C.POOL])
            if tf.local_step % 16 == 0:
                tf.update_mean_count()
        torch.cuda.synchronize()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for i in range(args.frozen_steps):
            tf.train_step(*pool_dev[(40 + i) % C.POOL])
            if tf.local_step % 16 == 0:
                tf.update_mean_count()
        f1.record()
        torch.cuda.synchronize()
        fms = f0.elapsed_time(f1) / args.frozen_steps
        frozen = {"value": C.RAYS_PER_GPU / (fms * 1e-3), "unit": "rays/s", "ms_per_step": fms, "samples_per_step": float(tf.step_counter[:, 0].float().mean().item()),
                  "note": "round 1's workload, for continuity only: marches through the analytic scene's frozen occupancy grid, no refresh"}
        del tf

    def shutdown():
        """Captured graphs hold peer-memory work: drop them before the communicator goes, and never let teardown hang the job."""
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
    total_rays = world * C.RAYS_PER_GPU * args.steps
    hbm, tc_burst, tc_sus, how = peaks()
    traffic = ncu_traffic()
    calls_of = {}
    # every timed entry point against its roofline: mean device time per call inside the eager pass, algorithmic work of one call
    per, rows = {}, {}
    for k in kt.timed_names():
        calls = kt.calls_per_step(k)
        t_call = kt.mean_ms(k) * 1e-3
        per[k] = kt.mean_ms(k) * calls
        if k in PEER_TIMED:
            rows[k] = {"ms_per_call": round(kt.mean_ms(k), 4), "calls_per_step": round(calls, 4)}
            continue
        if k in ALG_BYTES:
            nbytes = ALG_BYTES[k] * samples_prof
        elif k in ALG_BYTES_PER_PARAM:
            nbytes = ALG_BYTES_PER_PARAM[k] * N_PARAMS
        else:
            nbytes = MLP_BYTES[k] * samples_prof
        row = {"ms_per_call": kt.mean_ms(k), "calls_per_step": calls, "hbm_GBs": nbytes / t_call / 1e9, "hbm_frac": nbytes / t_call / 1e9 / hbm}
        if k in MLP_BYTES:
            flops = MLP_FLOPS[k] * samples_prof / calls
            row.update({"tensor_TFLOPs": flops / t_call / 1e12, "tensor_frac_of_burst": flops / t_call / 1e12 / tc_burst})
        tk = traffic.get("entry_points", {}).get(k[:-5] if k.endswith("_rows") else k)
        if tk:                                            # measured by ncu on one step of this workload, scaled to this run's samples per step
            sc_ = 1.0 if k in ALG_BYTES_PER_PARAM else samples_prof / traffic["samples_per_step"]
            row.update({"algorithmic_MB": nbytes / 1e6, "ncu_dram_MB": tk["dram_bytes"] * sc_ / 1e6, "ncu_l2_MB": tk["l2_bytes"] * sc_ / 1e6,
                        "l2_GBs": tk["l2_bytes"] * sc_ / t_call / 1e9, "ncu_tensor_pipe_pct": tk["tensor_pipe_pct"]})
        rows[k] = {kk: round(vv, 4) for kk, vv in row.items()}
        calls_of[k] = max(calls, 1.0)
    top = max((k for k in per if k not in PEER_TIMED), key=per.get)
    r_ = rows[top]
    if top in MLP_BYTES:
        # arithmetic intensity of the MLP entry points: MLP_FLOPS / MLP_BYTES = 46 (field forward) .. 77 (one backward call) FLOP per byte, below the
        # ridge of this part (burst tensor peak / HBM peak = 252 FLOP/B): the roofline that bounds them is HBM; the tensor-pipe figures
        # (FLOPs / time against the burst peak, and ncu's pipe-active counter) ride along as north_star asks
        roof = {"kernel": top, "bound": "hbm", "achieved": r_["hbm_GBs"], "peak": hbm, "unit": "GB/s", "frac": r_["hbm_frac"],
                "tensor_TFLOPs": r_["tensor_TFLOPs"], "tensor_frac_of_burst": r_["tensor_frac_of_burst"], "tensor_peak_TFLOPs": tc_burst,
                "ncu_tensor_pipe_active_pct": r_.get("ncu_tensor_pipe_pct"),
                "note": "stored activations make the fused MLP kernels HBM-bound at training sizes (intensity %.0f FLOP/B < ridge %.0f)"
                        % (MLP_FLOPS[top] / calls_of[top] / MLP_BYTES[top], tc_burst * 1e12 / (hbm * 1e9))}
    else:
        roof = {"kernel": top, "bound": "hbm", "achieved": r_["hbm_GBs"], "peak": hbm, "unit": "GB/s", "frac": r_["hbm_frac"]}
    t_top = traffic.get("entry_points", {}).get(top[:-5] if top.endswith("_rows") else top)
    roof.update({"traffic": None if not t_top else round(t_top["dram_bytes"] * (1.0 if top in ALG_BYTES_PER_PARAM else samples_prof / traffic["samples_per_step"])),
                 "traffic_source": "profiles/r2_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum per call from the committed ncu --set full capture of this "
                                   "workload (profiles/r2_kernels.md), scaled by samples per step" if t_top else None,
                 "peak_source": how, "ms_per_launch": r_["ms_per_call"], "samples_per_step": samples_prof, "timing": timing_mode,
                 "ms_per_step_by_entry_point": {k: round(v, 4) for k, v in sorted(per.items(), key=lambda kv: -kv[1])}, "by_entry_point": rows})
    out = {
        "metric": "train rays/s", "value": total_rays / (ms * 1e-3), "unit": "rays/s", "n_gpus": world, "steps": args.steps,
        "warmup": warm, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "render": render, "train_262k": big, "dp_parity": dp_parity,
        "config": {"workload": C.WORKLOAD, "rays_per_gpu": C.RAYS_PER_GPU, "samples_per_step": samples, "pretrain_steps": C.PRETRAIN,
                   "pretrain_s": pretrain_s, "pool_batches": C.POOL, "pool_build_s": t_pool, "parallelism": f"dp{world}" if world > 1 else "single",
                   "cuda_graph": use_graph, "graph_captures_in_timed_region": captures_timed, "fused_glue": not args.no_fused, "optimizer": "torch" if args.torch_optim else "fused ScaledAdam",
                   "l2": "fresh ray batch every step; step working set (fp32 table + Adam moments + fp16 copy + grads ~ 260 MB) exceeds the 126 MB L2"},
        "clocks": clk,
        "e2e": {"value": total_rays / (ms_e2e * 1e-3), "unit": "rays/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": world * C.RAYS_PER_GPU * 9 * 4, "d2h_bytes_per_step": world * 4},
        "gpu_launches": int(round(launches_per_step * args.steps)),
        "render_own_model": render_own if render is not render_own else None, "render_windowed": render_w,
        "multi_object": multi, "multi_object_edit8": multi8, "density_refresh": refresh, "quality": quality, "quality_dp": quality_dp, "e2e_pixels": e2e_pixels, "frozen_bitfield": frozen,
        "roofline": roof,
    }
    if world == 1:
        out["cpu_baseline"] = cpu_baseline(scene, bits, cascade)
        # configs[0] (north_star: "the reference's PyTorch-only renderer path is timed on the box's own host cores in the same
        # run, with core count stated, as a reported baseline"): NeRFRenderer.run restated on torch CPU ops, 4096 rays x 128 samples
        from oracle import torch_renderer
        ro, rd = cases.camera_rays(C.RAYS_PER_GPU, 5)
        out["cpu_torch_renderer"] = torch_renderer.time_cpu(ro, rd, num_steps=128, runs=5, warmup=2)
    if out_fd is None:
        print(json.dumps(out))
    else:
        sys.stdout.flush()
        os.write(out_fd, (json.dumps(out) + "\n").encode())
    shutdown()


def run_reference(args):
    """Reference arm: the reference's own unmodified stack (bench_ref.py); rank 0 alone runs it."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    import bench_ref
    try:
        print(json.dumps(bench_ref.run(args)))
    except Exception as e:
        import traceback
        traceback.print_exc()
        print(json.dumps({"impl": "reference", "unavailable": f"{type(e).__name__}: {e}"[:300]}))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--psnr-steps", type=int, default=C.PRETRAIN, help="0 = skip the extra seeds of the fixed-iteration PSNR run")
    ap.add_argument("--psnr-seeds", type=int, default=len(C.QUALITY_SEEDS))
    ap.add_argument("--frames", type=int, default=3, help="800x800 frames of the render workload (0 = skip)")
    ap.add_argument("--multi-frames", type=int, default=2, help="640x480 frames of the multi-object composite workloads (0 = skip)")
    ap.add_argument("--big-rays", type=int, default=32768, help="rays per GPU of the configs[4] train leg")
    ap.add_argument("--big-steps", type=int, default=32, help="timed steps of the configs[4] train leg (0 = skip)")
    ap.add_argument("--frozen-steps", type=int, default=64, help="timed steps of round 1's frozen-occupancy workload, for continuity (0 = skip)")
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-pixels", action="store_true", help="skip the pixel-index end-to-end leg")
    ap.add_argument("--no-graph", action="store_true", help="launch the step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--torch-optim", action="store_true", help="torch.optim.Adam + GradScaler instead of the fused ScaledAdam")
    ap.add_argument("--no-fused", action="store_true", help="field network through the unfused operator modules")
    a = ap.parse_args()
    (run_reference if a.impl == "reference" else run_ours)(a)
