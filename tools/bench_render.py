"""Time Trainer.render modes on one GPU: single_pass vs depth-windowed (several window schedules) vs the reference-style loop.
  python tools/bench_render.py [train_steps]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import cases  # noqa: E402
import synthetic  # noqa: E402
from focnerf_b200.trainer import Trainer  # noqa: E402


def main(steps):
    dev = torch.device("cuda", 0)
    # round-2 protocol (bench_common): the learnt-grid model after `steps` iterations of the reference's loop
    import bench_common as C
    scene, cascade, grid, bits, thresh = C.scene_and_grid()
    tr = Trainer(bound=C.BOUND, device=dev, seed=0, use_graph=True, lr=C.LR, iters=C.ITERS, update_extra_interval=C.UPDATE_EXTRA_INTERVAL)
    pool_dev, _ = C.make_pool(scene, C.POOL, C.RAYS_PER_GPU, 1000, dev, pinned=False)
    for i in range(steps):
        tr.fit_step(*pool_dev[i % C.POOL])
    pose = synthetic.ring_poses(8, radius=2.2, seed=123)[3]
    o, d = synthetic.get_rays(pose, synthetic.intrinsics(800, 800), 800, 800)
    o, d = o.to(dev), d.to(dev)
    if os.environ.get("FOC_RENDER_SHARE"):                      # one rank's share of the frame at world = N, on this GPU alone
        from focnerf_b200.dist import cyclic_tile
        from focnerf_b200 import peer
        n_world = int(os.environ["FOC_RENDER_SHARE"])
        idx = cyclic_tile(800 * 800, 0, n_world).to(dev)
        oo, dd = o[idx].contiguous(), d[idx].contiguous()
        for mode in ("single_pass", "windowed"):
            for _ in range(3):
                tr.render(oo, dd, mode=mode, windows=(32, 64, 928))
            torch.cuda.synchronize()
            t = time.perf_counter()
            tr.render_frames([(oo, dd)] * 10, mode=mode, windows=(32, 64, 928))
            torch.cuda.synchronize()
            print(f"1/{n_world} share ({oo.shape[0]} rays), render_frames x 10, {mode}: {(time.perf_counter() - t) / 10 * 1e3:.3f} ms/frame "
                  f"(overflows {tr.render_overflows})", flush=True)
    for mode in ("single_pass", "windowed"):
        for _ in range(3):
            tr.render(o, d, mode=mode, windows=(32, 64, 928))
        torch.cuda.synchronize()
        t = time.perf_counter()
        tr.render_frames([(o, d)] * 5, mode=mode, windows=(32, 64, 928))
        torch.cuda.synchronize()
        print(f"whole frame, render_frames x 5, {mode}: {(time.perf_counter() - t) / 5 * 1e3:.3f} ms/frame (overflows {tr.render_overflows})", flush=True)
    if os.environ.get("FOC_RENDER_QUICK"):
        return
    ref = None
    for name, kw in [("single_pass", dict(mode="single_pass")),
                     ("windowed 48+976", dict(mode="windowed", windows=(48, 976))),
                     ("windowed 32+992", dict(mode="windowed", windows=(32, 992))),
                     ("windowed 64+960", dict(mode="windowed", windows=(64, 960))),
                     ("windowed 24+48+952", dict(mode="windowed", windows=(24, 48, 952))),
                     ("windowed 32+64+928", dict(mode="windowed", windows=(32, 64, 928))),
                     ("windowed 16+32+64+912", dict(mode="windowed", windows=(16, 32, 64, 912))),
                     ("loop", dict(mode="loop"))]:
        out = tr.render(o, d, **kw)
        torch.cuda.synchronize()
        t = time.perf_counter()
        for _ in range(3):
            out = tr.render(o, d, **kw)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t) / 3 * 1e3
        img = torch.nan_to_num(out[0]).clone()
        if ref is None:
            ref = img
        if os.environ.get("FOC_RENDER_BREAKDOWN") == "1" and name in ("single_pass", "windowed 48+976"):
            breakdown(lambda: tr.render(o, d, **kw), name)
        print(f"{name:28s} {ms:8.2f} ms/frame  {0.64 / ms * 1e3:7.1f} Mrays/s   max|img - single_pass| = {float((img - ref).abs().max()):.2e}", flush=True)


def breakdown(fn, title):
    from torch.profiler import ProfilerActivity, profile
    fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        fn()
        torch.cuda.synchronize()
    print(f"==== {title}")
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=70))


def multi_breakdown():
    dev = torch.device("cuda", 0)
    from focnerf_b200.multiobject import MultiObjectRenderer
    from focnerf_b200.network import NeRFNetwork
    import bench_common as C
    o, d = (t.to(dev) for t in C.multi_rays())
    aabb = torch.tensor([-1.0] * 3 + [1.0] * 3, device=dev)
    models = C.multi_object_tables([NeRFNetwork(bound=1.0).to(dev).eval() for _ in range(4)], lambda m: m.encoder.embeddings)
    r = MultiObjectRenderer(models, aabb, num_steps=512)
    breakdown(lambda: r.render(o[:65536], d[:65536], chunk=16384), "multi-object, 65536 rays x 4 objects x 512 samples")


if __name__ == "__main__":
    if os.environ.get("FOC_RENDER_BREAKDOWN") == "1":
        multi_breakdown()
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 1024)
