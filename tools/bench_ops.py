"""Per-operator timing on realistic inputs (the march output of the synthetic scene), for kernel experiments.
  python tools/bench_ops.py [--rays 4096]
Prints mean device time (CUDA events, 20 reps after 5 warm-ups, L2 flushed between reps) per operator / variant."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch

import cases
from focnerf_b200 import raymarching
from focnerf_b200.gridencoder.grid import GridEncoder
from focnerf_b200.trainer import Trainer


def timeit(fn, reps=20, warm=5, flush=None):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(600000)                # keep the GPU busy while the host enqueues, so host latency is not timed
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=4096)
    a = ap.parse_args()
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    tr = Trainer(bound=1.0, density_grid=torch.from_numpy(grid))
    o, d = cases.camera_rays(a.rays, 5)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    nears, fars = raymarching.near_far_from_aabb(o, d, tr.aabb, tr.min_near)
    flush = torch.empty(192 << 20, dtype=torch.uint8, device="cuda")
    counter = torch.zeros(2, dtype=torch.int32, device="cuda")

    def march():
        counter.zero_()
        return raymarching.march_rays_train(o, d, 1.0, tr.density_bitfield, tr.cascade, 128, nears, fars, counter, 180224, True, 128, False, 0.0, 1024)
    xyzs, dirs, deltas, rays = march()
    M = xyzs.shape[0]
    print(f"samples {int(counter[0])} buffer {M}")
    print(f"march_rays_train            {timeit(march, flush=flush):8.1f} us")
    enc = tr.model.encoder
    with torch.autocast("cuda", dtype=torch.float16):
        print(f"grid fwd                    {timeit(lambda: enc(xyzs, bound=1.0), flush=flush):8.1f} us")
        out = enc(xyzs, bound=1.0)
        g = torch.randn_like(out)
        g[int(counter[0]):] = 0                   # padding rows of the sample buffer carry zero gradient in training
    from focnerf_b200.gridencoder.grid import _backend as gb, FOC_BLC
    import numpy as np
    emb16 = enc.embeddings.detach().half()
    gtab = torch.zeros_like(emb16)
    x01 = ((xyzs + 1.0) / 2.0).contiguous()
    B, L, C, D = xyzs.shape[0], enc.num_levels, enc.level_dim, 3
    Sc = np.log2(enc.per_level_scale)

    def kbwd():
        gb.grid_encode_backward(g, x01, emb16, enc.offsets, gtab, B, D, C, L, Sc, enc.base_resolution, None, None, enc.gridtype_id,
                                enc.align_corners, enc.interp_id, FOC_BLC)

    def run(label):
        ts = []
        for _ in range(12):
            gtab.zero_(); flush.zero_(); gtab.zero_()
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(600000)
            a_.record(); kbwd(); b_.record(); torch.cuda.synchronize()
            ts.append(a_.elapsed_time(b_) * 1e3)
        ts.sort()
        print(f"grid bwd kernel {label:40s} {ts[len(ts) // 2]:8.1f} us")
    run("runs + scatter")
    if True:
        from focnerf_b200 import _lib
        with _lib.KernelTimer(["foc_grid_encode_backward", "foc_grid_encode_forward", "foc_march_rays_train"]) as kt:
            for _ in range(10):
                flush.zero_(); torch.cuda._sleep(2000000); march()
                with torch.autocast("cuda", dtype=torch.float16):
                    out2 = enc(xyzs, bound=1.0)
                enc.embeddings.grad = None; out2.backward(g)
            torch.cuda.synchronize()
        for k in kt.events:
            print(f"{k:36s} {kt.mean_ms(k) * 1e3:8.1f} us per call")


if __name__ == "__main__":
    main()
