"""MLP kernel timing vs batch size (fixed cost vs per-tile cost).  python tools/bench_mlp.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from focnerf_b200.ffmlp.ffmlp import _backend as fb
from focnerf_b200._lib import lib


def t_us(fn, reps=15, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        torch.cuda._sleep(300000)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    for L, name in ((2, "sigma"), (3, "colour")):
        nparam = 64 * (32 + 64 * (L - 1) + 16)
        w = (torch.randn(nparam, device="cuda") * 0.1).half()
        for tiles in (148, 592, 1409, 2818, 14090):
            B = tiles * 128
            x = torch.randn(B, 32, device="cuda").half()
            fbuf = torch.empty(L, B, 64, device="cuda", dtype=torch.half)
            out = torch.empty(B, 16, device="cuda", dtype=torch.half)
            g = torch.randn(B, 16, device="cuda").half()
            gi = torch.empty(B, 32, device="cuda", dtype=torch.half)
            gw = torch.empty(nparam, device="cuda", dtype=torch.half)
            f = t_us(lambda: fb.ffmlp_forward(x, w, B, 32, 16, 64, L, 0, 6, fbuf, out))
            fi = t_us(lambda: fb.ffmlp_inference(x, w, B, 32, 16, 64, L, 0, 6, None, out))
            b = t_us(lambda: fb.ffmlp_backward(g, x, w, fbuf, B, 32, 16, 64, L, 0, 6, True, None, gi, gw))
            print(f"{name:6s} tiles {tiles:6d}  fwd {f:8.1f} us  infer {fi:8.1f} us  bwd(+finish) {b:8.1f} us   per tile: fwd {f / tiles * 1e3:6.1f} ns  bwd {b / tiles * 1e3:6.1f} ns")


if __name__ == "__main__":
    main()
