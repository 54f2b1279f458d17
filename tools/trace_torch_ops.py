"""List the torch (non-libfoc) kernels of one eager train step with the CPU op chain that launched each one.
  python tools/trace_torch_ops.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import cases  # noqa: E402
import bench  # noqa: E402
from focnerf_b200.trainer import Trainer  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    tr = Trainer(bound=1.0, density_grid=torch.from_numpy(grid), device=dev, seed=0, use_graph=False)
    pool_dev, _ = bench.make_pool(scene, 4, bench.RAYS_PER_GPU, 1000, dev)
    for i in range(20):
        tr.train_step(*pool_dev[i % 4])
        if tr.local_step % 16 == 0:
            tr.update_mean_count()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        tr.train_step(*pool_dev[1])
        torch.cuda.synchronize()
    n = 0
    for ev in prof.events():
        for k in ev.kernels:
            n += 1
            chain, p = [], ev
            while p is not None and len(chain) < 6:
                chain.append(p.name)
                p = p.cpu_parent
            print(f"{k.duration:7.1f} us  {k.name[:70]:70s} <- {' <- '.join(chain)}")
    print(f"{n} kernels in the step")


if __name__ == "__main__":
    main()
