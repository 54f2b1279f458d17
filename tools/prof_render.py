"""One 800x800 frame between cudaProfilerStart/Stop, for `ncu --profile-from-start off` captures of the frame-sized kernels.
  python tools/prof_render.py [mode]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import cases  # noqa: E402
import bench  # noqa: E402
import synthetic  # noqa: E402
from focnerf_b200.trainer import Trainer  # noqa: E402

dev = torch.device("cuda", 0)
scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
tr = Trainer(bound=1.0, density_grid=torch.from_numpy(grid), device=dev, seed=0, use_graph=True)
pool_dev, _ = bench.make_pool(scene, 8, bench.RAYS_PER_GPU, 1000, dev)
for i in range(64):
    tr.train_step(*pool_dev[i % 8])
    if tr.local_step % 16 == 0:
        tr.update_mean_count()
pose = synthetic.ring_poses(8, radius=2.2, seed=123)[3]
o, d = synthetic.get_rays(pose, synthetic.intrinsics(800, 800), 800, 800)
o, d = o.to(dev), d.to(dev)
mode = sys.argv[1] if len(sys.argv) > 1 else "single_pass"
tr.render(o, d, mode=mode)
torch.cuda.synchronize()
torch.cuda.profiler.start()
tr.render(o, d, mode=mode)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
