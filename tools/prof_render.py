"""One 800x800 frame between cudaProfilerStart/Stop, for `ncu --profile-from-start off` captures of the frame-sized kernels.
  python tools/prof_render.py [mode]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import bench_common as C  # noqa: E402
import synthetic  # noqa: E402
from focnerf_b200.trainer import Trainer  # noqa: E402

dev = torch.device("cuda", 0)
scene, cascade, grid, bits, thresh = C.scene_and_grid()
# round-2 protocol: the learnt-grid model after the fixed-iteration run
tr = Trainer(bound=C.BOUND, device=dev, seed=0, use_graph=True, lr=C.LR, iters=C.ITERS, update_extra_interval=C.UPDATE_EXTRA_INTERVAL)
pool_dev, _ = C.make_pool(scene, C.POOL, C.RAYS_PER_GPU, 1000, dev, pinned=False)
for i in range(C.PRETRAIN):
    tr.fit_step(*pool_dev[i % C.POOL])
pose = synthetic.ring_poses(8, radius=2.2, seed=123)[3]
o, d = synthetic.get_rays(pose, synthetic.intrinsics(800, 800), 800, 800)
o, d = o.to(dev), d.to(dev)
mode = sys.argv[1] if len(sys.argv) > 1 else "single_pass"
tr.render(o, d, mode=mode)
tr.render(o, d, mode=mode)
torch.cuda.synchronize()
torch.cuda.profiler.start()
tr.render(o, d, mode=mode)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
