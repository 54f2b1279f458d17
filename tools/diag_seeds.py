"""Diagnostic: seed-to-seed spread of the fixed-iteration PSNR under candidate protocols (scratch grid, refresh every 16)."""
import json, math, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import numpy as np
import torch
import cases
from focnerf_b200.trainer import Trainer

dev = torch.device("cuda", 0)
scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)


def pool(n, seed):
    out = []
    for i in range(n):
        o, d = cases.camera_rays(4096, seed + i)
        o, d = torch.from_numpy(o).to(dev), torch.from_numpy(d).to(dev)
        out.append((o, d, scene.render_gt(o, d)))
    return out


def psnr(tr, o, d, gt):
    img = tr.render(o, d)[0]
    return -10 * math.log10(max(float(torch.mean((img.float() - gt) ** 2)), 1e-12))


P = pool(256, 1000)
eo, ed = cases.camera_rays(65536, 424242)
eo, ed = torch.from_numpy(eo).to(dev), torch.from_numpy(ed).to(dev)
egt = scene.render_gt(eo, ed)
marks = [128, 256, 512, 1024, 2048]


def run(seed, steps, lr, decay, npool):
    tr = Trainer(bound=1.0, density_grid=None, device=dev, seed=seed, use_graph=True, lr=lr)
    torch.manual_seed(1234 + seed)
    out = {}
    for i in range(steps):
        if i % 16 == 0:
            tr.update_extra_state()
        if decay:
            tr.optimizer.set_lr(lr * 0.1 ** min(i / steps, 1.0))
        tr.train_step(*P[i % npool])
        if (i + 1) in marks:
            out[i + 1] = psnr(tr, eo, ed, egt)
    return out


res = {}
for name, steps, lr, decay, npool in [("lr1e-3_decay_p32", 2048, 1e-3, True, 32), ("lr1e-3_decay_p256", 2048, 1e-3, True, 256),
                                      ("lr1e-2_decay_p32", 2048, 1e-2, True, 32), ("lr1e-2_decay_p256", 2048, 1e-2, True, 256),
                                      ("lr1e-2_decay_p256_1024", 1024, 1e-2, True, 256), ("lr1e-3_const_p256", 2048, 1e-3, False, 256)]:
    t0 = time.time()
    runs = [run(s, steps, lr, decay, npool) for s in range(5)]
    res[name] = {m: (float(np.mean([r[m] for r in runs])), float(np.std([r[m] for r in runs], ddof=1))) for m in marks if m <= steps}
    print(name, {m: f"{a:.2f}+-{b:.2f}" for m, (a, b) in res[name].items()}, f"{time.time() - t0:.1f}s", flush=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "diag_seeds.json"), "w"), indent=1)
