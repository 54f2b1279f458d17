"""Diagnostic (not part of the product or the bench): why does the fixed-iteration run stall near 14.8 dB?
Trains the bench scene under several settings, logs loss / PSNR curves, dumps GT and predicted frames as PNG."""
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np
import torch

import cases
import synthetic
from focnerf_b200.trainer import Trainer

OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)
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
    mse = float(torch.mean((img.float() - gt) ** 2))
    return -10 * math.log10(max(mse, 1e-12))


def save_png(name, img, H, W):
    from PIL import Image
    a = (img.clamp(0, 1).reshape(H, W, 3).cpu().numpy() * 255).astype(np.uint8)
    Image.fromarray(a).save(os.path.join(OUT, name))


P = pool(32, 1000)
eo, ed = cases.camera_rays(32768, 424242)
eo, ed = torch.from_numpy(eo).to(dev), torch.from_numpy(ed).to(dev)
egt = scene.render_gt(eo, ed)
print("white-frame PSNR on held-out rays:", -10 * math.log10(float(torch.mean((1.0 - egt) ** 2))), flush=True)
print("fraction of held-out rays not white:", float(((1.0 - egt).abs().amax(-1) > 1e-3).float().mean()), flush=True)
H = W = 400
pose = synthetic.ring_poses(8, radius=2.2, seed=123)[3]
fo, fd = synthetic.get_rays(pose, synthetic.intrinsics(H, W), H, W)
fo, fd = fo.to(dev), fd.to(dev)
fgt = scene.render_gt(fo, fd)
save_png("diag_gt.png", fgt, H, W)


def run(name, steps, marks, refresh=False, pool_=P, scratch=False, decay_iters=0, seed=0, **kw):
    t0 = time.time()
    tr = Trainer(bound=1.0, density_grid=None if scratch else torch.from_numpy(grid), device=dev, seed=seed, **kw)
    lr0 = kw.get("lr", 1e-2)
    log = {}
    losses = []
    for i in range(steps):
        if scratch and i % 16 == 0:
            tr.update_extra_state()
        if decay_iters:
            tr.optimizer.set_lr(lr0 * 0.1 ** min(i / decay_iters, 1.0))
        loss = tr.train_step(*pool_[i % len(pool_)])
        if i % 32 == 0:
            losses.append(float(loss.detach()))
        if not scratch and tr.local_step % 16 == 0:
            if refresh:
                tr.update_extra_state()
            else:
                tr.update_mean_count()
        if (i + 1) in marks:
            sc = getattr(tr.optimizer, "scale", None)
            log[i + 1] = {"psnr_heldout": round(psnr(tr, eo, ed, egt), 3), "psnr_train0": round(psnr(tr, *pool_[0]), 3),
                          "loss": float(loss), "mean_count": tr.mean_count, "scale": float(sc) if sc is not None else None}
            print(name, i + 1, log[i + 1], flush=True)
    img = tr.render(fo, fd)[0]
    save_png(f"diag_{name}.png", img.float(), H, W)
    print(name, "losses/32:", [round(x, 5) for x in losses[:64]], f"{time.time() - t0:.1f}s", flush=True)
    return log


marks = {64, 128, 256, 512, 1024, 2048, 4096, 8192}
res = {}
which = sys.argv[1:] or ["base", "eager_torch", "refresh", "lr3", "bigpool"]
if "base" in which:
    res["base"] = run("base", 4096, marks, use_graph=True)
if "eager_torch" in which:
    res["eager_torch"] = run("eager_torch", 1024, marks, use_graph=False, optimizer="torch", fused=False)
if "refresh" in which:
    res["refresh"] = run("refresh", 4096, marks, refresh=True, use_graph=True)
if "lr3" in which:
    res["lr3"] = run("lr3", 2048, marks, use_graph=True, lr=2e-3)
if "bigpool" in which:
    res["bigpool"] = run("bigpool", 4096, marks, pool_=pool(256, 5000), use_graph=True)
if "frozen_lr1e-3" in which:
    res["frozen_lr1e-3"] = run("frozen_lr1e-3", 4096, marks, use_graph=True, lr=1e-3)
if "scratch_lr1e-3" in which:
    res["scratch_lr1e-3"] = run("scratch_lr1e-3", 8192, marks, scratch=True, use_graph=True, lr=1e-3)
if "scratch_lr1e-2" in which:
    res["scratch_lr1e-2"] = run("scratch_lr1e-2", 4096, marks, scratch=True, use_graph=True, lr=1e-2)
if "scratch_lr1e-2_decay" in which:
    res["scratch_lr1e-2_decay"] = run("scratch_lr1e-2_decay", 4096, marks, scratch=True, use_graph=True, lr=1e-2, decay_iters=4096)
if "scratch_lr3e-3" in which:
    res["scratch_lr3e-3"] = run("scratch_lr3e-3", 4096, marks, scratch=True, use_graph=True, lr=3e-3)
if "seeds" in which:
    for sd in (1, 2):
        res[f"scratch_lr1e-3_s{sd}"] = run(f"scratch_lr1e-3_s{sd}", 4096, marks, scratch=True, use_graph=True, lr=1e-3, seed=sd)
json.dump(res, open(os.path.join(OUT, "diag_psnr.json"), "w"), indent=1)
