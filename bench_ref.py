"""Reference arm of bench.py: the REFERENCE's own code on the same B200, same workload, same protocol (bench_common.py).

What runs here is the reference itself, unmodified, through its own public API and stock code path:
  * L0  its CUDA extensions, compiled from /root/reference/{raymarching,gridencoder,ffmlp,freqencoder}/src for sm_100a
        (oracle/_ref/*/_ref_*.so, recipe oracle/build_ref.py);
  * L1  its Python operator wrappers raymarching.py / grid.py / ffmlp.py (oracle/_ref/py, verbatim copies);
  * L2  its network and renderer: nerf/network_ff.py NeRFNetwork -> legacy/nerf/renderer.py run_cuda / update_extra_state;
  * L3  its training loop: legacy/nerf/utils.py Trainer.train_one_epoch / train_step (update_extra_state every 16 steps,
        autocast, GradScaler, Adam, per-step LambdaLR, loss.item(), tqdm) fed by an in-memory loader of synthetic batches.
Absent from the reference tree and therefore supplied (plain torch, oracle/ref_shims): `encoding.get_encoder` and the SH-4
direction encoder.  Third-party imports the callers never use on this path are stubs (oracle/stubs).  EMA is off (ema_decay=None:
the legacy Trainer updates it once per epoch and uses it for evaluation only).

Nothing from focnerf_b200 is imported or mapped: asserted from sys.modules and /proc/self/maps before the JSON line is printed.
The reference has no CPU implementation of this path (every operator `.cuda()`s its inputs) and no multi-GPU path.
"""
import argparse
import math
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import bench_common as C  # noqa: E402


def _assert_reference_only():
    bad = [m for m in sys.modules if m == "focnerf_b200" or m.startswith("focnerf_b200.")]
    assert not bad, f"reference arm imported product modules: {bad}"
    with open("/proc/self/maps") as f:
        assert "libfoc_b200" not in f.read(), "reference arm mapped libfoc_b200.so"


class PoolLoader:
    """What nerf/provider.py's DataLoader hands to train_one_epoch: dicts with rays_o / rays_d [1, N, 3] and images [1, N, 3] on the
    device (provider.py collates on the GPU).  host=True: every batch starts in pinned host memory and is copied inside the loop."""
    batch_size = 1

    def __init__(self, pool, first, count, device, host=False):
        self.pool, self.first, self.count, self.device, self.host = pool, first, count, device, host

    def __len__(self):
        return self.count

    def __iter__(self):
        for i in range(self.count):
            o, d, gt = self.pool[(self.first + i) % len(self.pool)]
            if self.host:
                o, d, gt = (t.to(self.device, non_blocking=True) for t in (o, d, gt))
            yield {"rays_o": o.unsqueeze(0), "rays_d": d.unsqueeze(0), "images": gt.unsqueeze(0)}


def make_trainer(dev, seed, iters, workspace=None):
    """main_nerf.py:417-480 for `--ff --cuda_ray --fp16` (the FOC fork's defaults: lr 1e-3, density_thresh 10, min_near 0.2)."""
    import torch
    import torch.optim as optim
    from nerf.network_ff import NeRFNetwork
    from nerf.utils import Trainer
    torch.manual_seed(seed)
    model = NeRFNetwork(encoding="hashgrid", bound=C.BOUND, cuda_ray=True, density_scale=1, min_near=C.MIN_NEAR,
                        density_thresh=C.DENSITY_THRESH, bg_radius=-1)
    opt = argparse.Namespace(patch_size=1, rand_pose=-1, color_space="srgb", update_extra_interval=C.UPDATE_EXTRA_INTERVAL,
                             dt_gamma=C.DT_GAMMA, max_steps=C.MAX_STEPS, iters=iters, lr=C.LR)
    trainer = Trainer("ngp", opt, model, device=dev, workspace=workspace, criterion=torch.nn.MSELoss(reduction="none"),
                      optimizer=lambda m: torch.optim.Adam(m.get_params(opt.lr), betas=(0.9, 0.99), eps=1e-15), ema_decay=None, fp16=True,
                      lr_scheduler=lambda o: optim.lr_scheduler.LambdaLR(o, lambda it: 0.1 ** min(it / opt.iters, 1)),
                      scheduler_update_every_step=True, use_checkpoint="scratch", use_tensorboardX=False, mute=True)
    trainer.error_map = None                              # Trainer.train() takes it from the dataset (nerf/utils.py:644); no error-map sampling here
    return trainer


def render_rays(trainer, o, d):
    """Trainer.test_step's call (nerf/utils.py:593-607) -> image [N,3], depth [N] through NeRFRenderer.render / run_cuda (inference)."""
    import torch
    trainer.model.eval()
    with torch.no_grad(), torch.cuda.amp.autocast(enabled=True):
        out = trainer.model.render(o.unsqueeze(0), d.unsqueeze(0), staged=True, bg_color=1, perturb=False, **vars(trainer.opt))
    return out["image"].reshape(-1, 3), out["depth"].reshape(-1)


def render_bench(trainer, dev, frames):
    import torch
    o, d = C.frame_rays()
    o, d = o.contiguous().to(dev), d.contiguous().to(dev)
    n = o.shape[0]
    render_rays(trainer, o, d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(frames):
        image, depth = render_rays(trainer, o, d)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / frames
    return {"metric": "800x800 render Mrays/s", "value": n / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_frame": ms, "frames": frames, "rays": n,
            "coverage": C.coverage_of(image), "mean_image": float(image.mean().item()),
            "workload": "configs[2]: one 800x800 test pose, NeRFRenderer.run_cuda inference loop (march_rays / composite_rays, T_thresh 1e-4, "
                        "max_steps 1024) on the model this arm trained"}


def multi_bench(dev, K, frames, offsets):
    """configs[3] / [4]-edit on the reference stack: COMBINED.py's own MultiTrainer.run / best_densities_and_colors_v3 /
    image_depth_generation (+ editable.py's ray offsets), extracted verbatim by oracle/foc_extract.py, over K resident
    nerf/network_ff.py objects, in the reference's 4096-ray batches, whole 640x480 frames."""
    import torch
    from nerf.network_ff import NeRFNetwork
    from oracle import foc_extract
    models = [NeRFNetwork(encoding="hashgrid", bound=C.BOUND, cuda_ray=False, density_scale=1, min_near=C.MIN_NEAR,
                          density_thresh=C.DENSITY_THRESH, bg_radius=-1).to(dev).eval() for _ in range(K)]
    C.multi_object_tables(models, lambda m: m.encoder.embeddings)
    o, d = C.multi_rays()
    o, d = o.to(dev), d.to(dev)
    n = o.shape[0]
    comp = foc_extract.Compositor(models, num_steps=512, offsets=offsets)
    comp.render(o[:8192], d[:8192])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(frames):
        image, depth = comp.render(o, d)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / frames
    return {"metric": "640x480 multi-object composite Mrays/s", "value": n / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_frame": ms, "frames": frames,
            "rays": n, "objects": K, "samples_per_ray": 512, "offsets": offsets is not None, "field_evaluations_per_s": n * K * 512 / (ms * 1e-3),
            "mean_image": float(image.mean().item()), "mean_depth": float(depth.mean().item()),
            "workload": f"COMBINED.py's own per-image loop body (oracle/foc_extract.py) over {K} resident nerf/network_ff.py objects, whole frame in "
                        "4096-ray batches (the reference also reloads every checkpoint from disk per image: not timed)"}


def run(args):
    """-> the JSON dict for `bench.py --impl reference` (rank 0, one GPU)."""
    import warnings
    warnings.filterwarnings("ignore")
    import torch
    from oracle import ref_py
    ref_py.activate_reference()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    scene, cascade, grid, bits, thresh = C.scene_and_grid()
    t_pool = time.time()
    pool_dev, pool_host = C.make_pool(scene, C.POOL, C.RAYS_PER_GPU, 1000, dev)
    t_pool = time.time() - t_pool
    eo, ed, egt = C.heldout(scene, dev)
    ws_dir = tempfile.mkdtemp(prefix="foc_ref_ws_")

    # ---- fixed-iteration run = pretraining of the timed model (seed 0) ------------------------------------------------------
    clocks = C.ClockSampler(0)
    clocks.start()
    trainer = make_trainer(dev, C.QUALITY_SEEDS[0], C.ITERS, workspace=ws_dir)
    t0 = time.time()
    trainer.train_one_epoch(PoolLoader(pool_dev, 0, C.PRETRAIN, dev))
    torch.cuda.synchronize()
    pretrain_s = time.time() - t0
    psnrs = [C.psnr_of(render_rays(trainer, eo, ed)[0], egt)]

    # ---- timed region: K more steps of the same loop -------------------------------------------------------------------------
    warm = max(3, args.warmup)
    trainer.train_one_epoch(PoolLoader(pool_dev, C.PRETRAIN, warm, dev))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    trainer.train_one_epoch(PoolLoader(pool_dev, C.PRETRAIN + warm, args.steps, dev))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    samples = float(trainer.model.step_counter[:, 0].float().mean().item())
    # end to end: pinned host batches, H2D inside the loop; the loop itself reads loss.item() every step (nerf/utils.py:872)
    trainer.train_one_epoch(PoolLoader(pool_host, 7, 3, dev, host=True))
    torch.cuda.synchronize()
    t0e, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0e.record()
    trainer.train_one_epoch(PoolLoader(pool_host, 10, args.steps, dev, host=True))
    t1e.record()
    torch.cuda.synchronize()
    ms_e2e = t0e.elapsed_time(t1e)
    clk = clocks.stop()
    rays = C.RAYS_PER_GPU * args.steps
    val, val_e2e = rays / (ms * 1e-3), rays / (ms_e2e * 1e-3)

    render = render_bench(trainer, dev, args.frames) if args.frames > 0 else None
    # the reference's own checkpoint format, written by its own save_checkpoint: the other arm renders THIS model too
    shared = None
    try:
        trainer.save_checkpoint(full=False, best=False)
        path = sorted(os.listdir(os.path.join(ws_dir, "checkpoints")))[-1]
        shutil.copyfile(os.path.join(ws_dir, "checkpoints", path), C.SHARED_CKPT)
        shared = C.SHARED_CKPT
    except Exception as e:                                # never fail the arm over the convenience file
        shared = f"not written: {type(e).__name__}: {e}"

    # ---- quality: the other seeds of the fixed-iteration run -----------------------------------------------------------------
    quality = None
    if args.psnr_steps > 0:
        for s in C.QUALITY_SEEDS[1:args.psnr_seeds]:
            tq = make_trainer(dev, s, C.ITERS)
            tq.train_one_epoch(PoolLoader(pool_dev, 0, C.PRETRAIN, dev))
            psnrs.append(C.psnr_of(render_rays(tq, eo, ed)[0], egt))
            del tq
            torch.cuda.empty_cache()
    m, sd = C.mean_sd(psnrs)
    quality = {"psnr_db": m, "psnr_sd_db": sd, "psnr_per_seed": psnrs, "seeds": list(C.QUALITY_SEEDS[:len(psnrs)]), "steps": C.PRETRAIN,
               "eval_rays": C.EVAL_RAYS, "note": "fresh model per seed, the protocol's first PRETRAIN steps from an empty density grid, held-out rays"}

    multi, multi8 = None, None
    if getattr(args, "multi_frames", 0) > 0:
        multi = multi_bench(dev, 4, args.multi_frames, None)
        multi8 = multi_bench(dev, 8, 1, [C.EDIT_OFFSET] + [None] * 7)
    shutil.rmtree(ws_dir, ignore_errors=True)
    _assert_reference_only()
    return {
        "impl": "reference", "metric": "train rays/s", "value": val, "unit": "rays/s", "n_gpus": 1, "steps": args.steps, "warmup": warm,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {"workload": C.WORKLOAD, "rays_per_gpu": C.RAYS_PER_GPU, "samples_per_step": samples, "pretrain_steps": C.PRETRAIN,
                   "pretrain_s": pretrain_s, "pool_batches": C.POOL, "pool_build_s": t_pool,
                   "note": "the reference's own stack, unmodified: CUDA extensions (sources compiled for sm_100a), L1 wrappers, nerf/network_ff.py, "
                           "legacy/nerf/renderer.py, legacy/nerf/utils.py Trainer.train_one_epoch; no product module imported or mapped; the reference has "
                           "no CPU implementation of this path and no multi-GPU path, so this arm always runs on one GPU"},
        "clocks": clk, "render": render, "multi_object": multi, "multi_object_edit8": multi8, "quality": quality, "shared_checkpoint": shared,
        "cpu_baseline": {"value": val, "unit": "rays/s", "cores": 0, "kind": "reference",
                         "sample": f"{args.steps} train steps of 4096 rays through the reference's own Trainer on its own CUDA extensions (GPU; the reference has no CPU path)"},
        "e2e": {"value": val_e2e, "unit": "rays/s", "ms_per_step": ms_e2e / args.steps, "h2d_bytes_per_step": C.RAYS_PER_GPU * 36, "d2h_bytes_per_step": 4},
    }
