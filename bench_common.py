"""What both arms of bench.py share: the workload protocol, the synthetic batches, the clock sampler, the evaluation rays.
Imports neither focnerf_b200 nor the reference -- `bench.py --impl reference` must not load the product library.

THE PROTOCOL (configs[1]; the reference's own recipe, main_nerf.py + legacy/nerf/utils.py:831-870)
  * model: hash grid 16 levels x 2 features, 2^19 entries, base 16, finest 2048 x bound; FFMLP 32-64-64-16 / 32-64-64-64-16 (network_ff.py)
  * 4096 rays per step (main_nerf.py:28) drawn from a pool of POOL pre-generated batches of the synthetic scene (analytic GT)
  * Adam(lr 1e-3, betas (0.9, 0.99), eps 1e-15) (main_nerf.py:26,472), LambdaLR 0.1 ** min(iter / ITERS, 1) stepped every
    iteration (main_nerf.py:25,477; ITERS = the run length PRETRAIN), fp16 autocast + GradScaler, MSE, white background, perturb=True
  * the density grid starts EMPTY and `update_extra_state` runs whenever global_step % 16 == 0 (nerf/utils.py:853-856): the
    bitfield the marcher reads is the learnt one, and the refresh is INSIDE every timed region
  * PRETRAIN steps of exactly that loop bring the model to its steady state before anything is timed; the same PRETRAIN steps
    are the "fixed-iteration run" of the quality gate (PSNR on EVAL_RAYS held-out rays, QUALITY_SEEDS seeds)
"""
import math
import os
import subprocess
import sys
import threading

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

RAYS_PER_GPU = 4096
POOL = 256                      # distinct pre-generated ray batches (1 M rays): held-out views are then interpolations, seed spread < 0.1 dB
PRETRAIN = 1024                 # steps of the protocol before the timed region = the fixed-iteration run of the quality gate
# `--iters` = the length of the run, as the reference's recipe is used (main_nerf.py:25,477: lr decays to 0.1 x lr at the last
# iteration and stays there).  Measured on this scene (tools/diag_seeds.py, 5 seeds): seed-to-seed sd of the held-out PSNR is 0.05 dB
# with the schedule ending at the run length against 0.2-0.3 dB with an effectively constant lr -- a 0.1 dB gate needs the former.
LR, ITERS, UPDATE_EXTRA_INTERVAL = 1e-3, PRETRAIN, 16
QUALITY_SEEDS = (0, 1, 2, 3, 4)
EVAL_RAYS = 65536
BOUND, MIN_NEAR, DT_GAMMA, MAX_STEPS, DENSITY_THRESH, T_THRESH = 1.0, 0.2, 0.0, 1024, 10.0, 1e-4
SHARED_CKPT = os.environ.get("FOC_SHARED_CKPT", "/tmp/foc_b200_shared_ckpt.pth")
WORKLOAD = ("configs[1]: single-object hash-grid NeRF train step (16 levels, 2^19 table, F=2, 64-wide FFMLPs, 4096 rays/batch per GPU, "
            "cuda_ray, fp16, Adam lr 1e-3) in the reference's training loop: learnt density grid refreshed every 16 steps inside the "
            f"timed region, model in its steady state after {PRETRAIN} steps from scratch")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed regions (B200_PROFILING.md).  20 ms period; the caller keeps
    it running across warm-up + timed + e2e loops so that short runs still collect samples."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0, period_ms=20):
        self.index, self.rows, self.proc, self.period = index, [], None, period_ms

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index),
                                          "-lms", str(self.period)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def scene_and_grid():
    """-> (scene, cascade, analytic density grid [C, H^3] numpy, its bitfield, threshold)."""
    import cases
    return cases.scene_bitfield(BOUND)


def make_pool(scene, n_batches, rays, seed, device, pinned=True):
    """Pre-generated batches: device copies for `value`, pinned host copies for `e2e`."""
    import cases
    import torch
    dev, host = [], []
    for i in range(n_batches):
        o, d = cases.camera_rays(rays, seed + i)
        o, d = torch.from_numpy(o).to(device), torch.from_numpy(d).to(device)
        gt = scene.render_gt(o, d)
        dev.append((o, d, gt))
        if pinned:
            host.append(tuple(t.cpu().pin_memory() for t in (o, d, gt)))
    return dev, host


def heldout(scene, device, n=EVAL_RAYS, seed=424242):
    import cases
    import torch
    o, d = cases.camera_rays(n, seed)
    o, d = torch.from_numpy(o).to(device), torch.from_numpy(d).to(device)
    return o, d, scene.render_gt(o, d)


def psnr_of(image, gt):
    import torch
    mse = float(torch.mean((image.float().reshape(gt.shape) - gt) ** 2).item())
    return -10.0 * math.log10(max(mse, 1e-12))


def coverage_of(image):
    """Fraction of pixels that are not background-white (same definition in both arms)."""
    return float(((1.0 - image.float()).amax(-1) > 0.02).float().mean().item())


def mean_sd(xs):
    m = sum(xs) / len(xs)
    sd = math.sqrt(sum((x - m) ** 2 for x in xs) / (len(xs) - 1)) if len(xs) > 1 else 0.0
    return m, sd


def frame_rays(H=800, W=800):
    """configs[2]: one 800x800 test pose of the synthetic scene -> rays_o, rays_d [H*W, 3] (CPU tensors)."""
    import synthetic
    pose = synthetic.ring_poses(8, radius=2.2, seed=123)[3]
    return synthetic.get_rays(pose, synthetic.intrinsics(H, W), H, W)


def multi_rays(H=480, W=640):
    """configs[3]/[4]: the 640x480 table-top view (fx = fy = 554, radius 1.8) -> rays_o, rays_d [H*W, 3] (CPU tensors)."""
    import synthetic
    pose = synthetic.ring_poses(8, radius=1.8, seed=321)[2]
    o, d = synthetic.get_rays(pose, (554.0, 554.0, W / 2.0, H / 2.0), H, W)
    return o.contiguous(), d.contiguous()


def multi_object_tables(models, embeddings_of):
    """K independently seeded object fields (the workload does not depend on the weights: the reference composites with no early
    termination); tables drawn wide so densities differ from place to place.  Same seeds in both arms -> the same objects."""
    import torch
    for k, m in enumerate(models):
        torch.manual_seed(4000 + k)                  # after construction: FFMLP.reset_parameters re-seeds the global generator (ffmlp.py:139)
        with torch.no_grad():
            emb = embeddings_of(m)
            emb.copy_(torch.randn_like(emb) * 1.5)
    return models


EDIT_OFFSET = (0.01, 0.01, 0.60)                     # editable.py:77-79 defaults (--x_offset / --y_offset / --z_offset), applied to object 0
