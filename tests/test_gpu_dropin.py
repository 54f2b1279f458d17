"""-m gpu: "the reference's callers run unchanged" (BASELINE.json north_star; SURVEY.md 8 rows a17 / b / f4).

The files under oracle/_ref/py are byte-for-byte copies of the reference's own Python (oracle/build_ref.py: PY_FILES).  Here they
run on top of focnerf_b200.install_aliases():
  * nerf/network_ff.py NeRFNetwork  ->  legacy/nerf/renderer.py run_cuda (training and inference branches), update_extra_state
  * legacy/nerf/utils.py Trainer.train_one_epoch (the reference's whole training loop)
  * the reference's L1 operator wrappers (raymarching.py, grid.py, ffmlp.py, freq.py) over this repo's `_backend` objects
and are compared with this repo's own Trainer on the same parameters (reference checkpoint format, both directions).
"""
import argparse
import os
import warnings

import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu
warnings.filterwarnings("ignore")


def _callers():
    from oracle import ref_py
    if not ref_py.present():
        pytest.skip("oracle/_ref/py not present (built by oracle/build_ref.py in the build container)")
    ref_py.activate_callers()
    from nerf.network_ff import NeRFNetwork
    from nerf.utils import Trainer as RefTrainer
    return NeRFNetwork, RefTrainer


def _ref_model(NeRFNetwork, seed=0):
    torch.manual_seed(seed)
    m = NeRFNetwork(encoding="hashgrid", bound=1, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10, bg_radius=-1).cuda()
    # a trained-like table so that the composite sees real structure
    with torch.no_grad():
        m.encoder.embeddings.copy_(torch.randn_like(m.encoder.embeddings) * 0.3)
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    m.density_grid.copy_(torch.from_numpy(grid).cuda())
    m.density_bitfield.copy_(torch.from_numpy(bits).cuda())
    return m


def _rel(a, b):
    """max |a - b| / max |b| over the finite entries; non-finite entries must coincide (rays that miss the box have near = far =
    FLT_MAX, so the reference's own `(depth - near) / (far - near)` is NaN there, in both stacks)."""
    a, b = a.detach().float().cpu().numpy().astype(np.float64), b.detach().float().cpu().numpy().astype(np.float64)
    fa, fb = np.isfinite(a), np.isfinite(b)
    assert np.array_equal(fa, fb), "non-finite entries differ"
    return float(np.abs(a[fa] - b[fb]).max() / (np.abs(b[fb]).max() + 1e-20))


def test_reference_modules_resolve_to_this_package():
    NeRFNetwork, _ = _callers()
    import raymarching, gridencoder, ffmlp, encoding, activation          # noqa: E401  what the reference's files import
    import focnerf_b200
    assert raymarching is focnerf_b200.raymarching and gridencoder is focnerf_b200.gridencoder and ffmlp is focnerf_b200.ffmlp
    m = NeRFNetwork(encoding="hashgrid", bound=1, cuda_ray=True)
    assert type(m.encoder).__module__.startswith("focnerf_b200.") and type(m.sigma_net).__module__.startswith("focnerf_b200.")
    import nerf.renderer
    assert os.path.realpath(nerf.renderer.__file__).endswith("oracle/_ref/py/nerf/renderer.py")


def test_run_cuda_unmodified_matches_trainer():
    """One training-branch and one inference-branch run_cuda of the reference's renderer on this repo's operators vs Trainer."""
    NeRFNetwork, _ = _callers()
    from focnerf_b200.trainer import Trainer
    m = _ref_model(NeRFNetwork)
    tr = Trainer(bound=1.0, use_graph=False)
    missing, unexpected = tr.load_checkpoint({"model": m.state_dict(), "mean_count": 0, "mean_density": 0.0})     # reference format in
    assert not missing and not unexpected, (missing, unexpected)
    o, d = cases.camera_rays(2048, 77)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()

    m.train(); tr.model.train()
    with torch.autocast("cuda", dtype=torch.float16):
        ref = m.run_cuda(o, d, dt_gamma=0, bg_color=1, perturb=False, force_all_rays=False, max_steps=1024)
        img, depth, ws = tr.render_train(o, d, bg_color=1.0, perturb=False)
    assert int(m.step_counter[0, 0]) == int(tr.step_counter[0, 0]) > 0          # same sample count from the same bitfield: bit-exact march
    assert _rel(img, ref["image"]) <= 1e-3 and _rel(ws, ref["weights_sum"]) <= 1e-3 and _rel(depth, ref["depth"]) <= 1e-3

    m.eval()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
        ref = m.run_cuda(o, d, dt_gamma=0, bg_color=1, perturb=False, max_steps=1024)
    for mode, tol in (("loop", 1e-3), ("single_pass", 2e-3)):
        img, depth, ws = tr.render(o, d, mode=mode)
        assert _rel(img, ref["image"]) <= tol, mode
        assert _rel(depth, ref["depth"]) <= 5e-3, mode


def test_reference_trainer_loop_unmodified_trains_on_this_stack():
    """legacy/nerf/utils.py Trainer.train_one_epoch, unmodified, over this repo's operators: update_extra_state every 16 steps from an
    empty grid, autocast, GradScaler, Adam, per-step LambdaLR.  The loss must fall and the model must land where this repo's own
    Trainer lands on the same batches (same protocol, different code on top of the same kernels)."""
    NeRFNetwork, RefTrainer = _callers()
    import bench_common as C
    from focnerf_b200.trainer import Trainer
    dev = torch.device("cuda")
    scene = cases.scene_bitfield(1.0)[0]
    pool, _ = C.make_pool(scene, 16, 4096, 1000, dev, pinned=False)
    eo, ed, egt = C.heldout(scene, dev, n=8192)

    torch.manual_seed(0)
    model = NeRFNetwork(encoding="hashgrid", bound=1, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10, bg_radius=-1)
    opt = argparse.Namespace(patch_size=1, rand_pose=-1, color_space="srgb", update_extra_interval=16, dt_gamma=0.0, max_steps=1024, iters=30000, lr=1e-3)
    rt = RefTrainer("ngp", opt, model, device=dev, workspace=None, criterion=torch.nn.MSELoss(reduction="none"),
                    optimizer=lambda mm: torch.optim.Adam(mm.get_params(opt.lr), betas=(0.9, 0.99), eps=1e-15), ema_decay=None, fp16=True,
                    lr_scheduler=lambda o_: torch.optim.lr_scheduler.LambdaLR(o_, lambda it: 0.1 ** min(it / opt.iters, 1)),
                    scheduler_update_every_step=True, use_checkpoint="scratch", use_tensorboardX=False, mute=True)
    rt.error_map = None

    class Loader:
        batch_size = 1

        def __len__(self):
            return 192

        def __iter__(self):
            for i in range(192):
                o, d, gt = pool[i % len(pool)]
                yield {"rays_o": o[None], "rays_d": d[None], "images": gt[None]}

    rt.train_one_epoch(Loader())
    assert rt.global_step == 192 and model.iter_density == 12 and model.mean_count > 0
    assert rt.stats["loss"][-1] < 0.03                     # epoch-average loss; an untrained model sits at ~0.09
    model.eval()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
        img_ref = model.render(eo[None], ed[None], staged=True, bg_color=1, perturb=False, dt_gamma=0.0, max_steps=1024)["image"].reshape(-1, 3)
    psnr_ref = C.psnr_of(img_ref, egt)

    tr = Trainer(bound=1.0, lr=1e-3, iters=30000, seed=0, use_graph=True)
    for i in range(192):
        tr.fit_step(*pool[i % len(pool)])
    psnr_own = C.psnr_of(tr.render(eo, ed)[0], egt)
    assert psnr_ref > 20.0 and psnr_own > 20.0, (psnr_ref, psnr_own)
    assert abs(psnr_ref - psnr_own) < 1.0, (psnr_ref, psnr_own)       # early training is nearly seed-independent (tools/diag_seeds.py)

    # checkpoint interchange, reference -> this repo: the reference-trained model rendered by Trainer gives the reference's image
    tr2 = Trainer(bound=1.0, use_graph=False)
    tr2.load_checkpoint({"model": model.state_dict(), "mean_count": model.mean_count, "mean_density": model.mean_density})
    assert abs(C.psnr_of(tr2.render(eo, ed, mode="loop")[0], egt) - psnr_ref) < 0.05
    # ... and this repo -> reference: Trainer.save_checkpoint's file loads through the reference's own load_checkpoint
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "ngp.pth")
        tr.save_checkpoint(path)
        torch.manual_seed(1)
        model2 = NeRFNetwork(encoding="hashgrid", bound=1, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10, bg_radius=-1)
        rt2 = RefTrainer("ngp", opt, model2, device=dev, workspace=None, criterion=torch.nn.MSELoss(reduction="none"), fp16=True,
                         use_checkpoint="scratch", use_tensorboardX=False, mute=True)
        rt2.load_checkpoint(path, model_only=True)
        model2.eval()
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
            img2 = model2.render(eo[None], ed[None], staged=True, bg_color=1, perturb=False, dt_gamma=0.0, max_steps=1024)["image"].reshape(-1, 3)
        assert abs(C.psnr_of(img2, egt) - psnr_own) < 0.05


def test_reference_l1_wrappers_over_this_backend():
    """The reference's own raymarching.py / grid.py / ffmlp.py / freq.py (unmodified) with this repo's `_backend` objects as their
    pybind modules: same positional calls, same results as this repo's wrappers."""
    from oracle import ref_py
    if not ref_py.present():
        pytest.skip("oracle/_ref/py not present")
    import focnerf_b200.raymarching as own_rm
    from focnerf_b200.raymarching.backend import _backend as rm_be
    from focnerf_b200.gridencoder.grid import GridEncoder, _backend as ge_be
    from focnerf_b200.ffmlp.ffmlp import FFMLP, _backend as ff_be
    from focnerf_b200.freqencoder.freq import FreqEncoder, _backend as fq_be
    L1 = {p: ref_py.load_l1_wrapper(p, be) for p, be in (("raymarching", rm_be), ("gridencoder", ge_be), ("ffmlp", ff_be), ("freqencoder", fq_be))}
    c = cases.march_case(512, 5, 1.0)
    o, d = torch.from_numpy(c["rays_o"]).cuda(), torch.from_numpy(c["rays_d"]).cuda()
    aabb, bits = torch.from_numpy(c["aabb"]).cuda(), torch.from_numpy(c["bitfield"]).cuda()
    rm = L1["raymarching"]
    n1, f1 = rm.near_far_from_aabb(o, d, aabb, 0.2)
    n2, f2 = own_rm.near_far_from_aabb(o, d, aabb, 0.2)
    assert torch.equal(n1, n2) and torch.equal(f1, f2)
    cnt1, cnt2 = torch.zeros(2, dtype=torch.int32, device="cuda"), torch.zeros(2, dtype=torch.int32, device="cuda")
    x1, d1, dl1, r1 = rm.march_rays_train(o, d, 1.0, bits, 1, 128, n1, f1, cnt1, 0, False, 128, False, 0, 1024)
    x2, d2, dl2, r2 = own_rm.march_rays_train(o, d, 1.0, bits, 1, 128, n1, f1, cnt2, 0, False, 128, False, 0, 1024)
    assert torch.equal(cnt1, cnt2) and torch.equal(r1, r2) and torch.equal(x1, x2) and torch.equal(dl1, dl2)
    sig = torch.rand(x1.shape[0], device="cuda") * 20
    rgb = torch.rand(x1.shape[0], 3, device="cuda")
    for a, b in zip(rm.composite_rays_train(sig, rgb, dl1, r1, 1e-4), own_rm.composite_rays_train(sig, rgb, dl1, r1, 1e-4)):
        assert torch.equal(a, b)
    coords = torch.randint(0, 128, (1000, 3), dtype=torch.int32, device="cuda")
    assert torch.equal(rm.morton3D_invert(rm.morton3D(coords)), coords)

    torch.manual_seed(3)
    g_own = GridEncoder(desired_resolution=2048).cuda()
    g_ref = L1["gridencoder"].GridEncoder(desired_resolution=2048).cuda()
    g_ref.load_state_dict(g_own.state_dict())
    x = (torch.rand(4096, 3, device="cuda") * 2 - 1)
    with torch.autocast("cuda", dtype=torch.float16):
        assert torch.equal(g_ref(x, bound=1), g_own(x, bound=1))
    f_own, f_ref = FFMLP(32, 16, 64, 2).cuda(), L1["ffmlp"].FFMLP(32, 16, 64, 2).cuda()
    h = torch.randn(1000, 32, device="cuda")
    with torch.autocast("cuda", dtype=torch.float16):
        assert torch.equal(f_ref(h), f_own(h))
    e_own, e_ref = FreqEncoder(3, 4).cuda(), L1["freqencoder"].FreqEncoder(3, 4).cuda()
    assert torch.equal(e_ref(x), e_own(x))


def test_shadow_copies_follow_load_state_dict():
    """after enable_fp16_side_buffers every forward gathers from the fp16 shadow; a state_dict loaded behind the optimiser's
    back must reach it (version-checked refresh in GridEncoder / FFMLP)."""
    from focnerf_b200.trainer import Trainer
    tr = Trainer(bound=1.0, use_graph=False)
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    tr.set_density_grid(torch.from_numpy(grid))
    o, d = cases.camera_rays(1024, 3)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    before = tr.render(o, d)[0].clone()
    sd = {k: v.clone() for k, v in tr.model.state_dict().items()}
    sd["encoder.embeddings"] = torch.randn_like(sd["encoder.embeddings"]) * 0.5
    sd["color_net.weights"] = sd["color_net.weights"] * -1.0
    tr.model.load_state_dict(sd)                           # plain nn.Module API, not Trainer.load_checkpoint
    after = tr.render(o, d)[0]
    assert (after - before).abs().max() > 1e-2
    assert torch.equal(tr.model.encoder._shadow16, tr.model.encoder.embeddings.detach().half())
    assert torch.equal(tr.model.color_net._shadow16, tr.model.color_net.weights.detach().half())
