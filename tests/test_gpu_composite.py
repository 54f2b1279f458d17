"""-m gpu: compositors vs the CPU oracle and the reference extension.  Float results: 1e-3 relative (BASELINE.json),
measured against the scale of each tensor (atol = 1e-3 * max|ref|) because individual entries cross zero."""
import numpy as np
import pytest
import torch

import cases
import ref_ext
import runners

pytestmark = pytest.mark.gpu
RTOL = 1e-3


def ours():
    from focnerf_b200.raymarching.backend import _backend
    return _backend


def close(a, b, what):
    scale = float(np.abs(b).max()) + 1e-12
    err = float(np.abs(a - b).max())
    assert err <= RTOL * scale, f"{what}: max err {err:.3e} vs scale {scale:.3e}"


@pytest.mark.parametrize("bound", [1.0, 2.0])
def test_composite_train_vs_oracle_and_reference(bound):
    import oracle
    c = cases.composite_case(768, 2, bound)
    got = runners.composite_train(ours(), c)
    ws, dep, img = oracle.composite_rays_train_forward(c["sigmas"], c["rgbs"], c["deltas"], c["rays"])
    gs, gc = oracle.composite_rays_train_backward(c["grad_ws"], c["grad_image"], c["sigmas"], c["rgbs"], c["deltas"], c["rays"], ws, img)
    for k, v in (("ws", ws), ("depth", dep), ("image", img), ("grad_sigmas", gs), ("grad_rgbs", gc)):
        close(got[k], v, f"oracle {k}")
    # rays with no samples produce exact zeros
    empty = c["rays"][:, 2] == 0
    assert np.all(got["ws"][c["rays"][empty, 0]] == 0) and np.all(got["image"][c["rays"][empty, 0]] == 0)
    ref = ref_ext.load("raymarching")
    if ref is not None:
        r = runners.composite_train(ref, c)
        for k in ("ws", "depth", "image", "grad_sigmas", "grad_rgbs"):
            close(got[k], r[k], f"reference {k}")


def test_composite_train_autograd_and_overflow_rays():
    """Through the autograd Function; a ray whose slice overruns M is treated as empty (raymarching.cu:521-528)."""
    from focnerf_b200 import raymarching
    import oracle
    c = cases.composite_case(256, 8, 1.0)
    rays = c["rays"].copy()
    last = np.argmax(rays[:, 1] + rays[:, 2])
    rays[last, 2] += 10_000                                   # overruns M
    sig = runners.T(c["sigmas"]).requires_grad_()
    rgb = runners.T(c["rgbs"]).requires_grad_()
    ws, dep, img = raymarching.composite_rays_train(sig, rgb, runners.T(c["deltas"]), runners.T(rays), 1e-4)
    (img * runners.T(c["grad_image"])).sum().add((ws * runners.T(c["grad_ws"])).sum()).backward()
    ows, odep, oimg = oracle.composite_rays_train_forward(c["sigmas"], c["rgbs"], c["deltas"], rays)
    ogs, ogc = oracle.composite_rays_train_backward(c["grad_ws"], c["grad_image"], c["sigmas"], c["rgbs"], c["deltas"], rays, ows, oimg)
    close(img.detach().cpu().numpy(), oimg, "image"); close(sig.grad.cpu().numpy(), ogs, "grad_sigmas"); close(rgb.grad.cpu().numpy(), ogc, "grad_rgbs")
    assert float(ws[rays[last, 0]]) == 0.0


def test_composite_multi_vs_oracle():
    import oracle
    from focnerf_b200 import foc
    c = cases.multi_case(K=4, N=300, T=64)
    o, d, nears, fars, xyzs = foc.sample_uniform(runners.T(c["rays_o"]), runners.T(c["rays_d"]), runners.T(c["aabb"]), c["T"])
    on, of = oracle.near_far_from_aabb(c["rays_o"], c["rays_d"], c["aabb"], 0.2)
    assert np.array_equal(nears.cpu().numpy().view(np.uint32), on.view(np.uint32))
    for bg in (1.0, 0.0):
        img, dep, ws = foc.composite_multi(runners.T(c["sigmas"]), runners.T(c["rgbs"]), nears, fars, bg=bg)
        oi, od, ow = oracle.foc_composite(c["sigmas"], c["rgbs"], on, of, bg)
        hit = on < 1e30
        close(img.cpu().numpy()[hit], oi[hit], "image4"); close(ws.cpu().numpy()[hit], ow[hit], "ws"); close(dep.cpu().numpy()[hit], od[hit], "depth")
    # torch restatement of COMBINED.py:141-200, 247-251 on the same tensors (the reference's own eager math, on the GPU)
    sig, rgb = runners.T(c["sigmas"]), runners.T(c["rgbs"])
    best_s, best_c = sig[0].clone(), rgb[0].clone()
    for k in range(1, c["K"]):
        m = sig[k] > best_s
        best_c = torch.where(m.unsqueeze(-1), rgb[k], best_c); best_s = torch.maximum(sig[k], best_s)
    z = nears[:, None] + (fars - nears)[:, None] * torch.linspace(0, 1, c["T"], device="cuda")[None]
    deltas = torch.cat([z[:, 1:] - z[:, :-1], ((fars - nears) / c["T"])[:, None]], -1)
    alphas = 1 - torch.exp(-deltas * best_s)
    Ts = torch.cumprod(torch.cat([torch.ones_like(alphas[:, :1]), 1 - alphas + 1e-15], -1), -1)[:, :-1]
    w = alphas * Ts
    im4 = torch.cat([(w[..., None] * best_c).sum(1), (w * best_s).sum(1, keepdim=True)], -1) + (1 - w.sum(1, keepdim=True)) * 1.0
    img, dep, ws = foc.composite_multi(sig, rgb, nears, fars, bg=1.0)
    hit = (nears < 1e30).cpu().numpy()
    close(img.cpu().numpy()[hit], im4.clamp(0, 1).cpu().numpy()[hit], "torch restatement")


def test_sample_uniform_offset_and_masks():
    from focnerf_b200 import foc
    c = cases.multi_case(K=3, N=200, T=48, seed=6)
    ro, rd, aabb = runners.T(c["rays_o"]), runners.T(c["rays_d"]), runners.T(c["aabb"])
    o2, d2, n2, f2, x2 = foc.sample_uniform(ro, rd, aabb, c["T"], offset=c["offset"], normalize_d=True)
    # editable.py:443-471 + COMBINED.py:460-474 in eager torch
    eo = ro + runners.T(c["offset"]); ed = torch.nn.functional.normalize(rd, dim=-1)
    assert torch.equal(o2, eo)
    np.testing.assert_allclose(d2.cpu().numpy(), ed.cpu().numpy(), rtol=0, atol=2.5e-7)   # 1-2 ulp vs torch's norm reduction
    hit = n2 < 1e30
    z = n2[:, None] + (f2 - n2)[:, None] * torch.linspace(0, 1, c["T"], device="cuda")[None]
    ex = torch.min(torch.max(o2[:, None] + d2[:, None] * z[..., None], aabb[:3]), aabb[3:])
    np.testing.assert_allclose(x2[hit].cpu().numpy(), ex[hit].cpu().numpy(), rtol=0, atol=2e-6)
    # per-object colour masks: object 1 sampled along the shifted rays
    _, _, n0, f0, _ = foc.sample_uniform(ro, rd, aabb, c["T"])
    sig, rgb = runners.T(c["sigmas"]), runners.T(c["rgbs"])
    obj_n, obj_f = torch.stack([n0, n2, n0]), torch.stack([f0, f2, f0])
    img, _, _ = foc.composite_multi(sig, rgb, n0, f0, bg=1.0, obj_nears=obj_n, obj_fars=obj_f)
    masked = rgb.clone()
    for k in range(3):
        zk = obj_n[k][:, None] + (obj_f[k] - obj_n[k])[:, None] * torch.linspace(0, 1, c["T"], device="cuda")[None]
        dk = torch.cat([zk[:, 1:] - zk[:, :-1], ((obj_f[k] - obj_n[k]) / c["T"])[:, None]], -1)
        ak = 1 - torch.exp(-dk * sig[k])
        wk = ak * torch.cumprod(torch.cat([torch.ones_like(ak[:, :1]), 1 - ak + 1e-15], -1), -1)[:, :-1]
        masked[k] = rgb[k] * (wk > 1e-10).unsqueeze(-1)
    img2, _, _ = foc.composite_multi(sig, masked, n0, f0, bg=1.0)
    ok = ((n0 < 1e30) & (n2 < 1e30)).cpu().numpy()
    close(img.cpu().numpy()[ok], img2.cpu().numpy()[ok], "in-kernel masks vs pre-masked colours")


def test_composite_mse_matches_unfused():
    """composite_rays_train_mse == composite_rays_train + background mix + mse_loss, value and gradients; the fused
    backward zero-fills the rows it does not own (no pre-zeroed buffers)."""
    from focnerf_b200 import raymarching as rm
    c = cases.composite_case(600, 5)
    rays = c["rays"].copy()
    assert np.all(np.diff(rays[:, 1]) >= 0)                        # ray-ordered slices: the fused op's contract
    dev = "cuda"
    sig = torch.from_numpy(c["sigmas"]).to(dev).requires_grad_(True)
    rgb = torch.from_numpy(c["rgbs"]).to(dev).requires_grad_(True)
    deltas, rays_t = torch.from_numpy(c["deltas"]).to(dev), torch.from_numpy(rays).to(dev)
    N = rays.shape[0]
    gt = torch.rand(N, 3, device=dev, generator=torch.Generator(device=dev).manual_seed(3))
    for bg in (1.0, 0.3):
        for t in (sig, rgb):
            t.grad = None
        ws, depth, image = rm.composite_rays_train(sig, rgb, deltas, rays_t)
        loss_a = torch.nn.functional.mse_loss(image + (1 - ws).unsqueeze(-1) * bg, gt)
        (loss_a * 1024.0).backward()
        ga = (sig.grad.clone(), rgb.grad.clone())
        for t in (sig, rgb):
            t.grad = None
        loss_b, image_b, ws_b, depth_b = rm.composite_rays_train_mse(sig, rgb, deltas, rays_t, gt, bg)
        (loss_b * 1024.0).backward()
        assert torch.allclose(loss_a, loss_b, rtol=1e-5, atol=1e-8), (float(loss_a), float(loss_b))
        assert torch.equal(image, image_b) and torch.equal(ws, ws_b) and torch.equal(depth, depth_b)
        for a, b in zip(ga, (sig.grad, rgb.grad)):
            assert torch.isfinite(b).all()
            assert torch.allclose(a, b, rtol=1e-4, atol=1e-6 * float(a.abs().max())), float((a - b).abs().max())


def test_composite_mse_hc_equals_sigmoid_then_mse():
    """composite_rays_train_mse_hc(sigma, hc) == composite_rays_train_mse(sigma, rgb_sigmoid(hc)): same loss / image bits, the same
    grad_sigma bits and the same grad_hc bits (columns 3..15 and rows without a gradient are zero)."""
    from focnerf_b200 import raymarching as rm
    from focnerf_b200.field_glue import rgb_sigmoid
    c = cases.composite_case(600, 8)
    rays = c["rays"].copy()
    dev = "cuda"
    M, N = c["sigmas"].shape[0], rays.shape[0]
    g = torch.Generator(device=dev).manual_seed(11)
    deltas, rays_t = torch.from_numpy(c["deltas"]).to(dev), torch.from_numpy(rays).to(dev)
    gt = torch.rand(N, 3, device=dev, generator=g)
    hc0 = (torch.randn(M, 16, device=dev, generator=g) * 2).half()
    for bg, seed_grad in ((1.0, 1024.0), (0.3, 65536.0)):
        sig_a = torch.from_numpy(c["sigmas"]).to(dev).requires_grad_(True)
        hc_a = hc0.clone().requires_grad_(True)
        loss_a, image_a, ws_a, depth_a = rm.composite_rays_train_mse(sig_a, rgb_sigmoid(hc_a), deltas, rays_t, gt, bg)
        loss_a.backward(gradient=torch.tensor(seed_grad, device=dev))
        sig_b = torch.from_numpy(c["sigmas"]).to(dev).requires_grad_(True)
        hc_b = hc0.clone().requires_grad_(True)
        loss_b, image_b, ws_b, depth_b = rm.composite_rays_train_mse_hc(sig_b, hc_b, deltas, rays_t, gt, bg)
        loss_b.backward(gradient=torch.tensor(seed_grad, device=dev))
        assert torch.equal(loss_a, loss_b) and torch.equal(image_a, image_b) and torch.equal(ws_a, ws_b) and torch.equal(depth_a, depth_b)
        assert torch.equal(sig_a.grad, sig_b.grad)
        assert hc_b.grad.dtype == torch.float16 and torch.equal(hc_a.grad, hc_b.grad)
        assert float(hc_b.grad[:, 3:].abs().max()) == 0.0 and float(hc_b.grad[:, :3].abs().max()) > 0.0
