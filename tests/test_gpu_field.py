"""-m gpu: the one-kernel field evaluation (csrc/field_fwd.cu, foc_field_forward) and the head prologues of the MLP backward
(foc_ffmlp_backward_head) against the unfused chain they replace -- foc_ffmlp_forward x 2 + foc_field_mid_forward + foc_rgb_sigmoid_forward
and their backward kernels, each of which is itself tested against the oracle and the reference extension (test_gpu_mlp.py).
Same operations at the same rounding points: every output must be BIT-IDENTICAL."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _nets(seed):
    from focnerf_b200.ffmlp import FFMLP
    sn, cn = FFMLP(32, 16, 64, 2).cuda(), FFMLP(32, 3, 64, 3).cuda()
    g = torch.Generator(device="cuda").manual_seed(seed)
    with torch.no_grad():
        for n in (sn, cn):
            n.weights.copy_((torch.rand(n.weights.shape, generator=g, device="cuda") * 2 - 1) * 0.35)
    for n in (sn, cn):
        n.enable_fp16_side_buffers()
    return sn, cn


def _inputs(B, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    enc = (torch.randn(B, 32, generator=g, device="cuda") * 0.5).half()
    d = torch.randn(B, 3, generator=g, device="cuda")
    d = d / d.norm(dim=-1, keepdim=True)
    gs = torch.randn(B, generator=g, device="cuda") * 1e-2
    gr = torch.randn(B, 3, generator=g, device="cuda") * 1e-2
    return enc, d, gs, gr


def _unfused(sn, cn, enc, d):
    from focnerf_b200.field_glue import field_mid, rgb_sigmoid
    h = sn.call_fused(enc, enc.requires_grad)
    sigma, cin = field_mid(h, d)
    return sigma, rgb_sigmoid(cn.call_fused(cin, cin.requires_grad))


def _bits(t):
    return t.detach().contiguous().view(torch.uint8)


# 7 tiles (< one CTA's 6 groups + 1), 905 tiles (several per group, ragged tail), exactly one wave, one more than a wave
@pytest.mark.parametrize("tiles", [1, 7, 905, 148 * 6, 148 * 6 + 1])
def test_field_forward_and_backward_bit_identical_to_unfused_chain(tiles):
    from focnerf_b200.field_glue import field_eval
    B = 128 * tiles
    enc, d, gs, gr = _inputs(B, 100 + tiles)
    out = {}
    for fused in (False, True):
        sn, cn = _nets(7)
        sn.train(); cn.train()
        e = enc.clone().requires_grad_()
        with torch.autocast("cuda", dtype=torch.float16):
            sigma, rgb = field_eval(e, d, sn, cn, True) if fused else _unfused(sn, cn, e, d)
            ((sigma * gs).sum() + (rgb * gr).sum()).backward()
        out[fused] = (sigma, rgb, e.grad, sn._grad16.clone(), cn._grad16.clone())
    for name, a, b in zip(("sigma", "rgb", "grad_enc", "grad_w_sigma", "grad_w_colour"), out[False], out[True]):
        assert a.dtype == b.dtype and a.shape == b.shape, name
        nbad = int((_bits(a) != _bits(b)).sum())
        assert nbad == 0, f"{name}: {nbad} bytes differ between the fused field kernel and the unfused chain"
    assert torch.isfinite(out[True][0]).all() and float(out[True][3].float().abs().max()) > 0


def test_field_inference_and_gradient_accumulation():
    """Inference (no stored activations) gives the training forward's bits; two backward passes before one optimiser step add up in
    grad16 (`.grad` semantics); sigma without a gradient (g_sigma = None) is handled."""
    from focnerf_b200.field_glue import field_eval
    B = 128 * 300
    enc, d, gs, gr = _inputs(B, 5)
    sn, cn = _nets(9)
    sn.train(); cn.train()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
        s0, c0 = field_eval(enc, d, sn, cn, False)
    e = enc.clone().requires_grad_()
    with torch.autocast("cuda", dtype=torch.float16):
        s1, c1 = field_eval(e, d, sn, cn, True)
        assert torch.equal(s0, s1) and torch.equal(c0, c1)
        (c1 * gr).sum().backward()                                   # no gradient reaches sigma
    once_s, once_c = sn._grad16.clone(), cn._grad16.clone()
    assert float(once_c.float().abs().max()) > 0
    e2 = enc.clone().requires_grad_()
    with torch.autocast("cuda", dtype=torch.float16):
        s2, c2 = field_eval(e2, d, sn, cn, True)
        (c2 * gr).sum().backward()
    assert torch.equal(e.grad, e2.grad)
    twice = cn._grad16.float()
    assert torch.allclose(twice, 2 * once_c.float(), rtol=2e-3, atol=1e-6)          # fp16 rounding of the running sum


def test_trainer_step_with_and_without_the_fused_field_gives_the_same_losses(monkeypatch):
    """Through Trainer.train_step: the fused field is a drop-in for the four-launch chain."""
    import cases
    from focnerf_b200.trainer import Trainer
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    o, dd = cases.camera_rays(4096, 31)
    o, dd = torch.from_numpy(o).cuda(), torch.from_numpy(dd).cuda()
    gt = scene.render_gt(o, dd)
    losses = {}
    for fused in ("0", "1"):
        monkeypatch.setenv("FOC_FIELD_FUSED", fused)
        tr = Trainer(bound=1.0, density_grid=torch.from_numpy(grid), seed=4, use_graph=False)
        assert tr.model._field_fused == (fused == "1")
        ls = []
        for _ in range(3):
            torch.manual_seed(123)                                   # same march jitter in both runs
            ls.append(float(tr.train_step(o, dd, gt)))
        losses[fused] = ls
    # step 1 is bit-identical; later steps differ only through the fp16 atomics of the grid backward (arrival order)
    assert losses["0"][0] == losses["1"][0]
    assert np.allclose(losses["0"], losses["1"], rtol=2e-3)


def test_graph_step_with_device_side_row_count(monkeypatch):
    """The graph-replayed step with a fixed 1.5x allocation whose kernels read the valid row count from device memory (FOC_ROWS=1,
    include/foc_b200.h "device-side row counts") against the same step on mean-sized, fully processed buffers (FOC_ROWS=0): same
    samples, same losses (up to the fp16 atomics of the grid backward), the published count is the step's total rounded up to 128,
    and a drifting sample count does not re-capture the graph."""
    import cases
    from focnerf_b200.trainer import Trainer
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    batches = []
    for i in range(6):
        o, dd = cases.camera_rays(4096, 700 + i)
        o, dd = torch.from_numpy(o).cuda(), torch.from_numpy(dd).cuda()
        batches.append((o, dd, scene.render_gt(o, dd)))
    runs = {}
    for rows in ("0", "1"):
        monkeypatch.setenv("FOC_ROWS", rows)
        tr = Trainer(bound=1.0, density_grid=torch.from_numpy(grid), seed=2, use_graph=True, lr=1e-2)
        torch.manual_seed(55)
        losses, counts = [], []
        for it in range(48):
            losses.append(float(tr.train_step(*batches[it % len(batches)])))
            if tr.local_step % 16 == 0:
                tr.update_mean_count()
            counts.append(int(tr.step_counter[(tr.local_step - 1) % 16, 0]))
        assert tr._rows_mode() == (rows == "1")
        runs[rows] = (losses, counts, tr)
    l0, c0, _ = runs["0"]
    l1, c1, tr1 = runs["1"]
    assert c0[:16] == c1[:16]                                        # eager steps before the first capture: identical marches
    assert np.allclose(l0, l1, rtol=5e-2, atol=1e-4), (l0[-4:], l1[-4:])
    assert np.mean(l1[-8:]) < 0.5 * np.mean(l1[:4])                   # it trains
    g_rows = int(tr1._graph[5])
    m_alloc = tr1._graph[0][0]
    assert g_rows == min(m_alloc + 128, (c1[-1] + 127) // 128 * 128) and g_rows % 128 == 0
    assert m_alloc >= int(1.4 * tr1.mean_count)
    assert tr1.capture_count == 1                                     # sample counts moved every step; one capture served them all
