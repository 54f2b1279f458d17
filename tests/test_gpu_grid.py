"""-m gpu: hash-grid encoder vs the CPU oracle and the reference extension."""
import numpy as np
import pytest
import torch

import cases
import ref_ext
import runners

pytestmark = pytest.mark.gpu


def ours():
    from focnerf_b200.gridencoder.grid import _backend
    return _backend


def rel(a, b):
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max()) / (float(np.abs(b).max()) + 1e-20)


def frac_bad(a, b, rtol):
    """Fraction of entries further than rtol * max|b| apart (inf / nan patterns must agree).  Used where the CPU
    oracle's exp2f() can differ from the device's in the last bit of a non-integer level scale: a point within one ulp
    of a cell boundary then lands in the neighbouring cell, which changes dy_dx (discontinuous there) but not the
    encoding (continuous)."""
    a, b = a.astype(np.float64), b.astype(np.float64)
    fin = np.isfinite(a) & np.isfinite(b)
    assert np.array_equal(np.isfinite(a), np.isfinite(b))
    scale = float(np.abs(b[fin]).max()) + 1e-20
    return float((np.abs(a[fin] - b[fin]) > rtol * scale).mean())


@pytest.mark.parametrize("half", [True, False])
@pytest.mark.parametrize("res", [2048, None])
def test_grid_vs_oracle(half, res):
    import oracle
    c = cases.grid_case(4096, 3, desired_resolution=res)
    got = runners.grid_fwd_bwd(ours(), c, half=half)
    out, dy = oracle.grid_encode_forward(c["inputs"], c["embeddings"], c["offsets"], c["per_level_scale"], c["H"], True, half=half)
    assert rel(got["out"].astype(np.float32), out.astype(np.float32)) <= 1e-3
    assert np.all(got["out"][:, 0] == 0) and np.all(got["out"][:, 1] == 0)            # out-of-range points encode to zero
    assert frac_bad(got["dy_dx"], dy, 2e-3) <= 1e-3
    L = c["offsets"].shape[0] - 1
    g_lbc = np.ascontiguousarray(c["grad"].reshape(-1, L, 2).transpose(1, 0, 2))
    acc = oracle.grid_encode_backward(g_lbc, c["inputs"], c["offsets"], c["per_level_scale"], c["H"], 2, half=half)
    # fp16 atomics: order-dependent rounding; fp32: host-vs-device exp2f moves interpolation weights at the 1e-4 level
    assert rel(got["grad_emb"], acc) <= (2e-2 if half else 5e-4)
    gi = oracle.grid_input_backward(g_lbc.astype(np.float16 if half else np.float32), dy, c["inputs"].shape[0], 3, 2, L, half=half)
    assert frac_bad(got["grad_inputs"], gi, 2e-2 if half else 1e-4) <= 2e-3


@pytest.mark.parametrize("half", [True, False])
def test_grid_vs_reference(half):
    ref = ref_ext.load("gridencoder")
    if ref is None:
        pytest.skip("oracle/_ref/gridencoder not built")
    c = cases.grid_case(65536, 13)
    a, b = runners.grid_fwd_bwd(ours(), c, half=half), runners.grid_fwd_bwd(ref, c, half=half)
    if half:
        nbad = int((a["out"].view(np.uint16) != b["out"].view(np.uint16)).sum())
        assert nbad == 0, f"{nbad} of {a['out'].size} fp16 encodings differ from the reference bitwise"
    assert rel(a["out"].astype(np.float32), b["out"].astype(np.float32)) <= 1e-3
    assert rel(a["dy_dx"].astype(np.float32), b["dy_dx"].astype(np.float32)) <= 1e-3
    assert rel(a["grad_emb"], b["grad_emb"]) <= (3e-2 if half else 1e-4)
    assert rel(a["grad_inputs"], b["grad_inputs"]) <= (1e-2 if half else 1e-4)


def test_grid_module_blc_layout_matches_lbc():
    """GridEncoder (point-major [B, L*C] kernel layout) == permuted reference-layout output; backward == LBC backward."""
    from focnerf_b200.gridencoder import GridEncoder
    c = cases.grid_case(8192, 5)
    enc = GridEncoder(desired_resolution=2048).cuda()
    with torch.no_grad():
        enc.embeddings.copy_(runners.T(c["embeddings"]))
    x = (runners.T(c["inputs"]) * 2 - 1)
    with torch.autocast("cuda", dtype=torch.float16):
        y = enc(x, bound=1)
        (y.float() * runners.T(c["grad"])).sum().backward()
    y = y.detach()
    c2 = dict(c); c2["inputs"] = ((x + 1) / 2).cpu().numpy()
    r = runners.grid_fwd_bwd(ours(), c2, half=True, calc_grad_inputs=False)
    exp = np.ascontiguousarray(r["out"].transpose(1, 0, 2)).reshape(x.shape[0], -1)
    assert np.array_equal(y.cpu().numpy().view(np.uint16), exp.view(np.uint16))
    assert enc.embeddings.grad.dtype == torch.float32
    assert rel(enc.embeddings.grad.cpu().numpy(), r["grad_emb"]) <= 3e-2


def test_grad_total_variation_vs_oracle():
    import oracle
    from focnerf_b200.gridencoder import GridEncoder
    # power-of-two level scales: exp2f is exact on host and device, so every point falls in the same cell on both
    c = cases.grid_case(20000, 7, with_oob=False, desired_resolution=None)
    enc = GridEncoder().cuda()
    with torch.no_grad():
        enc.embeddings.copy_(runners.T(c["embeddings"]))
    enc.embeddings.grad = torch.zeros_like(enc.embeddings)
    enc.grad_total_variation(1e-3, inputs=runners.T(c["inputs"]) * 2 - 1, bound=1)
    x01 = ((runners.T(c["inputs"]) * 2 - 1 + 1) / 2).cpu().numpy()
    acc = oracle.grad_total_variation(x01, c["embeddings"], c["offsets"], c["per_level_scale"], c["H"], 1e-3)
    assert rel(enc.embeddings.grad.cpu().numpy(), acc) <= 1e-3


def test_grad_total_variation_vs_reference():
    """K14 (gridencoder.cu:506-645) against the reference extension's own grad_total_variation on the same cells: both ACCUMULATE
    into the passed gradient with fp32 atomics, so the comparison is within the fp32 accumulation-order tolerance."""
    ref = ref_ext.load("gridencoder")
    if ref is None:
        pytest.skip("oracle/_ref/gridencoder not built")
    for res in (None, 2048):                                  # power-of-two and non-integer level scales
        c = cases.grid_case(100000, 17, with_oob=False, desired_resolution=res)
        x, emb, off = runners.T(c["inputs"]), runners.T(c["embeddings"]), runners.T(c["offsets"])
        B, D, C, L = x.shape[0], 3, emb.shape[1], off.shape[0] - 1
        S, H = float(np.log2(c["per_level_scale"])), c["H"]
        seed = torch.from_numpy(np.random.RandomState(5).normal(0, 1e-6, c["embeddings"].shape).astype(np.float32)).cuda()
        ga, gb = seed.clone(), seed.clone()
        ours().grad_total_variation(x, emb, ga, off, 1e-3, B, D, C, L, S, H, 0, False)
        ref.grad_total_variation(x, emb, gb, off, 1e-3, B, D, C, L, S, H, 0, False)
        a, b = ga.cpu().numpy(), gb.cpu().numpy()
        assert float(np.abs(b - seed.cpu().numpy()).max()) > 0                       # the reference did add something
        assert rel(a, b) <= 1e-4, rel(a, b)
        # per element: |a - b| <= 1e-3 * max(|b|, eps), eps = 1e-3 of the tensor's scale (entries that cancel to ~0 carry no relative meaning)
        eps = 1e-3 * float(np.abs(b).max())
        assert np.all(np.abs(a - b) <= 1e-3 * np.maximum(np.abs(b), eps))
