"""CPU: the oracle (oracle/*.c) against (1) golden vectors produced by the REFERENCE's own CUDA extensions on a B200
(tests/golden/*.npz, written by tests/golden/make_golden.py) and (2) its own invariants.  No GPU, no CUDA library."""
import os

import numpy as np
import pytest

import cases
import runners_cpu as rc

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def gold(name):
    p = os.path.join(GOLD, name + ".npz")
    if not os.path.exists(p):
        pytest.skip(f"{name}.npz not generated yet (run tests/golden/make_golden.py on the GPU box)")
    return np.load(p)


def rel(a, b):
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max()) / (float(np.abs(b).max()) + 1e-20)


def bits_equal(a, b):
    return a.shape == b.shape and np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


@pytest.mark.parametrize("name", ["march_b1", "march_b2"])
def test_oracle_march_matches_reference_golden(name):
    import oracle
    from golden.make_golden import GOLDEN_CASES
    g, k = gold(name), GOLDEN_CASES[name]
    c = cases.march_case(k["n"], k["seed"], k["bound"], k["dt_gamma"])
    assert bits_equal(c["nears"], g["nears"]) and bits_equal(c["fars"], g["fars"])
    r = rc.march_train(c)
    assert np.array_equal(r["counts"], g["counts"]) and np.array_equal(r["counter"], g["counter"])
    for key in ("xyzs", "dirs", "deltas"):
        assert bits_equal(r[key], g[key]), key
    N = c["rays_o"].shape[0]
    x, d, dl = oracle.march_rays(N, 1, np.arange(N, dtype=np.int32), c["nears"].copy(), c["rays_o"], c["rays_d"], c["bound"],
                                 c["bitfield"], c["cascade"], c["H"], c["nears"], c["fars"], np.zeros(N, np.float32),
                                 dt_gamma=c["dt_gamma"], max_steps=c["max_steps"])
    assert bits_equal(x, g["infer_xyzs"]) and bits_equal(dl, g["infer_deltas"])


def test_oracle_composite_matches_reference_golden():
    import oracle
    from golden.make_golden import GOLDEN_CASES
    g, k = gold("composite"), GOLDEN_CASES["composite"]
    c = cases.composite_case(k["n"], k["seed"], k["bound"])
    ws, dep, img = oracle.composite_rays_train_forward(c["sigmas"], c["rgbs"], c["deltas"], c["rays"])
    gs, gc = oracle.composite_rays_train_backward(c["grad_ws"], c["grad_image"], c["sigmas"], c["rgbs"], c["deltas"], c["rays"], ws, img)
    for a, key in ((ws, "ws"), (dep, "depth"), (img, "image"), (gs, "grad_sigmas"), (gc, "grad_rgbs")):
        assert rel(a, g[key]) <= 1e-3, key                       # __expf vs expf: tolerance, never bitwise


@pytest.mark.parametrize("half", [True, False])
def test_oracle_grid_matches_reference_golden(half):
    import oracle
    from golden.make_golden import GOLDEN_CASES
    g, k = gold("grid_f16" if half else "grid_f32"), GOLDEN_CASES["grid"]
    c = cases.grid_case(k["B"], k["seed"])
    exact = "device_level_scales" in g.files
    if exact:                                   # the level scales the GPU's exp2f produced -> bitwise comparison is meaningful
        oracle.grid_set_level_scales(g["device_level_scales"])
    try:
        out, dy = oracle.grid_encode_forward(c["inputs"], c["embeddings"], c["offsets"], c["per_level_scale"], c["H"], True, half=half)
    finally:
        oracle.grid_set_level_scales(None)
    if half:
        nbad = int((out.view(np.uint16) != g["out"].view(np.uint16)).sum())
        # with the host exp2f, level scales differ from the device's in the last bit and ~8 % of fp16 roundings flip
        assert nbad <= (0 if exact else out.size * 15 // 100), f"{nbad}/{out.size} fp16 encodings differ bitwise"
    assert rel(out.astype(np.float32), g["out"].astype(np.float32)) <= 1e-3
    assert rel(dy.astype(np.float32), g["dy_dx"].astype(np.float32)) <= 2e-3
    L = c["offsets"].shape[0] - 1
    g_lbc = np.ascontiguousarray(c["grad"].reshape(-1, L, 2).transpose(1, 0, 2))
    acc = oracle.grid_encode_backward(g_lbc, c["inputs"], c["offsets"], c["per_level_scale"], c["H"], 2, half=half).reshape(-1)
    ref = np.zeros_like(acc); ref[g["grad_idx"]] = g["grad_val"]
    # fp32: host-vs-device exp2f moves the interpolation weights at the 1e-4 level (see device_level_scales above)
    assert rel(acc, ref) <= (3e-2 if half else 5e-4)


@pytest.mark.parametrize("name", ["mlp_sigma", "mlp_color"])
def test_oracle_mlp_matches_reference_golden(name):
    import oracle
    from golden.make_golden import GOLDEN_CASES
    g, k = gold(name), GOLDEN_CASES[name]
    c = cases.mlp_case(**k)
    out, fb = oracle.ffmlp_forward(c["inputs"], c["weights"], c["input_dim"], 16, 64, c["num_layers"])
    # backward on the REFERENCE's stored activations: its fp16 accumulation flips the ReLU mask of a few near-zero units,
    # which changes whole gradient rows (a discontinuity, not an arithmetic difference)
    gw, bb, gi = oracle.ffmlp_backward(c["grad"], c["inputs"], c["weights"], g["fb"].astype(np.float16), c["input_dim"], 16, 64,
                                      c["num_layers"], calc_grad_inputs=True)
    # the reference accumulates in fp16 (ffmlp.cu:564): its distance from the exact oracle is a few fp16 ulps of the scale
    assert rel(out.astype(np.float32), g["out"]) <= 1e-2
    assert rel(fb.astype(np.float32), g["fb"]) <= 1e-2
    assert rel(gi.astype(np.float32), g["grad_inputs"]) <= 2e-2
    assert rel(gw, g["grad_weights"]) <= 2e-2


def test_oracle_freq_matches_reference_golden():
    import oracle
    from golden.make_golden import GOLDEN_CASES
    g, k = gold("freq"), GOLDEN_CASES["freq"]
    x = np.random.RandomState(k["seed"]).uniform(-1, 1, (k["B"], 3)).astype(np.float32)
    np.testing.assert_allclose(oracle.freq_encode_forward(x, k["deg"]), g["out"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(oracle.freq_encode_backward(g["grad"], g["out"], 3, k["deg"]), g["grad_inputs"], rtol=1e-4, atol=1e-4)


# ---------------------------------------------------------------- invariants that need no fixture
def test_march_invariants():
    c = cases.march_case(256, 41, 2.0, 1.0 / 128)
    r = rc.march_train(c)
    assert r["counter"][1] == c["rays_o"].shape[0] and r["counter"][0] == r["counts"].sum()
    miss = c["nears"] > 1e30
    assert miss.any() and np.all(r["counts"][miss] == 0)
    assert np.all(np.abs(r["xyzs"]) <= c["bound"]) and np.all(r["deltas"][:, 0] > 0)
    assert np.all(r["counts"] <= c["max_steps"])


def test_composite_invariants():
    import oracle
    c = cases.composite_case(256, 42, 1.0)
    ws, dep, img = oracle.composite_rays_train_forward(c["sigmas"], c["rgbs"], c["deltas"], c["rays"])
    assert np.all(ws >= 0) and np.all(ws <= 1 + 1e-5) and np.all(img <= 1 + 1e-5)
    # linearity in colour: compositing 2*rgb doubles the image, leaves weights unchanged
    ws2, _, img2 = oracle.composite_rays_train_forward(c["sigmas"], 2 * c["rgbs"], c["deltas"], c["rays"])
    np.testing.assert_allclose(img2, 2 * img, rtol=1e-6, atol=1e-7); assert np.array_equal(ws2, ws)
    # finite-difference check of grad_sigmas on a few samples
    gs, _ = oracle.composite_rays_train_backward(c["grad_ws"], c["grad_image"], c["sigmas"], c["rgbs"], c["deltas"], c["rays"], ws, img)
    loss = lambda s: float((oracle.composite_rays_train_forward(s, c["rgbs"], c["deltas"], c["rays"])[2].astype(np.float64) * c["grad_image"]).sum()
                           + (oracle.composite_rays_train_forward(s, c["rgbs"], c["deltas"], c["rays"])[0].astype(np.float64) * c["grad_ws"]).sum())
    ray = c["rays"][np.argmax(c["rays"][:, 2])]
    for i in (int(ray[1]), int(ray[1]) + 1):
        if c["sigmas"][i] > 100:
            continue
        s1, s0 = c["sigmas"].copy(), c["sigmas"].copy()
        h = max(1e-2, 1e-2 * float(c["sigmas"][i]))
        s1[i] += h; s0[i] -= h
        assert abs((loss(s1) - loss(s0)) / (2 * h) - gs[i]) <= 5e-2 * max(1.0, abs(gs[i]))


def test_foc_composite_selection_rule():
    import oracle
    c = cases.multi_case(K=3, N=64, T=32)
    on, of = oracle.near_far_from_aabb(c["rays_o"], c["rays_d"], c["aabb"], 0.2)
    img, dep, ws = oracle.foc_composite(c["sigmas"], c["rgbs"], on, of, 1.0)
    # object order matters only through ties: swapping objects 0 and 1 changes the colour picked at the tied sample 3
    sw = c["sigmas"][[1, 0, 2]], c["rgbs"][[1, 0, 2]]
    img2, _, ws2 = oracle.foc_composite(sw[0], sw[1], on, of, 1.0)
    np.testing.assert_allclose(ws2, ws, rtol=1e-6)
    assert np.all(img >= 0) and np.all(img <= 1)


def test_torch_renderer_restatement_matches_the_c_oracle():
    """oracle/torch_renderer.py (configs[0], the reference's PyTorch-only path on CPU): its torch restatements of the two CUDA-only
    operators agree with the C oracle, and its K = 1 render equals the oracle's FOC composite of the same densities / colours."""
    import torch
    import foc_eager
    import oracle
    from oracle import torch_renderer as tr
    o, d = cases.camera_rays(300, 4)
    aabb = np.array([-1] * 3 + [1] * 3, np.float32)
    n, f = oracle.near_far_from_aabb(o, d, aabb, 0.2)
    to, td, ta = torch.from_numpy(o), torch.from_numpy(d), torch.from_numpy(aabb)
    tn, tf = tr.near_far(to, td, ta)
    assert np.array_equal(tn.numpy(), n) and np.array_equal(tf.numpy(), f)
    x = np.random.RandomState(0).randn(257, 3).astype(np.float32)
    np.testing.assert_allclose(tr.freq_encode(torch.from_numpy(x), 4).numpy(), oracle.freq_encode_forward(x, 4), atol=2e-6, rtol=0)
    torch.manual_seed(1)
    model = tr.FieldNetwork().eval()
    with torch.no_grad():
        sig, rgb = foc_eager.run_object(model, tr.near_far, to, td, ta, 32)
        img, dep, ws = tr.render(model, to, td, ta, 32)
    hit = n < 1e30
    oi, od, ow = oracle.foc_composite(sig.numpy()[None], rgb.numpy()[None], n, f, 1.0)
    np.testing.assert_allclose(img.numpy()[hit], oi[hit, :3], atol=2e-5, rtol=0)
    np.testing.assert_allclose(ws.numpy()[hit], ow[hit], atol=2e-5, rtol=0)
    np.testing.assert_allclose(dep.numpy()[hit], od[hit], atol=2e-5, rtol=0)


@pytest.mark.parametrize("input_dim,hidden,num_layers", [(144, 16, 1), (48, 64, 2), (32, 128, 3), (16, 256, 2)])
def test_oracle_mlp_general_shapes_against_numpy(input_dim, hidden, num_layers):
    """The golden vectors only hold the field's 64-wide nets.  The oracle's code path for every other shape the reference module accepts
    (ffmlp.py:99-123; the shapes csrc/mlp_generic.cu serves, tests/test_gpu_mlp.py checks against this oracle) is pinned here against a
    plain numpy restatement: y = W_out relu(W_h ... relu(W_in x)), fp16 storage between layers, and the analytic gradients of that chain."""
    import oracle
    g = np.random.RandomState(5)
    B = 64
    std = np.sqrt(3 / hidden)
    n = hidden * (input_dim + hidden * (num_layers - 1) + 16)
    x = g.normal(0, 1, (B, input_dim)).astype(np.float16)
    w = g.uniform(-std, std, n).astype(np.float16)
    grad = g.normal(0, 1e-2, (B, 16)).astype(np.float16)
    out, fb = oracle.ffmlp_forward(x, w, input_dim, 16, hidden, num_layers)
    gw, bb, gi = oracle.ffmlp_backward(grad, x, w, fb, input_dim, 16, hidden, num_layers, calc_grad_inputs=True)
    # numpy chain
    mats, off = [], 0
    for o, i in [(hidden, input_dim)] + [(hidden, hidden)] * (num_layers - 1) + [(16, hidden)]:
        mats.append(w[off:off + o * i].astype(np.float64).reshape(o, i)); off += o * i
    acts, h = [], x.astype(np.float64)
    for m in mats[:-1]:
        h = np.maximum(h @ m.T, 0).astype(np.float16).astype(np.float64)
        acts.append(h)
    want = (h @ mats[-1].T).astype(np.float16)
    # the oracle rounds double -> float -> half (orc_mlp.c: (float)s, then rh()), numpy double -> half: a double rounding can differ by one
    # fp16 ulp in rare elements, which then moves later layers by a few ulps
    close = lambda a, b: np.allclose(a.astype(np.float64), b.astype(np.float64), rtol=4e-3, atol=1e-4)
    assert close(out, want) and float(np.mean(out.view(np.uint16) != want.view(np.uint16))) < 0.05
    for l in range(num_layers):
        assert close(fb[l], acts[l])
    acts = [fb[l].astype(np.float64) for l in range(num_layers)]        # the backward below goes through the ORACLE's stored activations
    # backward through the stored activations
    closeg = lambda a, b: np.allclose(a.astype(np.float64), b.astype(np.float64), rtol=4e-3, atol=2e-6)
    d = grad.astype(np.float64)
    gws = [None] * len(mats)
    gws[-1] = d.T @ acts[-1]
    for l in range(num_layers - 1, -1, -1):
        d = ((d @ mats[l + 1]).astype(np.float16).astype(np.float64) * (acts[l] > 0)).astype(np.float16).astype(np.float64)
        assert closeg(bb[num_layers - 1 - l], d), l
        d = bb[num_layers - 1 - l].astype(np.float64)                  # continue from the oracle's own (double-rounded) values
        prev = acts[l - 1] if l > 0 else x.astype(np.float64)
        gws[l] = d.T @ prev
    assert closeg(gi, (d @ mats[0]).astype(np.float16))
    np.testing.assert_allclose(gw, np.concatenate([m.reshape(-1) for m in gws]), rtol=1e-10, atol=1e-12)
