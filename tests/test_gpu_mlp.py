"""-m gpu: fused tcgen05 MLP vs the CPU oracle and the reference extension (wmma, fp16 accumulate), plus the
frequency / SH encoders.  Tolerance: 1e-3 of each tensor's scale vs the exact-accumulation oracle; the reference's own
fp16-accumulation noise is larger than that, so against it the bound is the reference's distance from the oracle."""
import numpy as np
import pytest
import torch

import cases
import ref_ext
import runners

pytestmark = pytest.mark.gpu


def ours():
    from focnerf_b200.ffmlp.ffmlp import _backend
    return _backend


def rel(a, b):
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max()) / (float(np.abs(b).max()) + 1e-20)


def _oracle(c, calc_grad_inputs=True, fb_for_backward=None):
    """fb_for_backward: the stored activations of the implementation under test.  The backward is a function of those
    (ReLU mask + wgrad operand): a unit whose exact pre-activation is ~0 can be 0 in one implementation and 1 fp16 ulp in
    another, which switches a whole gradient row on or off -- a discontinuity, not an arithmetic error."""
    import oracle
    out, fb = oracle.ffmlp_forward(c["inputs"], c["weights"], c["input_dim"], 16, 64, c["num_layers"])
    fbb = fb if fb_for_backward is None else fb_for_backward.astype(np.float16)
    gw, bb, gi = oracle.ffmlp_backward(c["grad"], c["inputs"], c["weights"], fbb, c["input_dim"], 16, 64, c["num_layers"],
                                      calc_grad_inputs=calc_grad_inputs)
    return dict(out=out.astype(np.float32), fb=fb.astype(np.float32), grad_weights=gw, grad_inputs=None if gi is None else gi.astype(np.float32))


@pytest.mark.parametrize("input_dim,num_layers,B", [(32, 2, 128), (32, 2, 1280), (32, 3, 1280), (64, 2, 640), (32, 3, 128 * 700), (32, 2, 128 * 1001)])
def test_mlp_vs_oracle(input_dim, num_layers, B):
    c = cases.mlp_case(B, input_dim, 16, num_layers)
    got = runners.mlp_fwd_bwd(ours(), c)
    exp = _oracle(c, fb_for_backward=got["fb"])
    assert rel(got["out"], exp["out"]) <= 1e-3, "outputs"
    assert rel(got["out_inf"], exp["out"]) <= 1e-3, "inference outputs"
    assert rel(got["fb"], exp["fb"]) <= 1e-3, "forward_buffer"
    assert rel(got["grad_inputs"], exp["grad_inputs"]) <= 2e-3, "grad_inputs"
    assert rel(got["grad_weights"], exp["grad_weights"]) <= 2e-3, "grad_weights"


@pytest.mark.parametrize("input_dim,num_layers", [(32, 2), (32, 3)])
def test_mlp_vs_reference(input_dim, num_layers):
    ref = ref_ext.load("ffmlp")
    if ref is None:
        pytest.skip("oracle/_ref/ffmlp not built")
    ref.allocate_splitk(num_layers + 1)
    c = cases.mlp_case(8192, input_dim, 16, num_layers, seed=14)
    a, b = runners.mlp_fwd_bwd(ours(), c), runners.mlp_fwd_bwd(ref, c)
    exp_a, exp_b = _oracle(c, fb_for_backward=a["fb"]), _oracle(c, fb_for_backward=b["fb"])
    for k in ("out", "fb", "grad_inputs", "grad_weights"):
        ref_noise = rel(b[k], exp_b[k])                     # the reference's own fp16-accumulation error vs exact
        ours_err = rel(a[k], exp_a[k])
        assert ours_err <= 2e-3, f"{k}: ours vs exact {ours_err:.2e}"
        assert ours_err <= ref_noise + 1e-3, f"{k}: ours {ours_err:.2e} should be at least as close to exact as the reference {ref_noise:.2e}"
    # forward results: ours vs the reference directly, bounded by the reference's own distance from exact
    for k in ("out", "fb"):
        assert rel(a[k], b[k]) <= 2 * rel(b[k], exp_b[k]) + 1e-3, k


def test_ffmlp_module_matches_reference_wrapper_semantics():
    """FFMLP pads by 128 even when B % 128 == 0 and slices [:B, :output_dim] (ffmlp.py:157-165)."""
    import oracle
    from focnerf_b200.ffmlp import FFMLP
    net = FFMLP(32, 3, 64, 3).cuda()
    assert net.weights.numel() == 11264 and net.weights.dtype == torch.float32
    x = torch.randn(256, 32, device="cuda").half().requires_grad_()
    net.train()
    with torch.autocast("cuda", dtype=torch.float16):         # cast_inputs=torch.half only acts under autocast (ffmlp.py:18)
        y = net(x)
    assert y.shape == (256, 3)
    y.float().sum().backward()
    assert x.grad.shape == x.shape and net.weights.grad.shape == net.weights.shape
    out, _ = oracle.ffmlp_forward(x.detach().cpu().numpy(), net.weights.detach().half().cpu().numpy(), 32, 16, 64, 3)
    assert rel(y.detach().float().cpu().numpy(), out[:, :3].astype(np.float32)) <= 1e-3
    net.eval()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
        y2 = net(x.detach())
    assert torch.equal(y2, y.detach())


def _mlp_run(be, c, hidden, act=0):
    """runners.mlp_fwd_bwd for any hidden width / activation."""
    T = runners.T
    x, w, g = T(c["inputs"]), T(c["weights"]), T(c["grad"])
    B, IN, L = x.shape[0], c["input_dim"], c["num_layers"]
    fb = torch.empty(L, B, hidden, dtype=torch.float16, device="cuda")
    out = torch.empty(B, 16, dtype=torch.float16, device="cuda")
    be.ffmlp_forward(x, w, B, IN, 16, hidden, L, act, 6, fb, out)
    out_inf = torch.empty(B, 16, dtype=torch.float16, device="cuda")
    be.ffmlp_inference(x, w, B, IN, 16, hidden, L, act, 6, torch.empty(B, hidden, dtype=torch.float16, device="cuda"), out_inf)
    bb = torch.zeros(L, B, hidden, dtype=torch.float16, device="cuda")
    gi = torch.zeros(B, IN, dtype=torch.float16, device="cuda")
    gw = torch.zeros_like(w)
    be.ffmlp_backward(g, x, w, fb, B, IN, 16, hidden, L, act, 6, True, bb, gi, gw)
    torch.cuda.synchronize()
    return dict(out=out.float().cpu().numpy(), out_inf=out_inf.float().cpu().numpy(), fb=fb.float().cpu().numpy(), bb=bb.float().cpu().numpy(),
                grad_inputs=gi.float().cpu().numpy(), grad_weights=gw.float().cpu().numpy())


@pytest.mark.parametrize("input_dim,hidden,num_layers,act", [(16, 16, 2, 0), (144, 16, 2, 0), (48, 64, 3, 0), (32, 32, 2, 0), (32, 128, 3, 0),
                                                             (64, 256, 2, 0), (32, 128, 5, 0), (32, 64, 2, 3), (48, 32, 2, 6), (16, 64, 2, 1)])
def test_mlp_generic_shapes_vs_oracle(input_dim, hidden, num_layers, act):
    """Everything ffmlp.py:99-123 accepts besides the field's own shapes (hidden 16 / 32 / 128 / 256, input_dim = 16 m, e.g. the 144 -> 16 -> 16
    feature encoder and the 48-wide colour input of nerf/network_tcnn.py:504-546; activations relu / exponential / sigmoid / none) runs on the
    CUDA-core kernels of csrc/mlp_generic.cu -- same storage roundings, checked against the same oracle.  (32 / 64 -> 64 with a non-ReLU
    activation stays on the tcgen05 kernel: the (32, 64, 2, sigmoid) case.)"""
    import math
    import oracle
    g = np.random.RandomState(77)
    B = 128 * 5
    n = hidden * (input_dim + hidden * (num_layers - 1) + 16)
    std = math.sqrt(3 / hidden)
    c = dict(inputs=g.normal(0, 1, (B, input_dim)).astype(np.float16), weights=g.uniform(-std, std, n).astype(np.float16),
             grad=g.normal(0, 1e-2, (B, 16)).astype(np.float16), input_dim=input_dim, num_layers=num_layers)
    got = _mlp_run(ours(), c, hidden, act)
    out, fb = oracle.ffmlp_forward(c["inputs"], c["weights"], input_dim, 16, hidden, num_layers, activation=act)
    gw, bb, gi = oracle.ffmlp_backward(c["grad"], c["inputs"], c["weights"], got["fb"].astype(np.float16), input_dim, 16, hidden, num_layers,
                                      activation=act, calc_grad_inputs=True)
    tol = 1e-3 if act in (0, 6) else 3e-3            # __expf in the kernels vs expf in the oracle; the derivative through stored fp16 values
    assert rel(got["out"], out.astype(np.float32)) <= tol
    assert rel(got["out_inf"], out.astype(np.float32)) <= tol
    assert rel(got["fb"], fb.astype(np.float32)) <= tol
    if not (hidden == 64 and input_dim in (32, 64)):   # the tensor-core kernel keeps dL/dact in shared memory and does not write backward_buffer
        assert rel(got["bb"], bb.astype(np.float32)) <= 2 * tol
    assert rel(got["grad_inputs"], gi.astype(np.float32)) <= 2 * tol
    assert rel(got["grad_weights"], gw) <= 2 * tol


def test_ffmlp_module_generic_width_trains():
    """The module surface for a non-field shape: FFMLP(144, 16, 16, 2) (network_tcnn.py:504-515) forward + backward through autograd."""
    from focnerf_b200.ffmlp import FFMLP
    net = FFMLP(144, 16, 16, 2).cuda().train()
    x = torch.randn(300, 144, device="cuda").half().requires_grad_()
    with torch.autocast("cuda", dtype=torch.float16):
        y = net(x)
    assert y.shape == (300, 16)
    y.float().pow(2).sum().backward()
    assert torch.isfinite(net.weights.grad).all() and float(net.weights.grad.abs().max()) > 0 and x.grad.shape == x.shape
    w = net.weights.detach().half().float()
    h = torch.relu(x.detach().float() @ w[:144 * 16].view(16, 144).t()).half().float()
    h = torch.relu(h @ w[144 * 16:144 * 16 + 256].view(16, 16).t()).half().float()
    want = (h @ w[144 * 16 + 256:].view(16, 16).t())
    assert float((y.detach().float() - want).abs().max()) <= 2e-3 * float(want.abs().max())


def test_unsupported_shapes_raise():
    """What the reference module rejects is rejected (ffmlp.py:104-107) -- and so is a hidden width its kernels do not have."""
    from focnerf_b200.ffmlp import FFMLP
    for args in ((32, 3, 48, 2), (20, 3, 64, 2), (32, 17, 64, 2), (32, 3, 64, 1)):
        with pytest.raises(AssertionError):
            FFMLP(*args)


def test_freq_and_sh_encoders():
    import oracle
    from focnerf_b200.freqencoder.freq import _backend as fb
    from focnerf_b200.shencoder import sh_encode
    x = np.random.RandomState(0).uniform(-1, 1, (5000, 3)).astype(np.float32)
    for deg in (4, 10):
        out, gi, g = runners.freq(fb, x, deg)
        eo = oracle.freq_encode_forward(x, deg)
        np.testing.assert_allclose(out, eo, rtol=0, atol=2e-3 if deg == 10 else 1e-5)      # __sinf absolute error grows with |arg|
        np.testing.assert_allclose(gi, oracle.freq_encode_backward(g, out, 3, deg), rtol=1e-4, atol=1e-3)
        ref = ref_ext.load("freqencoder")
        if ref is not None:
            ro, rgi, _ = runners.freq(ref, x, deg)
            assert np.array_equal(out.view(np.uint32), ro.view(np.uint32)), "freq forward not bit-exact vs reference"
            np.testing.assert_allclose(gi, rgi, rtol=1e-5, atol=1e-5)
    d = x / np.linalg.norm(x, axis=-1, keepdims=True)
    np.testing.assert_allclose(sh_encode(runners.T(d)).cpu().numpy(), oracle.sh4_encode(d), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("num_layers", [2, 3])
def test_side_buffers_and_split_weight_gradient_reduction(num_layers):
    """FFMLP.enable_fp16_side_buffers + call_fused (weights read from the fp16 shadow, gradient written in place) gives the bits of the
    autograd path; with a WgradSide the reduction runs through foc_ffmlp_backward(grad_weights = NULL) + foc_ffmlp_wgrad_finish on a
    side stream and still gives the same bits."""
    from focnerf_b200.ffmlp import FFMLP
    from focnerf_b200.ffmlp.ffmlp import WgradSide
    torch.manual_seed(4)
    B = 128 * 300
    x0 = torch.randn(B, 32, device="cuda").half()
    go = (torch.randn(B, 16, device="cuda") * 1e-2).half()
    outs = []
    for mode in ("autograd", "side_buffers", "side_stream"):
        net = FFMLP(32, 16, 64, num_layers).cuda().train()
        x = x0.clone().requires_grad_()
        if mode != "autograd":
            g16, s16 = net.enable_fp16_side_buffers()
            if mode == "side_stream":
                net._wgrad_side = WgradSide()
        with torch.autocast("cuda", dtype=torch.float16):
            y = net.call_fused(x, True)
        y.backward(go)
        if mode == "side_stream":
            net._wgrad_side.join()
        torch.cuda.synchronize()
        gw = net.weights.grad.half() if mode == "autograd" else g16.clone()
        if mode != "autograd":
            assert net.weights.grad is None
        outs.append((y.detach().clone(), x.grad.clone(), gw))
    for o in outs[1:]:
        for a, b in zip(outs[0], o):
            assert torch.equal(a, b)
