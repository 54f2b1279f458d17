"""-m gpu: ScaledAdam (csrc/optim.cu) against the pair it replaces on this path, torch.optim.Adam + torch.amp.GradScaler
(main_nerf.py:211,472; nerf/utils.py:866-868): same parameters after a sequence of steps that includes an overflow step
(skipped, scale halved) and scale growth; fp16 side buffers (shadow current, gradient zeroed)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SIZES = [100003, 4096, 12]          # one size that is not a multiple of the 4-wide vector path


def _make(seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return [torch.randn(n, device="cuda", generator=g) * 0.1 for n in SIZES]


@pytest.mark.parametrize("grad_dtype", [torch.float32, torch.float16])
def test_scaled_adam_matches_torch_adam_gradscaler(grad_dtype):
    from focnerf_b200.optim import ScaledAdam
    init = _make()
    pa = [torch.nn.Parameter(t.clone()) for t in init]
    pb = [torch.nn.Parameter(t.clone()) for t in init]
    ours = ScaledAdam(pa, lr=1e-2, betas=(0.9, 0.99), eps=1e-15, growth_interval=5)
    side = None
    if grad_dtype == torch.float16:                                 # half gradients arrive through attached side buffers
        side = [torch.zeros(n, dtype=torch.half, device="cuda") for n in SIZES]
        for p, g16 in zip(pa, side):
            ours.attach(p, grad16=g16)
    ref = torch.optim.Adam(pb, lr=1e-2, betas=(0.9, 0.99), eps=1e-15, fused=True)
    scaler = torch.amp.GradScaler("cuda", growth_interval=5)
    scaler.scale(torch.zeros(1, device="cuda"))                     # initialises the scale tensor
    gen = torch.Generator(device="cuda").manual_seed(1)
    for it in range(23):
        grads = [torch.randn(n, device="cuda", generator=gen) * 1e-3 for n in SIZES]
        if it in (7, 8, 15):
            grads[it % 3][3] = float("inf") if it != 8 else float("nan")
        s_ours, s_ref = float(ours.scale.item()), scaler.get_scale()
        assert s_ours == s_ref, (it, s_ours, s_ref)
        for i, (p, q, g) in enumerate(zip(pa, pb, grads)):
            scaled = (g * s_ref).to(grad_dtype)                     # what backward of (loss * scale) leaves behind
            q.grad = scaled.float().clone()
            if side is None:
                p.grad = scaled.clone()
            else:
                side[i].copy_(scaled)
        ours.step()
        scaler.step(ref)
        scaler.update()
    assert int(ours.step_count.item()) == 20                        # three skipped steps
    for p, q in zip(pa, pb):
        assert torch.allclose(p, q, rtol=2e-5, atol=1e-7), float((p - q).abs().max())
    for m, st in zip(ours.exp_avg, [ref.state[q] for q in pb]):
        assert torch.allclose(m, st["exp_avg"], rtol=1e-5, atol=1e-9)
    assert all(float(q.abs().max()) < 10 for q in pb)


def test_scaled_adam_side_buffers():
    from focnerf_b200.optim import ScaledAdam
    p = torch.nn.Parameter(_make(3)[0].clone())
    p0 = p.detach().clone()
    g16, s16 = torch.zeros(p.numel(), dtype=torch.half, device="cuda"), torch.empty(p.numel(), dtype=torch.half, device="cuda")
    opt = ScaledAdam([p], lr=1e-2)
    opt.attach(p, grad16=g16, shadow16=s16)
    assert torch.equal(s16, p0.half())
    g16.copy_((torch.randn_like(p0) * 1e-3 * 65536.0).half())
    assert opt.grad_tensors()[0] is g16
    opt.step()
    assert p.grad is None and not torch.equal(p.detach(), p0)
    assert torch.equal(s16, p.detach().half())                      # shadow follows the update
    assert int(g16.count_nonzero()) == 0                            # gradient consumed and zeroed
    # an overflow step: parameters untouched, gradient still zeroed, scale halved
    p1 = p.detach().clone()
    g16.fill_(1.0); g16[5] = float("inf")
    opt.step()
    assert torch.equal(p.detach(), p1) and int(g16.count_nonzero()) == 0 and float(opt.scale.item()) == 32768.0
    # checkpoint round trip
    sd = opt.state_dict()
    opt2 = ScaledAdam([p], lr=1.0)
    opt2.load_state_dict(sd)
    assert float(opt2.lr.item()) == pytest.approx(1e-2) and int(opt2.step_count.item()) == 1 and float(opt2.scale.item()) == 32768.0
