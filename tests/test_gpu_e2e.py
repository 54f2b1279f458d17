"""-m gpu: the path end to end -- train steps reduce the loss, the render loop matches with either compaction,
and the reference's own caller code (legacy/nerf/renderer.py semantics restated in trainer.py) runs on the operators."""
import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu


def _trainer(bound=1.0, **kw):
    from focnerf_b200.trainer import Trainer
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(bound)
    tr = Trainer(bound=bound, density_grid=torch.from_numpy(grid), **kw)
    assert np.array_equal(tr.density_bitfield.cpu().numpy(), bits)
    return scene, tr


@pytest.mark.parametrize("optimizer", ["fused", "torch"])
def test_train_steps_reduce_loss(optimizer):
    scene, tr = _trainer(optimizer=optimizer)
    losses = []
    for it in range(60):
        o, d = cases.camera_rays(4096, 100 + it)
        o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        gt = scene.render_gt(o, d)
        losses.append(float(tr.train_step(o, d, gt)))
        if (it + 1) % 16 == 0:
            tr.update_mean_count()
    assert all(np.isfinite(losses)), losses
    assert np.mean(losses[-10:]) < 0.5 * np.mean(losses[:5]), (losses[:5], losses[-10:])


def test_render_loop_compaction_equivalence():
    scene, tr = _trainer()
    o, d = cases.camera_rays(20000, 7)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    a = tr.render(o, d, device_compaction=True, mode="loop")
    b = tr.render(o, d, device_compaction=False, mode="loop")
    for x, y in zip(a, b):
        assert torch.equal(torch.nan_to_num(x), torch.nan_to_num(y))


def test_render_single_pass_matches_loop():
    """The single-pass render (all samples marched at once, one compositor launch) reproduces the reference-style loop."""
    scene, tr = _trainer()
    for it in range(48):                                           # a partly trained field, so early termination is exercised
        o, d = cases.camera_rays(4096, 300 + it)
        o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        tr.train_step(o, d, scene.render_gt(o, d))
        if (it + 1) % 16 == 0:
            tr.update_mean_count()
    o, d = cases.camera_rays(30000, 9)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    a = tr.render(o, d, mode="single_pass")
    b = tr.render(o, d, mode="loop")
    assert float(b[2].max()) > 0.5                                 # something was hit
    for x, y, tol in zip(a, b, (2e-3, 2e-3, 2e-3)):
        x, y = torch.nan_to_num(x), torch.nan_to_num(y)
        assert torch.allclose(x, y, atol=tol, rtol=0), float((x - y).abs().max())


def test_render_speculative_sizing_is_exact():
    """Trainer.render(single_pass) sizes the sample buffers from the previous frame and reads the exact total only after the frame
    is enqueued; render_frames does one read for a whole path.  Both must give the bits of the exactly-sized frame, also when the
    guess was too small (the frame is then rendered again)."""
    scene, tr = _trainer()
    for it in range(32):
        o, d = cases.camera_rays(4096, 500 + it)
        o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        tr.train_step(o, d, scene.render_gt(o, d))
    assert tr._render_rows_ok()
    o, d = cases.camera_rays(30000, 9)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    o2, d2 = cases.camera_rays(30000, 10)
    o2, d2 = torch.from_numpy(o2).cuda(), torch.from_numpy(d2).cuda()
    N = o.shape[0]
    exact = [t.clone() for t in tr.render(o, d)]                   # first frame of this size: no capacity yet -> exact path
    exact2 = None
    assert N in tr._render_cap and tr.render_overflows == 0
    same = lambda a, b: all(torch.equal(torch.nan_to_num(x), torch.nan_to_num(y)) for x, y in zip(a, b))
    assert same(tr.render(o, d), exact)                            # speculative path, the guess fits
    assert tr.render_overflows == 0
    tr._render_cap[N] = 16384                                      # far too small: most rays dropped by the march -> second, exact pass
    assert same(tr.render(o, d), exact)
    assert tr.render_overflows == 1 and tr._render_cap[N] > 16384
    import os
    os.environ["FOC_RENDER_SPECULATIVE"] = "0"
    try:
        exact2 = [t.clone() for t in tr.render(o2, d2)]
    finally:
        del os.environ["FOC_RENDER_SPECULATIVE"]
    outs = tr.render_frames([(o, d), (o2, d2), (o, d)])
    assert same(outs[0], exact) and same(outs[1], exact2) and same(outs[2], exact)
    tr._render_cap[N] = 16384
    outs = tr.render_frames([(o, d), (o2, d2)])                    # both overflow, both re-rendered after the single read
    assert same(outs[0], exact) and same(outs[1], exact2) and tr.render_overflows >= 3


def test_render_windowed_speculative_sizing_is_exact():
    """mode="windowed" with per-window capacities learnt from the previous frame (no host read inside the frame) gives the bits of the
    exactly-sized windowed render, also when a window's guess was too small, and through render_frames."""
    scene, tr = _trainer()
    for it in range(32):
        o, d = cases.camera_rays(4096, 700 + it)
        o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        tr.train_step(o, d, scene.render_gt(o, d))
    o, d = cases.camera_rays(30000, 9)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    wins = (16, 48, 960)
    same = lambda a, b: all(torch.equal(torch.nan_to_num(x), torch.nan_to_num(y)) for x, y in zip(a, b))
    exact = [t.clone() for t in tr.render(o, d, mode="windowed", windows=wins)]          # first frame: exact sizes, one read per window
    keys = [k for k in tr._render_cap if isinstance(k, tuple)]
    assert len(keys) == 1 and len(tr._render_cap[keys[0]]) == 3
    assert same(tr.render(o, d, mode="windowed", windows=wins), exact) and tr.render_overflows == 0
    tr._render_cap[keys[0]] = [16384, 16384, 16384]                                      # too small for the first windows
    assert same(tr.render(o, d, mode="windowed", windows=wins), exact) and tr.render_overflows == 1
    outs = tr.render_frames([(o, d), (o, d)], mode="windowed", windows=wins)
    assert same(outs[0], exact) and same(outs[1], exact)
    single = tr.render(o, d)                                                             # and it is the same picture as the single pass
    assert float((torch.nan_to_num(single[0]) - torch.nan_to_num(exact[0])).abs().max()) <= 2e-3


@pytest.mark.parametrize("windows", [(48, 976), (8, 24, 64, 928), (1024,)])
def test_render_windowed_matches_loop(windows):
    """Depth-windowed render (recorded march, K10 recurrence per window, alive flags) reproduces the reference-style loop."""
    scene, tr = _trainer()
    for it in range(48):
        o, d = cases.camera_rays(4096, 300 + it)
        o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        tr.train_step(o, d, scene.render_gt(o, d))
        if (it + 1) % 16 == 0:
            tr.update_mean_count()
    o, d = cases.camera_rays(30000, 9)
    o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    a = tr.render(o, d, mode="windowed", windows=windows)
    b = tr.render(o, d, mode="loop")
    assert float(b[2].max()) > 0.5
    for x, y in zip(a, b):
        x, y = torch.nan_to_num(x), torch.nan_to_num(y)
        assert torch.allclose(x, y, atol=2e-3, rtol=0), float((x - y).abs().max())


def test_forward_fused_matches_forward():
    """field_glue.cu path == the module-by-module wiring of network_ff.py:51-74, values and parameter gradients."""
    from focnerf_b200.network import NeRFNetwork
    torch.manual_seed(3)
    net = NeRFNetwork(bound=1.0).cuda()
    with torch.no_grad():
        net.encoder.embeddings.uniform_(-0.5, 0.5)         # the 1e-4 init leaves the MLP outputs degenerate
    B = 128 * 37
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.rand(B, 3, device="cuda", generator=g) * 2 - 1
    d = torch.nn.functional.normalize(torch.randn(B, 3, device="cuda", generator=g), dim=-1)
    ws, wc = torch.randn(B, device="cuda", generator=g), torch.randn(B, 3, device="cuda", generator=g)
    grads, outs = [], []
    for fn in (net.forward, net.forward_fused):
        net.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.float16):
            sigma, rgb = fn(x, d)
        ((sigma.float() * ws).sum() * 1e-2 + (rgb.float() * wc).sum()).backward()
        outs.append((sigma.detach().float(), rgb.detach().float()))
        grads.append([p.grad.detach().float().clone() for p in net.parameters()])
    assert outs[1][0].dtype == torch.float32 and outs[1][1].shape == (B, 3)
    assert torch.allclose(outs[0][0], outs[1][0], rtol=1e-5, atol=1e-7)
    assert torch.allclose(outs[0][1], outs[1][1], rtol=0, atol=1e-3)
    for a, b in zip(*grads):
        rel = float((a - b).norm() / (a.norm() + 1e-20))
        assert rel < 2e-2, rel                              # fp16 gradient rounding (different add order in the unfused cat/slice)


@pytest.mark.parametrize("pipeline", [False, True])
def test_graph_step_trains_and_tracks_counter(pipeline):
    """use_graph=True: the step replayed as one CUDA graph still trains, keeps the step_counter ring current, and
    re-captures when the sample-buffer quantum changes.  pipeline=True defers each optimiser sweep into the next replay."""
    from focnerf_b200.trainer import Trainer
    scene, cascade, grid, bits, thresh = cases.scene_bitfield(1.0)
    tr = Trainer(bound=1.0, density_grid=torch.from_numpy(grid), use_graph=True, pipeline=pipeline)
    losses, captured = [], set()
    for it in range(80):
        o, d = cases.camera_rays(4096, 100 + it)
        o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        gt = scene.render_gt(o, d)
        losses.append(float(tr.train_step(o, d, gt)))
        if tr._graph is not None:
            captured.add(tr._graph[0][0])
        if (it + 1) % 16 == 0:
            tr.update_mean_count()
    assert tr._graph is not None and len(captured) >= 1
    assert all(np.isfinite(losses)), losses
    # pipelined replay: the last step's gradients are still pending; flush() applies them exactly once
    assert tr._pending == pipeline
    steps_before = int(tr.optimizer.step_count.item())
    tr.flush()
    assert not tr._pending and int(tr.optimizer.step_count.item()) == steps_before + (1 if pipeline else 0)
    tr.flush()
    assert int(tr.optimizer.step_count.item()) == 80                # one optimiser step per train_step, none lost or doubled
    assert np.mean(losses[-10:]) < 0.5 * np.mean(losses[:5]), (losses[:5], losses[-10:])
    cnt = tr.step_counter[:, 0].cpu().numpy()
    assert (cnt > 0).all()
    # a forced quantum change re-captures
    tr.mean_count = max(captured) + 3 * Trainer.GRAPH_QUANTUM
    tr.train_step(o, d, gt)                                          # (after the flush above: an eager step that leaves gradients pending)
    tr.train_step(o, d, gt)
    assert tr._graph[0][0] > max(captured)


def test_update_extra_state_matches_reference_logic():
    """Density-grid refresh on the device (csrc/density_grid.cu) == the reference's steps (legacy/nerf/renderer.py:445-531)
    restated with torch ops on the same jittered positions: masked EMA-max, mean with -1 regions as 0, packbits at
    min(mean, density_thresh)."""
    from focnerf_b200 import raymarching as rm
    from focnerf_b200._lib import check, lib, ptr, stream
    scene, tr = _trainer()
    for it in range(20):                                            # a little training so the field is not flat
        o, d = cases.camera_rays(4096, 500 + it)
        o, d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        tr.train_step(o, d, scene.render_gt(o, d))
    H, n = tr.grid_size, tr.grid_size ** 3
    # cell positions: kernel vs the reference formula through morton3D_invert
    idx = torch.arange(n, dtype=torch.int32, device="cuda")
    noise = torch.rand(n, 3, device="cuda")
    xyz = torch.empty(n, 3, device="cuda")
    check(lib().foc_density_cell_positions(None, n, H, 1.0, ptr(noise), ptr(xyz), stream()), "cell_positions")
    coords = rm.morton3D_invert(idx).float()
    hgs = 1.0 / H
    want = (2 * coords / (H - 1) - 1) * (1.0 - hgs) + (noise * 2 - 1) * hgs
    assert torch.allclose(xyz, want, atol=1e-6)
    # one full update against the torch restatement, same RNG stream
    grid0 = tr.density_grid.clone()
    grid0[0, ::7] = -1.0                                             # some untrained (-1) cells must stay untouched
    tr.density_grid.copy_(grid0)
    torch.manual_seed(77)
    tr.update_extra_state(decay=0.95)
    torch.manual_seed(77)
    noise = torch.rand(n, 3, device="cuda")
    check(lib().foc_density_cell_positions(None, n, H, 1.0, ptr(noise), ptr(xyz), stream()), "cell_positions")
    tr.model.eval()
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
        tmp = tr.model.density(xyz)["sigma"].reshape(1, -1).float()
    valid = (grid0 >= 0) & (tmp >= 0)
    want_grid = torch.where(valid, torch.maximum(grid0 * 0.95, tmp), grid0)
    assert torch.equal(tr.density_grid, want_grid)
    mean = float(want_grid.clamp(min=0).mean())
    assert abs(float(tr.mean_density) - mean) <= 1e-5 * max(mean, 1e-6)
    want_bits = rm.packbits(want_grid, min(mean, tr.density_thresh))
    differing = int((want_bits ^ tr.density_bitfield).to(torch.int32).ne(0).sum())
    assert differing <= 2                                            # cells within float rounding of the mean threshold
    # partial updates run and keep the grid consistent
    tr.iter_density = 16
    tr.update_extra_state()
    assert torch.isfinite(tr.density_grid).all() and int(tr.density_bitfield.count_nonzero()) > 0


def test_partial_refresh_static_shapes_and_graph_replay():
    """Steady-state refresh (renderer.py:487-516) with static shapes: (a) the cells re-sampled from the occupied set ARE occupied
    and spread over it like the reference's `nonzero(...)[randint]` draw; (b) only sampled cells change, by the EMA-max rule;
    (c) the graph-replayed refresh gives the eager one's bits from the same RNG state; (d) two trainers with the same seed stay
    bit-identical (what data-parallel replicas rely on: no broadcast of the bitfield)."""
    def fresh(use_graph):
        scene, tr = _trainer(use_graph=use_graph, seed=3)
        tr.iter_density = 16
        return tr
    a, b = fresh(False), fresh(True)
    grid0 = a.density_grid.clone()
    for tr in (a, b):
        torch.manual_seed(11)
        tr.update_extra_state()                                       # b: warm-up + capture + first replay
    # (b) EMA-max rule on the cells that changed; untouched cells keep their value
    changed = a.density_grid != grid0
    assert 0.2 < float(changed.float().mean()) < 0.6                   # ~ 1 - exp(-0.25) random + the occupied draws, some of them unchanged
    assert bool((a.density_grid[changed] >= grid0[changed] * 0.95 - 1e-6).all())
    # (a) occupied draws: every occupied cell of a SMALL occupied set gets re-sampled (m = H^3/4 draws over ~1e5 cells)
    occ0 = grid0[0] > 0
    assert float(changed[0][occ0].float().mean()) > 0.9
    # (c) graph == eager, given the same generator state.  The capture run consumed the generator once more (warm-up), so compare a
    # second pair of calls started from a common state
    b.density_grid.copy_(a.density_grid); b.density_bitfield.copy_(a.density_bitfield)
    for tr in (a, b):
        torch.manual_seed(12)
        tr.update_extra_state()
    assert torch.equal(a.density_grid, b.density_grid) and torch.equal(a.density_bitfield, b.density_bitfield)
    assert b._refresh_graph is not None and a._refresh_graph is None
    # (d) same seed, same calls -> same bits
    c = fresh(True)
    c.density_grid.copy_(grid0)
    d = fresh(True)
    d.density_grid.copy_(grid0)
    for tr in (c, d):
        torch.manual_seed(13)
        tr.update_extra_state(); tr.update_extra_state()
    assert torch.equal(c.density_bitfield, d.density_bitfield) and torch.equal(c.density_grid, d.density_grid)


def test_mark_untrained_grid_vs_reference_logic():
    """csrc/density_grid.cu: k_mark_untrained against the reference's steps (legacy/nerf/renderer.py:380-443) restated with torch ops:
    same cells marked -1, up to cells whose frustum test falls within float rounding of a mask edge."""
    import synthetic
    from focnerf_b200 import raymarching as rm
    for bound, n_poses, radius in ((1.0, 5, 2.2), (2.0, 70, 3.0)):
        scene, tr = _trainer(bound)
        H, C = tr.grid_size, tr.cascade
        poses = synthetic.ring_poses(n_poses, radius=radius, seed=9)
        fx, fy, cx, cy = synthetic.intrinsics(800, 800, fovy_deg=30.0)
        grid0 = tr.density_grid.clone()
        got = tr.mark_untrained_grid(poses, (fx, fy, cx, cy))
        P = torch.from_numpy(poses).cuda()
        idx = torch.arange(H ** 3, dtype=torch.int32, device="cuda")
        coords = rm.morton3D_invert(idx).float()
        world = 2 * coords / (H - 1) - 1
        count = torch.zeros(C, H ** 3, device="cuda")
        margin = torch.full((C, H ** 3), 1e9, device="cuda")            # distance of each test from its mask edge, for the rounding band
        for cas in range(C):
            b = min(2 ** cas, bound)
            hgs = b / H
            xw = (world * (b - hgs)).unsqueeze(0)
            cam = (xw - P[:, :3, 3].unsqueeze(1)) @ P[:, :3, :3]
            z = cam[..., 2]
            ex = cx / fx * z + hgs * 2 - cam[..., 0].abs()
            ey = cy / fy * z + hgs * 2 - cam[..., 1].abs()
            count[cas] = ((z > 0) & (ex > 0) & (ey > 0)).sum(0)
            margin[cas] = torch.minimum(torch.minimum(z.abs(), ex.abs()), ey.abs()).amin(0)
        want = count == 0
        have = tr.density_grid == -1
        differ = want != have
        assert int((differ & (margin > 1e-5)).sum()) == 0                # any disagreement sits on a mask edge
        assert int(differ.sum()) <= 64
        assert got == int(have.sum()) and 0 < got < C * H ** 3
        assert torch.equal(tr.density_grid[~have], grid0[~have])         # seen cells untouched
