"""-m gpu: ray marching through the C ABI vs the CPU oracle and vs the reference's own extension (oracle/_ref).
Integer results (counts, offsets, cell-derived sample positions) must be BIT-EXACT (BASELINE.json north_star)."""
import numpy as np
import pytest
import torch

import cases
import ref_ext
import runners

pytestmark = pytest.mark.gpu


def ours():
    from focnerf_b200.raymarching.backend import _backend
    return _backend


def _oracle_march(c, M=None):
    import oracle
    N = c["rays_o"].shape[0]
    M = N * c["max_steps"] if M is None else M
    x, d, dl, rays, counter = oracle.march_rays_train(c["rays_o"], c["rays_d"], c["bound"], c["bitfield"], c["cascade"], c["H"],
                                                      c["nears"], c["fars"], c["noises"], M, c["dt_gamma"], c["max_steps"])
    out = runners.canonical_march(rays, x, d, dl, M)
    out["counter"] = counter
    return out


def _assert_march_equal(a, b):
    assert np.array_equal(a["counts"], b["counts"]), f"{(a['counts'] != b['counts']).sum()} rays differ in sample count"
    assert np.array_equal(a["counter"], b["counter"])
    for k in ("xyzs", "dirs", "deltas"):
        assert a[k].shape == b[k].shape
        assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)), f"{k} not bit-exact"


# dt_gamma == 0 takes the warp-per-ray constant-step walk; bound 2 / 4 make it span several 1024-element chunks
@pytest.mark.parametrize("bound,dt_gamma", [(1.0, 0.0), (2.0, 1.0 / 128), (1.0, 1.0 / 128), (2.0, 0.0), (4.0, 0.0)])
def test_march_train_vs_oracle(bound, dt_gamma):
    c = cases.march_case(512, 1, bound, dt_gamma)
    got = runners.march_train(ours(), c)
    _assert_march_equal(got, _oracle_march(c))
    # deterministic, ray-ordered slots: offsets are the exclusive scan of the counts
    assert np.array_equal(got["offsets"], np.concatenate([[0], np.cumsum(got["counts"])[:-1]]))


@pytest.mark.parametrize("bound,dt_gamma", [(1.0, 0.0), (2.0, 1.0 / 128), (2.0, 0.0), (4.0, 0.0)])
def test_march_train_vs_reference(bound, dt_gamma):
    ref = ref_ext.load("raymarching")
    if ref is None:
        pytest.skip("oracle/_ref/raymarching not built")
    c = cases.march_case(2048, 11, bound, dt_gamma)
    a, b = runners.march_train(ours(), c), runners.march_train(ref, c)
    _assert_march_equal(a, b)
    # both slot assignments tile [0, total) without gaps or overlaps (ours in ray order, the reference's in arrival order)
    for r in (a, b):
        iv = sorted((int(o), int(n)) for o, n in zip(r["offsets"], r["counts"]) if n > 0)
        pos = 0
        for o, n in iv:
            assert o == pos
            pos += n
        assert pos == int(r["counter"][0])


def test_march_large_batch_multi_block_scan():
    """> 32768 rays take the three-launch scan (frame-sized march_rays_all): same per-ray results as the same rays marched in
    single-CTA-scan sized pieces, offsets = exclusive scan of the counts, counter = (total, N)."""
    c = cases.march_case(70001, 21, 1.0)                              # not a multiple of the 4096-ray scan block
    n = c["rays_o"].shape[0]
    got = runners.march_train(ours(), c, M=n * 128)
    assert np.array_equal(got["offsets"], np.concatenate([[0], np.cumsum(got["counts"])[:-1]]))
    assert int(got["counter"][0]) == int(got["counts"].sum()) <= n * 128 and int(got["counter"][1]) == n
    pos = 0
    # pieces below 16384 rays also run WITHOUT the conservative coarse-occupancy ray rejection that frame-sized calls use
    for s0 in range(0, n, 15000):
        sub = {k: (v[s0:s0 + 15000] if isinstance(v, np.ndarray) and v.shape[:1] == (n,) else v) for k, v in c.items()}
        m = sub["rays_o"].shape[0]
        part = runners.march_train(ours(), sub, M=m * 160)
        assert np.array_equal(part["counts"], got["counts"][s0:s0 + m])
        tot = int(part["counter"][0])
        assert tot <= m * 160 and part["xyzs"].shape[0] == tot
        assert np.array_equal(part["xyzs"].view(np.uint32), got["xyzs"][pos:pos + tot].view(np.uint32))
        assert np.array_equal(part["deltas"].view(np.uint32), got["deltas"][pos:pos + tot].view(np.uint32))
        pos += tot
    assert pos == int(got["counter"][0]) == got["xyzs"].shape[0]


def test_march_train_truncated():
    """mean_count underestimates (M too small): kept rays are exact, nothing is written past M."""
    c = cases.march_case(512, 3, 1.0)
    full = runners.march_train(ours(), c)
    M = int(full["counter"][0]) // 2
    got = runners.march_train(ours(), c, M=M)
    assert np.array_equal(got["counts"], full["counts"])            # counts are still reported for every ray
    assert int((got["counts"][got["kept"]]).sum()) <= M
    ref_x = _oracle_march(c, M=M)
    assert np.array_equal(got["xyzs"].view(np.uint32), ref_x["xyzs"].view(np.uint32))


def test_near_far_packbits_morton_vs_oracle():
    import oracle
    c = cases.march_case(1024, 5, 1.0)
    n, f = runners.near_far(ours(), c)
    assert np.array_equal(n.view(np.uint32), c["nears"].view(np.uint32)) and np.array_equal(f.view(np.uint32), c["fars"].view(np.uint32))
    _, _, grid, bits, thresh = cases.scene_bitfield(1.0)
    assert np.array_equal(runners.packbits(ours(), grid, thresh), bits)
    coords = np.random.RandomState(0).randint(0, 128, (5000, 3))
    idx, back = runners.morton(ours(), coords)
    assert np.array_equal(idx, oracle.morton3D(coords)) and np.array_equal(back, coords)
    ref = ref_ext.load("raymarching")
    if ref is not None:
        rn, rf = runners.near_far(ref, c)
        assert np.array_equal(n.view(np.uint32), rn.view(np.uint32)) and np.array_equal(f.view(np.uint32), rf.view(np.uint32))
        assert np.array_equal(runners.packbits(ref, grid, thresh), bits)
        ridx, _ = runners.morton(ref, coords)
        assert np.array_equal(idx, ridx)


def _field(scene):
    def f(xyzs, dirs):
        s, c = scene.sigma_color(xyzs, dirs)
        return s, c
    return f


def test_infer_loop_vs_reference_and_oracle():
    """march_rays + composite_rays through the caller's loop: sample positions bit-exact at every iteration,
    accumulated image / depth / weights within 1e-3 relative."""
    scene = cases.scene_bitfield(1.0)[0]
    c = cases.march_case(1500, 21, 1.0)
    a = runners.infer_loop(ours(), c, _field(scene))
    ref = ref_ext.load("raymarching")
    if ref is not None:
        b = runners.infer_loop(ref, c, _field(scene))
        assert len(a["trace"]) == len(b["trace"])
        for ta, tb in zip(a["trace"], b["trace"]):
            assert ta["n_alive"] == tb["n_alive"] and ta["n_step"] == tb["n_step"]
            assert np.array_equal(ta["xyzs"].view(np.uint32), tb["xyzs"].view(np.uint32))
            assert np.array_equal(ta["deltas"].view(np.uint32), tb["deltas"].view(np.uint32))
            assert np.array_equal(ta["alive_after"], tb["alive_after"])
        for k in ("ws", "depth", "image"):
            np.testing.assert_allclose(a[k], b[k], rtol=1e-3, atol=1e-5)
    # oracle: first iteration of the loop
    import oracle
    N = c["rays_o"].shape[0]
    x, d, dl = oracle.march_rays(N, 1, np.arange(N, dtype=np.int32), c["nears"].copy(), c["rays_o"], c["rays_d"], 1.0, c["bitfield"],
                                 c["cascade"], c["H"], c["nears"], c["fars"], np.zeros(N, np.float32))
    assert np.array_equal(a["trace"][0]["xyzs"][:N].view(np.uint32), x.view(np.uint32))
    assert np.array_equal(a["trace"][0]["deltas"][:N].view(np.uint32), dl.view(np.uint32))


def test_compact_alive():
    from focnerf_b200 import raymarching
    g = torch.Generator().manual_seed(0)
    for n in (1, 31, 2048, 2049, 640000):
        v = torch.randint(-1, 5, (n,), generator=g, dtype=torch.int32)
        v = torch.where(v < 0, v, torch.arange(n, dtype=torch.int32)).cuda()
        out, n_out = raymarching.compact_alive(v)
        exp = v[v >= 0]
        assert int(n_out.item()) == exp.numel()
        assert torch.equal(out[:exp.numel()], exp)


@pytest.mark.parametrize("n_rays,bound,dt_gamma", [(3000, 1.0, 0.0), (40000, 1.0, 0.0), (2000, 2.0, 1.0 / 128)])
def test_depth_windows_partition_the_march(n_rays, bound, dt_gamma):
    """WindowedMarch: the windows [0,16), [16,40), [40,max) of every ray, concatenated, are bitwise the ray's samples from
    march_rays_all; a ray marked dead gets no samples; offsets are the exclusive scan of the window counts."""
    from focnerf_b200 import raymarching
    c = cases.march_case(n_rays, 31, bound, dt_gamma=dt_gamma)
    n_rays = c["rays_o"].shape[0]                                     # + the adversarial rays
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    o, d, nears, fars, bits = T(c["rays_o"]), T(c["rays_d"]), T(c["nears"]), T(c["fars"]), T(c["bitfield"])
    x, dr, dl, rays = raymarching.march_rays_all(o, d, c["bound"], bits, c["cascade"], c["H"], nears, fars, 128, c["dt_gamma"], c["max_steps"])
    rays = rays.cpu().numpy(); x = x.cpu().numpy(); dl = dl.cpu().numpy()
    wm = raymarching.WindowedMarch(o, d, c["bound"], bits, c["cascade"], c["H"], nears, fars, c["dt_gamma"], c["max_steps"])
    assert np.array_equal(wm.rays.cpu().numpy(), rays)
    dead = np.zeros(n_rays, bool); dead[::7] = True
    taken = np.zeros(n_rays, np.int64)
    for start, width in ((0, 16), (16, 24), (40, c["max_steps"] - 40)):
        if start == 16:
            wm.alive[torch.from_numpy(dead).cuda()] = 0
        w = wm.window(start, width)
        rw = wm.rays_w.cpu().numpy()
        exp = np.clip(rays[:, 2] - start, 0, width)
        if start >= 16:
            exp[dead] = 0
        assert np.array_equal(rw[:, 2], exp)
        assert np.array_equal(rw[:, 1], np.concatenate([[0], np.cumsum(exp)[:-1]]))
        if w is None:
            assert exp.sum() == 0
            continue
        wx, wd, wdl = (t.cpu().numpy() for t in w)
        assert wx.shape[0] >= exp.sum() and wx.shape[0] % 128 == 0
        assert not wdl[int(exp.sum()):].any()                          # padding rows read as zero
        src = np.concatenate([np.arange(rays[n, 1] + start, rays[n, 1] + start + exp[n]) for n in np.nonzero(exp)[0]])
        assert np.array_equal(wx[:len(src)].view(np.uint32), x[src].view(np.uint32))
        assert np.array_equal(wdl[:len(src)].view(np.uint32), dl[src].view(np.uint32))
        taken += exp
    assert np.array_equal(taken[~dead], rays[~dead, 2])


def test_sph_from_ray_vs_oracle_and_reference():
    """K2 (raymarching.cu:163-198): background-sphere (theta, phi).  The kernel states its roundings (FMA chains where nvcc contracts
    the reference's expressions); atan2f / sqrtf are the same device routines, so results are compared bit for bit with the
    reference extension and within 2 ulp-scale (1e-6 absolute on values in [-1, 1]) with the host libm oracle."""
    import oracle
    c = cases.march_case(4096, 21, 1.0, adversarial=False)
    o, d = runners.T(c["rays_o"]), runners.T(c["rays_d"])
    N = o.shape[0]
    for radius in (2.5, 8.0):
        got = torch.empty(N, 2, device="cuda")
        ours().sph_from_ray(o, d, radius, N, got)
        exp = oracle.sph_from_ray(c["rays_o"], c["rays_d"], radius)
        g = got.cpu().numpy()
        assert np.isfinite(g).all() and np.abs(g).max() <= 1.0 + 1e-6
        assert np.abs(g - exp).max() <= 2e-6, np.abs(g - exp).max()
        ref = ref_ext.load("raymarching")
        if ref is not None:
            r = torch.empty(N, 2, device="cuda")
            ref.sph_from_ray(o, d, radius, N, r)
            nbad = int((r.cpu().numpy().view(np.uint32) != g.view(np.uint32)).sum())
            assert nbad == 0, f"{nbad} of {g.size} coordinates differ bitwise from the reference extension"
