"""Drive a `_backend` (ours, or the reference's pybind module from oracle/_ref -- both expose the SAME positional
signatures) on the numpy cases of tests/cases.py and return numpy results.  GPU only."""
import numpy as np
import torch


def T(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return t if dtype is None else t.to(dtype)


def canonical_march(rays, xyzs, dirs, deltas, M):
    """Ray-ordered view of a march result: counts per ray id and each kept ray's samples concatenated in id order.
    The reference hands out slots in atomic arrival order (raymarching.cu:405-406), ours by prefix scan."""
    rays = rays[np.argsort(rays[:, 0], kind="stable")]
    counts = rays[:, 2].copy()
    keep = (counts > 0) & (rays[:, 1].astype(np.int64) + counts <= M)
    xs, ds, dl = [], [], []
    for off, cnt in rays[keep][:, 1:3]:
        xs.append(xyzs[off:off + cnt]); ds.append(dirs[off:off + cnt]); dl.append(deltas[off:off + cnt])
    cat = lambda l, w: np.concatenate(l, 0) if l else np.zeros((0, w), np.float32)
    return dict(counts=counts, kept=keep, xyzs=cat(xs, 3), dirs=cat(ds, 3), deltas=cat(dl, 2), offsets=rays[:, 1].copy())


def near_far(be, c):
    o, d, aabb = T(c["rays_o"]), T(c["rays_d"]), T(c["aabb"])
    N = o.shape[0]
    nears, fars = torch.empty(N, device="cuda"), torch.empty(N, device="cuda")
    be.near_far_from_aabb(o, d, aabb, N, 0.2, nears, fars)
    return nears.cpu().numpy(), fars.cpu().numpy()


def packbits(be, grid, thresh):
    g = T(grid.reshape(-1))
    N = g.numel() // 8
    out = torch.empty(N, dtype=torch.uint8, device="cuda")
    be.packbits(g, N, float(thresh), out)
    return out.cpu().numpy()


def morton(be, coords):
    cds = T(coords.astype(np.int32))
    idx = torch.empty(cds.shape[0], dtype=torch.int32, device="cuda")
    be.morton3D(cds, cds.shape[0], idx)
    back = torch.empty(cds.shape[0], 3, dtype=torch.int32, device="cuda")
    be.morton3D_invert(idx, cds.shape[0], back)
    return idx.cpu().numpy(), back.cpu().numpy()


def march_train(be, c, M=None):
    o, d, bits = T(c["rays_o"]), T(c["rays_d"]), T(c["bitfield"])
    nears, fars, noises = T(c["nears"]), T(c["fars"]), T(c["noises"])
    N = o.shape[0]
    M = N * c["max_steps"] if M is None else M
    xyzs, dirs, deltas = torch.zeros(M, 3, device="cuda"), torch.zeros(M, 3, device="cuda"), torch.zeros(M, 2, device="cuda")
    rays = torch.zeros(N, 3, dtype=torch.int32, device="cuda")
    counter = torch.zeros(2, dtype=torch.int32, device="cuda")
    be.march_rays_train(o, d, bits, c["bound"], c["dt_gamma"], c["max_steps"], N, c["cascade"], c["H"], M, nears, fars, xyzs, dirs,
                        deltas, rays, counter, noises)
    torch.cuda.synchronize()
    out = canonical_march(rays.cpu().numpy(), xyzs.cpu().numpy(), dirs.cpu().numpy(), deltas.cpu().numpy(), M)
    out["counter"] = counter.cpu().numpy()
    out["raw"] = (xyzs, dirs, deltas, rays)
    return out


def composite_train(be, c):
    sig, rgb, dl, rays = T(c["sigmas"]), T(c["rgbs"]), T(c["deltas"]), T(c["rays"])
    M, N = sig.shape[0], rays.shape[0]
    ws, dep, img = torch.empty(N, device="cuda"), torch.empty(N, device="cuda"), torch.empty(N, 3, device="cuda")
    be.composite_rays_train_forward(sig, rgb, dl, rays, M, N, 1e-4, ws, dep, img)
    gs, gc = torch.zeros(M, device="cuda"), torch.zeros(M, 3, device="cuda")
    be.composite_rays_train_backward(T(c["grad_ws"]), T(c["grad_image"]), sig, rgb, dl, rays, ws, img, M, N, 1e-4, gs, gc)
    return dict(ws=ws.cpu().numpy(), depth=dep.cpu().numpy(), image=img.cpu().numpy(), grad_sigmas=gs.cpu().numpy(),
                grad_rgbs=gc.cpu().numpy())


def infer_loop(be, c, field, max_iters=64, T_thresh=1e-4):
    """The caller-owned render loop (legacy/nerf/renderer.py:341-367) with an analytic `field(xyzs, dirs) -> sigma, rgb`.
    Returns per-iteration march outputs (bit-comparable) and the final accumulators."""
    o, d, bits = T(c["rays_o"]), T(c["rays_d"]), T(c["bitfield"])
    nears, fars = T(c["nears"]), T(c["fars"])
    N = o.shape[0]
    ws, dep, img = torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda"), torch.zeros(N, 3, device="cuda")
    alive = torch.arange(N, dtype=torch.int32, device="cuda")
    rays_t = nears.clone()
    trace, step, it = [], 0, 0
    while step < c["max_steps"] and it < max_iters:
        n_alive = alive.shape[0]
        if n_alive == 0:
            break
        n_step = max(min(N // n_alive, 8), 1)
        M = n_alive * n_step
        M += 128 - M % 128
        xyzs, dirs, deltas = torch.zeros(M, 3, device="cuda"), torch.zeros(M, 3, device="cuda"), torch.zeros(M, 2, device="cuda")
        noises = torch.zeros(n_alive, device="cuda")
        be.march_rays(n_alive, n_step, alive, rays_t, o, d, c["bound"], c["dt_gamma"], c["max_steps"], c["cascade"], c["H"], bits,
                      nears, fars, xyzs, dirs, deltas, noises)
        sig, rgb = field(xyzs, dirs)
        be.composite_rays(n_alive, n_step, T_thresh, alive, rays_t, sig.contiguous(), rgb.contiguous(), deltas, ws, dep, img)
        trace.append(dict(n_alive=n_alive, n_step=n_step, xyzs=xyzs.cpu().numpy(), deltas=deltas.cpu().numpy(),
                          alive_after=alive.cpu().numpy().copy()))
        alive = alive[alive >= 0]
        step += n_step; it += 1
    return dict(trace=trace, ws=ws.cpu().numpy(), depth=dep.cpu().numpy(), image=img.cpu().numpy(), rays_t=rays_t.cpu().numpy())


def grid_fwd_bwd(be, c, half=True, calc_grad_inputs=True):
    dt = torch.float16 if half else torch.float32
    x, emb, off = T(c["inputs"]), T(c["embeddings"], dt), T(c["offsets"])
    B, D, C, L = x.shape[0], 3, emb.shape[1], off.shape[0] - 1
    S, H = float(np.log2(c["per_level_scale"])), c["H"]
    out = torch.empty(L, B, C, dtype=dt, device="cuda")
    dy_dx = torch.empty(B, L * D * C, dtype=dt, device="cuda") if calc_grad_inputs else None
    be.grid_encode_forward(x, emb, off, out, B, D, C, L, S, H, dy_dx, 0, False, 0)
    grad = T(c["grad"], dt).view(B, L, C).permute(1, 0, 2).contiguous()
    ge = torch.zeros_like(emb)
    gi = torch.zeros(B, D, dtype=dt, device="cuda") if calc_grad_inputs else None
    be.grid_encode_backward(grad, x, emb, off, ge, B, D, C, L, S, H, dy_dx, gi, 0, False, 0)
    return dict(out=out.cpu().numpy(), dy_dx=None if dy_dx is None else dy_dx.cpu().numpy(), grad_emb=ge.float().cpu().numpy(),
                grad_inputs=None if gi is None else gi.float().cpu().numpy())


def mlp_fwd_bwd(be, c, calc_grad_inputs=True):
    x, w, g = T(c["inputs"]), T(c["weights"]), T(c["grad"])
    B, IN, L = x.shape[0], c["input_dim"], c["num_layers"]
    fb = torch.empty(L, B, 64, dtype=torch.float16, device="cuda")
    out = torch.empty(B, 16, dtype=torch.float16, device="cuda")
    be.ffmlp_forward(x, w, B, IN, 16, 64, L, 0, 6, fb, out)
    out_inf = torch.empty(B, 16, dtype=torch.float16, device="cuda")
    be.ffmlp_inference(x, w, B, IN, 16, 64, L, 0, 6, torch.empty(B, 64, dtype=torch.float16, device="cuda"), out_inf)
    bb = torch.zeros(L, B, 64, dtype=torch.float16, device="cuda")
    gi = torch.zeros(B, IN, dtype=torch.float16, device="cuda") if calc_grad_inputs else torch.zeros(1, dtype=torch.float16, device="cuda")
    gw = torch.zeros_like(w)
    be.ffmlp_backward(g, x, w, fb, B, IN, 16, 64, L, 0, 6, calc_grad_inputs, bb, gi, gw)
    torch.cuda.synchronize()
    return dict(out=out.float().cpu().numpy(), out_inf=out_inf.float().cpu().numpy(), fb=fb.float().cpu().numpy(),
                grad_inputs=gi.float().cpu().numpy(), grad_weights=gw.float().cpu().numpy())


def freq(be, x, deg):
    xi = T(x)
    B, D = xi.shape
    C = D + 2 * D * deg
    out = torch.empty(B, C, device="cuda")
    be.freq_encode_forward(xi, B, D, deg, C, out)
    g = torch.from_numpy(np.random.RandomState(9).normal(0, 1, (B, C)).astype(np.float32)).cuda()
    gi = torch.zeros(B, D, device="cuda")
    be.freq_encode_backward(g, out, B, D, deg, C, gi)
    return out.cpu().numpy(), gi.cpu().numpy(), g.cpu().numpy()
