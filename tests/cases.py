"""Seeded input builders shared by the parity tests, tests/golden/make_golden.py and __graft_entry__.smoke().
numpy in / numpy out, CPU only, deterministic: the oracle, the CUDA path and the reference extensions all see the
same bytes."""
import functools
import math
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for _p in (ROOT, HERE):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import synthetic  # noqa: E402


@functools.lru_cache(maxsize=4)
def scene_bitfield(bound=1.0, H=128, n_objects=4):
    """-> (scene, cascade, density_grid [C,H^3] float32 numpy, bitfield uint8 numpy, thresh)."""
    import oracle
    scene = synthetic.Scene(n_objects=n_objects, bound=bound, seed=0)
    cascade = 1 + math.ceil(math.log2(bound))
    grid = scene.density_grid(cascade, H=H).numpy().astype(np.float32)
    thresh = min(float(np.clip(grid, 0, None).mean()), 10.0)
    bits = oracle.packbits(grid.reshape(-1), thresh)
    return scene, cascade, grid, bits, thresh


def camera_rays(n, seed, bound=1.0, H=800, W=800, radius=None):
    """n random pixels over 8 ring poses -> rays_o, rays_d float32 [n,3]."""
    g = np.random.RandomState(seed)
    radius = radius if radius is not None else 2.2 * max(1.0, bound) / (2.0 if bound > 1 else 1.0)
    poses = synthetic.ring_poses(8, radius=radius, seed=seed)
    intr = synthetic.intrinsics(H, W)
    per = [n // 8 + (1 if i < n % 8 else 0) for i in range(8)]
    os_, ds_ = [], []
    for p, m in zip(poses, per):
        if m == 0:
            continue
        pix = torch.from_numpy(g.randint(0, H * W, size=m).astype(np.int64))
        o, d = synthetic.get_rays(p, intr, H, W, pix)
        os_.append(o.numpy()); ds_.append(d.numpy())
    return np.ascontiguousarray(np.concatenate(os_, 0), np.float32), np.ascontiguousarray(np.concatenate(ds_, 0), np.float32)


def adversarial_rays(bound=1.0):
    """Axis-aligned directions (infinite reciprocals), rays missing the box, origins inside the box, grazing rays."""
    o, d = [], []
    for ax in range(3):
        for sgn in (1.0, -1.0):
            for off in (0.0, 0.31, -0.2):
                oo = np.full(3, off, np.float32); oo[ax] = -sgn * 2.5 * bound
                dd = np.zeros(3, np.float32); dd[ax] = sgn
                o.append(oo); d.append(dd)
    o.append(np.array([3 * bound, 3 * bound, 0], np.float32)); d.append(np.array([0, 0, 1], np.float32))          # miss
    o.append(np.array([0, 5 * bound, 0], np.float32)); d.append(np.array([0.6, 0.8, 0], np.float32))              # miss
    o.append(np.array([0.1, 0.0, 0.1], np.float32)); d.append(np.array([0.577, 0.577, 0.578], np.float32))        # inside
    o.append(np.array([0.45 * bound, -0.05, 0.1], np.float32)); d.append(np.array([-0.8, 0.0, 0.6], np.float32))  # inside, occupied
    o.append(np.array([-2 * bound, bound * 0.999, 0], np.float32)); d.append(np.array([1, 0, 1e-4], np.float32))  # grazing a face
    o, d = np.stack(o), np.stack(d)
    d = d / np.linalg.norm(d, axis=-1, keepdims=True)
    return np.ascontiguousarray(o, np.float32), np.ascontiguousarray(d.astype(np.float32))


def march_case(n=512, seed=1, bound=1.0, dt_gamma=0.0, adversarial=True):
    import oracle
    scene, cascade, grid, bits, thresh = scene_bitfield(bound)
    o, d = camera_rays(n, seed, bound)
    if adversarial:
        ao, ad = adversarial_rays(bound)
        o, d = np.concatenate([o, ao]), np.concatenate([d, ad])
    aabb = np.array([-bound] * 3 + [bound] * 3, np.float32)
    nears, fars = oracle.near_far_from_aabb(o, d, aabb, 0.2)
    noises = np.random.RandomState(seed + 100).rand(o.shape[0]).astype(np.float32)
    return dict(rays_o=o, rays_d=d, aabb=aabb, nears=nears, fars=fars, noises=noises, bitfield=bits, cascade=cascade, H=128,
                bound=float(bound), dt_gamma=float(dt_gamma), max_steps=1024)


def composite_case(n=512, seed=2, bound=1.0):
    """Real march output + random sigmas / rgbs / upstream gradients."""
    import oracle
    c = march_case(n, seed, bound)
    N = c["rays_o"].shape[0]
    M = N * 1024
    xyzs, dirs, deltas, rays, counter = oracle.march_rays_train(c["rays_o"], c["rays_d"], c["bound"], c["bitfield"], c["cascade"],
                                                                c["H"], c["nears"], c["fars"], c["noises"], M, c["dt_gamma"], c["max_steps"])
    m = int(counter[0]); m += 128 - m % 128
    g = np.random.RandomState(seed + 7)
    sig = np.exp(g.normal(0, 2, m)).astype(np.float32)
    # a few very dense samples so that early termination (T < 1e-4) triggers mid-ray
    sig[g.rand(m) < 0.05] = 400.0
    rgbs = g.rand(m, 3).astype(np.float32)
    c.update(xyzs=xyzs[:m], dirs=dirs[:m], deltas=deltas[:m], rays=rays, M=m, sigmas=sig, rgbs=rgbs,
             grad_ws=g.normal(0, 1, N).astype(np.float32), grad_image=g.normal(0, 1, (N, 3)).astype(np.float32))
    return c


def grid_case(B=4096, seed=3, trained=True, with_oob=True, desired_resolution=2048):
    import oracle
    g = np.random.RandomState(seed)
    offsets, pls = oracle.grid_offsets(desired_resolution=desired_resolution)
    n = int(offsets[-1])
    emb = (g.normal(0, 0.1, (n, 2)) if trained else g.uniform(-1e-4, 1e-4, (n, 2))).astype(np.float32)
    x = g.rand(B, 3).astype(np.float32)
    if with_oob:
        x[0] = [1.5, 0.5, 0.5]; x[1] = [0.5, -0.01, 0.5]; x[2] = [0.0, 0.0, 0.0]; x[3] = [1.0, 1.0, 1.0]
    grad = g.normal(0, 1e-3, (B, 32)).astype(np.float32)
    return dict(inputs=x, embeddings=emb, offsets=offsets, per_level_scale=pls, H=16, grad=grad)


def mlp_case(B=512, input_dim=32, output_dim=16, num_layers=2, seed=4):
    import oracle
    g = np.random.RandomState(seed)
    n = oracle.ffmlp_num_params(input_dim, output_dim, 64, num_layers)
    std = math.sqrt(3 / 64)
    w = g.uniform(-std, std, n).astype(np.float16)
    x = g.normal(0, 1, (B, input_dim)).astype(np.float16)
    grad = g.normal(0, 1e-2, (B, 16)).astype(np.float16)
    return dict(inputs=x, weights=w, grad=grad, input_dim=input_dim, output_dim=output_dim, num_layers=num_layers)


def multi_case(K=4, N=300, T=64, seed=5):
    g = np.random.RandomState(seed)
    bound = 1.0
    o, d = camera_rays(N, seed, bound, H=480, W=640, radius=1.8)
    sig = np.exp(g.normal(0, 2, (K, N, T))).astype(np.float32) * (g.rand(K, N, T) < 0.3)
    sig[0, :, 3] = sig[1, :, 3]                     # exact ties: the earlier object must win
    rgbs = g.rand(K, N, T, 3).astype(np.float32)
    return dict(rays_o=o, rays_d=d, aabb=np.array([-1, -1, -1, 1, 1, 1], np.float32), sigmas=sig.astype(np.float32), rgbs=rgbs,
                K=K, N=N, T=T, offset=np.array([0.01, 0.01, 0.60], np.float32))
