"""Synthetic table-top scene shared by the tests and BOTH arms of bench.py: there is no dataset in the
reference tree, so rays, ground-truth pixels and the occupancy grid come from an analytic field.

Plain torch, device-agnostic (runs on CPU for the GPU-less tests); nothing here is on the timed path, and it lives outside
the product package so that the reference arm can use it without importing focnerf_b200.
Ray generation follows the reference's `get_rays` (nerf/utils.py:56-157: pixel centres +0.5, pinhole,
normalised directions rotated by the camera-to-world pose).
"""
import math

import numpy as np
import torch


# host-side Morton helpers, used only to BUILD synthetic occupancy grids (the operator `raymarching.morton3D` is a CUDA kernel)
def _spread3(v):
    v = v.to(torch.int64)
    v = (v * 0x00010001) & 0xFF0000FF
    v = (v * 0x00000101) & 0x0F00F00F
    v = (v * 0x00000011) & 0xC30C30C3
    v = (v * 0x00000005) & 0x49249249
    return v


def morton3d_torch(coords):
    """coords [n,3] ints in [0,1024) -> int64 Morton codes (x lowest bit), as raymarching.cu:56-71."""
    return _spread3(coords[:, 0]) | (_spread3(coords[:, 1]) << 1) | (_spread3(coords[:, 2]) << 2)


class Scene:
    """Union of K solid primitives (spheres / boxes) on a ground slab inside [-bound, bound]^3.

    sigma = `density` inside a primitive, 0 outside; colour = per-primitive albedo * (0.7 + 0.3 * lobe(dir)).
    """

    def __init__(self, n_objects=4, bound=1.0, density=50.0, seed=0):
        g = np.random.RandomState(seed)
        self.bound, self.density, self.K = float(bound), float(density), n_objects
        s = min(1.0, bound)
        ang = np.linspace(0, 2 * np.pi, n_objects, endpoint=False) + 0.3
        rad = 0.45 * s
        self.centers = np.stack([rad * np.cos(ang), np.full(n_objects, -0.05 * s), rad * np.sin(ang)], 1).astype(np.float32)
        self.radii = (g.uniform(0.15, 0.30, n_objects) * s).astype(np.float32)
        self.is_box = (np.arange(n_objects) % 2 == 1)
        self.albedo = g.uniform(0.2, 1.0, (n_objects + 1, 3)).astype(np.float32)   # last = ground slab
        self.slab_y = (-0.40 * s, -0.32 * s)
        self.slab_r = 0.85 * s

    # ---- analytic field -------------------------------------------------------------------------
    def _inside(self, x):
        """x [...,3] -> (inside_any [...], object id [...] (K = slab, -1 none))"""
        dev = x.device
        c = torch.from_numpy(self.centers).to(dev)
        r = torch.from_numpy(self.radii).to(dev)
        d = x.unsqueeze(-2) - c                                   # [...,K,3]
        in_sph = (d * d).sum(-1) < r * r
        in_box = (d.abs() < (0.8 * r).unsqueeze(-1)).all(-1)
        sel = torch.from_numpy(self.is_box).to(dev)
        inside_k = torch.where(sel, in_box, in_sph)               # [...,K]
        slab = (x[..., 1] > self.slab_y[0]) & (x[..., 1] < self.slab_y[1]) & (x[..., 0].abs() < self.slab_r) & (x[..., 2].abs() < self.slab_r)
        any_k = inside_k.any(-1)
        first = torch.argmax(inside_k.to(torch.int8), dim=-1)
        oid = torch.where(any_k, first, torch.where(slab, torch.full_like(first, self.K), torch.full_like(first, -1)))
        return any_k | slab, oid

    def sigma(self, x):
        ins, _ = self._inside(x)
        return ins.to(torch.float32) * self.density

    def sigma_color(self, x, d):
        ins, oid = self._inside(x)
        alb = torch.from_numpy(self.albedo).to(x.device)[oid.clamp(min=0)]
        lobe = 0.5 * (1.0 + d[..., 1:2].expand_as(alb))
        col = alb * (0.7 + 0.3 * lobe)
        return ins.to(torch.float32) * self.density, col * ins.unsqueeze(-1)

    # ---- occupancy grid (Morton order, layout of NeRFRenderer.density_grid: [cascade, 128^3]) ----
    def density_grid(self, cascade, H=128, device="cpu"):
        ar = torch.arange(H, device=device)
        xx, yy, zz = torch.meshgrid(ar, ar, ar, indexing="ij")
        coords = torch.stack([xx.reshape(-1), yy.reshape(-1), zz.reshape(-1)], -1)
        idx = morton3d_torch(coords)
        grid = torch.zeros(cascade, H ** 3, device=device)
        for c in range(cascade):
            b = min(2.0 ** c, self.bound)
            xyz = (2.0 * (coords.float() + 0.5) / H - 1.0) * b
            # a cell is occupied if any of its 8 corners or its centre is inside (conservative)
            s = self.sigma(xyz)
            hs = b / H
            for ox in (-hs, hs):
                for oy in (-hs, hs):
                    for oz in (-hs, hs):
                        s = torch.maximum(s, self.sigma(xyz + torch.tensor([ox, oy, oz], device=device)))
            grid[c, idx] = s
        return grid

    # ---- ground truth by dense quadrature ----------------------------------------------------------
    @torch.no_grad()
    def render_gt(self, rays_o, rays_d, n_quad=1024, bg=1.0, chunk=16384):
        out = []
        for h in range(0, rays_o.shape[0], chunk):
            o, d = rays_o[h:h + chunk].double(), rays_d[h:h + chunk].double()
            b = self.bound
            inv = 1.0 / d
            t0, t1 = (-b - o) * inv, (b - o) * inv
            near = torch.minimum(t0, t1).amax(-1).clamp(min=0.0)
            far = torch.maximum(t0, t1).amin(-1)
            hit = far > near
            far = torch.where(hit, far, near)
            u = (torch.arange(n_quad, device=o.device, dtype=torch.float64) + 0.5) / n_quad
            t = near[:, None] + (far - near)[:, None] * u[None]
            x = o[:, None] + d[:, None] * t[..., None]
            sg, col = self.sigma_color(x.float(), d[:, None].expand_as(x).float())
            dt = ((far - near) / n_quad)[:, None]
            alpha = 1.0 - torch.exp(-sg.double() * dt)
            T = torch.cumprod(torch.cat([torch.ones_like(alpha[:, :1]), 1.0 - alpha], 1), 1)[:, :-1]
            w = alpha * T
            rgb = (w[..., None] * col.double()).sum(1) + (1.0 - w.sum(1, keepdim=True)) * bg
            out.append(rgb.float())
        return torch.cat(out, 0)


def look_at_pose(theta, phi, radius):
    """Camera-to-world 4x4 (OpenGL-style axes as in the reference's orbit poses), looking at the origin."""
    eye = np.array([radius * math.cos(phi) * math.sin(theta), radius * math.sin(phi), radius * math.cos(phi) * math.cos(theta)], np.float64)
    fwd = -eye / np.linalg.norm(eye)
    up = np.array([0.0, 1.0, 0.0])
    right = np.cross(fwd, up); right /= np.linalg.norm(right)
    up2 = np.cross(right, fwd)
    pose = np.eye(4, dtype=np.float32)
    pose[:3, 0], pose[:3, 1], pose[:3, 2], pose[:3, 3] = right, up2, fwd, eye   # +z forward (nerf/utils.py:131)
    return pose


def ring_poses(n, radius=2.2, phi=0.45, seed=0):
    g = np.random.RandomState(seed)
    th = np.linspace(0, 2 * np.pi, n, endpoint=False) + g.uniform(0, 0.1)
    return np.stack([look_at_pose(t, phi + 0.1 * math.sin(3 * t), radius) for t in th], 0)


def intrinsics(H, W, fovy_deg=50.0):
    f = H / (2 * math.tan(math.radians(fovy_deg) / 2))
    return f, f, W / 2, H / 2


def get_rays(pose, intr, H, W, pix_idx=None, device="cpu"):
    """pose [4,4] c2w; returns rays_o, rays_d [n,3] for the given flat pixel indices (all pixels if None).
    Same arithmetic as nerf/utils.py:125-139: zs=1, xs=(i-cx)/fx, ys=(j-cy)/fy with i,j pixel centres (+0.5)."""
    fx, fy, cx, cy = intr
    pose = torch.as_tensor(pose, dtype=torch.float32, device=device)
    if pix_idx is None:
        pix_idx = torch.arange(H * W, device=device)
    i = (pix_idx % W).float() + 0.5
    j = torch.div(pix_idx, W, rounding_mode="floor").float() + 0.5
    dirs = torch.stack([(i - cx) / fx, (j - cy) / fy, torch.ones_like(i)], -1)
    dirs = dirs / dirs.norm(dim=-1, keepdim=True)
    rays_d = dirs @ pose[:3, :3].T
    rays_o = pose[:3, 3].expand_as(rays_d)
    return rays_o.contiguous(), rays_d.contiguous()
