"""TEST / BASELINE INFRASTRUCTURE (not product code): BASELINE.json configs[0], the reference's PyTorch-only renderer path
(`NeRFRenderer.run`, nerf/renderer.py:126-238, non-cuda_ray) restated with torch CPU ops so it can be timed on the GPU box's
host cores.  The reference itself cannot run on a CPU: every operator wrapper `.cuda()`s its inputs (raymarching.py:34-35,
freq.py:22), so the two CUDA operators this path touches are restated in torch:

  near_far       raymarching.cu:92-145   slab test, near clamped to min_near, misses -> (FLT_MAX, FLT_MAX)
  freq_encode    freqencoder.cu:30-94    [x, sin(2^f x), cos(2^f x)]_{f < degree}  (cos as sin(. + pi/2), as the kernel does)
  FieldNetwork   nerf/network.py:33-117  bias-free nn.Linear sigma net (num_layers 2, hidden 64, 1 + 15 outputs), trunc_exp,
                                         colour net (num_layers 3, hidden 64) on [freq(dirs), geo_feat], sigmoid;
                                         density() / color(mask=) as network.py:119-175
  render         nerf/renderer.py:126-238 through tests/foc_eager.py (run_object + image_depth with K = 1)

`get_encoder` is absent from the reference tree; the frequency degrees are vanilla NeRF's (10 for positions, 4
for directions).  Only bench.py's cpu legs and tests import this file.
"""
import math
import os
import sys
import time

import torch
import torch.nn as nn

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import foc_eager  # noqa: E402

FLT_MAX = 3.4028234663852886e38


def near_far(rays_o, rays_d, aabb, min_near=0.2):
    inv = 1.0 / rays_d
    lo = (aabb[:3] - rays_o) * inv
    hi = (aabb[3:] - rays_o) * inv
    tmin, tmax = torch.minimum(lo, hi), torch.maximum(lo, hi)
    near, far = tmin.max(dim=-1).values, tmax.min(dim=-1).values
    miss = near > far
    near = torch.clamp(near, min=min_near)
    near = torch.where(miss, torch.full_like(near, FLT_MAX), near)
    far = torch.where(miss, torch.full_like(far, FLT_MAX), far)
    return near, far


def freq_encode(x, degree):
    f = 2.0 ** torch.arange(degree, dtype=x.dtype, device=x.device)
    xf = x.unsqueeze(-1) * f                                            # [B, D, deg]
    s, c = torch.sin(xf), torch.sin(xf + math.pi / 2)
    # layout of freqencoder.cu:52-57: x, then per frequency (sin over D, cos over D)
    return torch.cat([x, torch.stack([s, c], dim=-1).permute(0, 2, 3, 1).reshape(x.shape[0], -1)], dim=-1)


class FieldNetwork(nn.Module):
    def __init__(self, bound=1.0, degree=10, degree_dir=4, num_layers=2, hidden_dim=64, geo_feat_dim=15, num_layers_color=3, hidden_dim_color=64):
        super().__init__()
        self.bound, self.degree, self.degree_dir = bound, degree, degree_dir
        in_dim, in_dir = 3 + 6 * degree, 3 + 6 * degree_dir
        dims = [in_dim] + [hidden_dim] * (num_layers - 1) + [1 + geo_feat_dim]
        self.sigma_net = nn.ModuleList([nn.Linear(a, b, bias=False) for a, b in zip(dims[:-1], dims[1:])])
        dims = [in_dir + geo_feat_dim] + [hidden_dim_color] * (num_layers_color - 1) + [3]
        self.color_net = nn.ModuleList([nn.Linear(a, b, bias=False) for a, b in zip(dims[:-1], dims[1:])])

    @staticmethod
    def _mlp(layers, h):
        for i, l in enumerate(layers):
            h = l(h)
            if i != len(layers) - 1:
                h = torch.relu(h)
        return h

    def density(self, x):
        h = self._mlp(self.sigma_net, freq_encode(x, self.degree))
        return {"sigma": torch.exp(h[..., 0]), "geo_feat": h[..., 1:]}

    def color(self, x, d, mask=None, geo_feat=None, **kw):
        if mask is not None:
            rgbs = torch.zeros(mask.shape[0], 3, dtype=x.dtype, device=x.device)
            if not mask.any():
                return rgbs
            d, geo_feat = d[mask], geo_feat[mask]
        h = torch.sigmoid(self._mlp(self.color_net, torch.cat([freq_encode(d, self.degree_dir), geo_feat], dim=-1)))
        if mask is not None:
            rgbs[mask] = h.to(rgbs.dtype)
            return rgbs
        return h


@torch.no_grad()
def render(model, rays_o, rays_d, aabb, num_steps=128, bg=1.0):
    """-> image [N,3], depth [N], weights_sum [N]  (the 4th image channel of the FOC composite is dropped)"""
    sig, rgb = foc_eager.run_object(model, near_far, rays_o, rays_d, aabb, num_steps)
    img, depth, ws = foc_eager.image_depth(near_far, rays_o, rays_d, aabb, sig, rgb, num_steps, bg)
    return img[:, :3], depth, ws


def time_cpu(rays_o, rays_d, num_steps=128, runs=20, warmup=3, threads=None):
    """configs[0]: N rays x num_steps samples on the host cores -> dict(rays_per_s, ms, cores, samples)."""
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    model = FieldNetwork().eval()
    aabb = torch.tensor([-1.0] * 3 + [1.0] * 3)
    o, d = torch.as_tensor(rays_o, dtype=torch.float32).cpu(), torch.as_tensor(rays_d, dtype=torch.float32).cpu()
    for _ in range(warmup):
        render(model, o, d, aabb, num_steps)
    t = time.perf_counter()
    for _ in range(runs):
        out = render(model, o, d, aabb, num_steps)
    dt = (time.perf_counter() - t) / runs
    return {"value": o.shape[0] / dt, "unit": "rays/s", "ms_per_batch": dt * 1e3, "cores": threads, "rays": int(o.shape[0]), "samples_per_ray": num_steps,
            "kind": "port", "finite": bool(torch.isfinite(out[0]).all()),
            "sample": f"{runs} renders of {o.shape[0]} rays x {num_steps} uniform samples, fp32 torch on {threads} host threads "
                      "(configs[0]: NeRFRenderer.run with frequency encoders and bias-free nn.Linear nets; the reference's own "
                      "operators are CUDA-only, restated in torch)"}
