"""Eager-PyTorch restatement of the reference's multi-object pipeline, op for op (test infrastructure + the timed
reference leg of configs[3]; never imported by the product):

  run_object        COMBINED.py:451-534  MultiTrainer.run (upsample_steps = 0, perturb = False): linspace z_vals, clipped xyzs,
                    model.density on every sample, cumprod weights, model.color on the `weights > 1e-10` subset
  modify_rays       editable.py:443-471  per-object origin offset + re-normalised direction
  select            COMBINED.py:247-251  best_densities_and_colors_v3 (running max, strict '>')
  image_depth       COMBINED.py:141-200  image_depth_generation: cumprod composite, 4th channel sum w*sigma, white/black
                    background, clamp

`model` needs .density(x) -> {'sigma', 'geo_feat'} and .color(x, d, mask=, geo_feat=) like nerf/network_ff.py:76-123;
`near_far(rays_o, rays_d, aabb, min_near)` is the raymarching operator of whichever stack is under test.
"""
import torch


def modify_rays(rays_o, rays_d, offset):
    if offset is None:
        return rays_o, rays_d
    off = torch.as_tensor(offset, dtype=rays_o.dtype, device=rays_o.device)
    return rays_o + off, torch.nn.functional.normalize(rays_d, dim=-1)


def z_values(nears, fars, T):
    z = torch.linspace(0.0, 1.0, T, device=nears.device).unsqueeze(0).expand(nears.shape[0], T)
    z = nears + (fars - nears) * z
    sample_dist = (fars - nears) / T
    deltas = torch.cat([z[..., 1:] - z[..., :-1], sample_dist * torch.ones_like(z[..., :1])], dim=-1)
    return z, deltas


def cumprod_weights(deltas, sigma):
    alphas = 1 - torch.exp(-deltas * sigma)
    shifted = torch.cat([torch.ones_like(alphas[..., :1]), 1 - alphas + 1e-15], dim=-1)
    return alphas * torch.cumprod(shifted, dim=-1)[..., :-1]


def run_object(model, near_far, rays_o, rays_d, aabb, T, min_near=0.2):
    """-> densities [N,T], rgbs [N,T,3]"""
    N = rays_o.shape[0]
    nears, fars = near_far(rays_o, rays_d, aabb, min_near)
    nears, fars = nears.unsqueeze(-1), fars.unsqueeze(-1)
    z, deltas = z_values(nears, fars, T)
    xyzs = rays_o.unsqueeze(-2) + rays_d.unsqueeze(-2) * z.unsqueeze(-1)
    xyzs = torch.min(torch.max(xyzs, aabb[:3]), aabb[3:])
    dens = model.density(xyzs.reshape(-1, 3))
    sigma = dens['sigma'].view(N, T)
    weights = cumprod_weights(deltas, sigma)
    dirs = rays_d.view(-1, 1, 3).expand_as(xyzs)
    mask = weights > 1e-10
    rgbs = model.color(xyzs.reshape(-1, 3), dirs.reshape(-1, 3), mask=mask.reshape(-1), geo_feat=dens['geo_feat'].view(N * T, -1))
    return sigma.float(), rgbs.view(N, T, 3).float()


def select(densities, max_densities, rgbs, best_rgbs):
    new_max = torch.maximum(densities, max_densities)
    new_best = torch.where(densities[..., None] > max_densities[..., None], rgbs, best_rgbs)
    return new_max, new_best


def image_depth(near_far, rays_o, rays_d, aabb, densities, rgbs, T, bg=1.0, min_near=0.2):
    """densities [N,T], rgbs [N,T,3] -> image [N,4], depth [N], weights_sum [N]"""
    nears, fars = near_far(rays_o, rays_d, aabb, min_near)
    nears, fars = nears.unsqueeze(-1), fars.unsqueeze(-1)
    z, deltas = z_values(nears, fars, T)
    weights = cumprod_weights(deltas, densities)
    ws = weights.sum(dim=-1)
    ori_z = ((z - nears) / (fars - nears)).clamp(0, 1)
    depth = torch.sum(weights * ori_z, dim=-1)
    image = torch.sum(weights.unsqueeze(-1) * rgbs, dim=-2)
    alpha = torch.sum(weights.unsqueeze(-1) * densities.unsqueeze(-1), dim=-2)
    image = torch.cat((image, alpha), dim=-1)
    image = image + (1 - ws).unsqueeze(-1) * bg
    return image.clamp(0, 1), depth, ws


@torch.no_grad()
def render_chunk(models, near_far, rays_o, rays_d, aabb, T, offsets=None, bg=1.0, min_near=0.2, autocast=True):
    """The reference's per-image loop body for one ray batch: K x run_object, running arg-max, composite."""
    offsets = offsets or [None] * len(models)
    best_s = best_c = None
    with torch.autocast("cuda", dtype=torch.float16, enabled=autocast):
        for m, off in zip(models, offsets):
            o, d = modify_rays(rays_o, rays_d, off)
            s, c = run_object(m, near_far, o, d, aabb, T, min_near)
            if best_s is None:
                best_s, best_c = s, c
            else:
                best_s, best_c = select(s, best_s, c, best_c)
    return image_depth(near_far, rays_o, rays_d, aabb, best_s, best_c, T, bg, min_near)
