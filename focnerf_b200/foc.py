"""FOC-NeRF multi-object composition over libfoc_b200 (no reference extension exists for this: COMBINED.py /
editable.py do it in eager PyTorch, materialising [K,N,T] tensors and a running max).

sample_uniform  : COMBINED.py:451-474 `run` sampling + editable.py:443-471 per-object offset.
composite_multi : COMBINED.py:247-251 selection + COMBINED.py:141-200 composite (+ the per-object colour mask of
                  COMBINED.py:515-526 when obj_nears/obj_fars are given).
"""
import torch

from ._lib import check, lib, ptr, require_cuda, stream


def sample_uniform(rays_o, rays_d, aabb, num_steps, min_near=0.2, offset=None, normalize_d=False):
    """-> (rays_o', rays_d', nears [N], fars [N], xyzs [N,T,3]) for one object; offset = (x,y,z) or None."""
    require_cuda(rays_o, rays_d, aabb)
    rays_o = rays_o.to(torch.float32).contiguous().view(-1, 3)
    rays_d = rays_d.to(torch.float32).contiguous().view(-1, 3)
    aabb = aabb.to(torch.float32).contiguous()
    N, dev = rays_o.shape[0], rays_o.device
    off = None if offset is None else torch.as_tensor(offset, dtype=torch.float32, device=dev).contiguous()
    out_o, out_d = torch.empty_like(rays_o), torch.empty_like(rays_d)
    nears, fars = torch.empty(N, dtype=torch.float32, device=dev), torch.empty(N, dtype=torch.float32, device=dev)
    xyzs = torch.empty(N, num_steps, 3, dtype=torch.float32, device=dev)
    check(lib().foc_sample_uniform(ptr(rays_o), ptr(rays_d), ptr(off), int(bool(normalize_d)), ptr(aabb), float(min_near), N,
                                   num_steps, ptr(out_o), ptr(out_d), ptr(nears), ptr(fars), ptr(xyzs), stream()), "sample_uniform")
    return out_o, out_d, nears, fars, xyzs


def composite_multi(sigmas, rgbs, nears, fars, bg=1.0, obj_nears=None, obj_fars=None, mask_thresh=1e-10, T_thresh=0.0):
    """sigmas [K,N,T], rgbs [K,N,T,3] -> image [N,4], depth [N], weights_sum [N]."""
    require_cuda(sigmas, rgbs, nears, fars)
    sigmas, rgbs = sigmas.to(torch.float32).contiguous(), rgbs.to(torch.float32).contiguous()
    nears, fars = nears.to(torch.float32).contiguous().view(-1), fars.to(torch.float32).contiguous().view(-1)
    K, N, T = sigmas.shape
    if obj_nears is not None:
        obj_nears, obj_fars = obj_nears.to(torch.float32).contiguous(), obj_fars.to(torch.float32).contiguous()
    dev = sigmas.device
    image = torch.empty(N, 4, dtype=torch.float32, device=dev)
    depth = torch.empty(N, dtype=torch.float32, device=dev)
    ws = torch.empty(N, dtype=torch.float32, device=dev)
    check(lib().foc_composite_multi(ptr(sigmas), ptr(rgbs), ptr(nears), ptr(fars), ptr(obj_nears), ptr(obj_fars), K, N, T,
                                    float(bg), float(mask_thresh), float(T_thresh), ptr(image), ptr(depth), ptr(ws), stream()),
          "composite_multi")
    return image, depth, ws
