"""Device-side batch construction (csrc/rays.cu): rays and ground truth from pixel indices.

`get_rays` keeps the reference's signature and result keys for the part of nerf/utils.py:56-157 the cuda_ray path uses (random pixel
sampling with N > 0, every pixel with N <= 0); `PixelDataset` is the resident form of a preloaded dataset (nerf/provider.py keeps
`poses` / `images` on the GPU with --preload) from which `batch()` builds (rays_o, rays_d, gt) for index lists chosen by the host."""
import torch

from ._lib import check, lib, ptr, require_cuda, stream


def pixel_batch(poses, intrinsics, H, W, pix=None, pose_idx=None, images=None, bg=1.0, out=None):
    """poses [P,4,4] f32 cuda; pix / pose_idx int32 [N] cuda (None: all H*W pixels of pose 0 / pose 0); images [P,H*W,C] f32 or None.
    -> rays_o [N,3], rays_d [N,3], gt [N,3] or None.  `out` = (rays_o, rays_d, gt) buffers to fill (a captured step's inputs)."""
    require_cuda(poses)
    poses = poses.to(torch.float32).contiguous().view(-1, 4, 4)
    N = H * W if pix is None else int(pix.shape[0])
    dev = poses.device
    if out is None:
        out = (torch.empty(N, 3, device=dev), torch.empty(N, 3, device=dev), torch.empty(N, 3, device=dev) if images is not None else None)
    rays_o, rays_d, gt = out
    for t in (pix, pose_idx):
        if t is not None and (t.dtype != torch.int32 or not t.is_contiguous() or t.shape[0] != N):
            raise RuntimeError("pixel_batch: pix / pose_idx must be contiguous int32 tensors of the same length")
    C = 0
    if images is not None:
        if images.dtype != torch.float32 or not images.is_contiguous() or images.shape[0] != poses.shape[0] or images.shape[-1] not in (3, 4):
            raise RuntimeError("pixel_batch: images must be a contiguous float32 [n_poses, H*W, 3|4] tensor")
        C = images.shape[-1]
    fx, fy, cx, cy = (float(v) for v in intrinsics)
    check(lib().foc_pixel_batch(ptr(poses), poses.shape[0], fx, fy, cx, cy, H, W, ptr(pose_idx), ptr(pix), N, ptr(rays_o), ptr(rays_d),
                                ptr(images), C, float(bg), ptr(gt) if images is not None else None, stream()), "pixel_batch")
    return rays_o, rays_d, (gt if images is not None else None)


@torch.no_grad()
def get_rays(poses, intrinsics, H, W, N=-1, error_map=None):
    """legacy/nerf/utils.py get_rays (the cuda_ray path's arguments): poses [B,4,4] -> {'rays_o', 'rays_d' [B,N,3], 'inds' [B,N]}.
    N > 0 draws N random pixels (torch.randint on the device, shared by the B poses, as the reference does); error_map sampling is the
    reference's training-time importance sampler and stays in torch there -- not on this path."""
    if error_map is not None:
        raise NotImplementedError("error_map sampling is outside the hot path (SURVEY.md 8: out of scope)")
    B, dev = poses.shape[0], poses.device
    if N > 0:
        N = min(N, H * W)
        inds = torch.randint(0, H * W, size=[N], device=dev)
    else:
        inds = torch.arange(H * W, device=dev)
        N = H * W
    pix = inds.to(torch.int32).repeat(B)
    pidx = torch.arange(B, device=dev, dtype=torch.int32).repeat_interleave(N)
    o, d, _ = pixel_batch(poses, intrinsics, H, W, pix, pidx)
    out = {"rays_o": o.view(B, N, 3), "rays_d": d.view(B, N, 3)}
    if N < H * W:
        out["inds"] = inds.expand(B, N)
    return out


class PixelDataset:
    """Poses, intrinsics and (optionally) images resident on the GPU; batch(pix, pose_idx) -> (rays_o, rays_d, gt)."""

    def __init__(self, poses, intrinsics, H, W, images=None, bg=1.0):
        self.poses = poses.to(torch.float32).contiguous().view(-1, 4, 4)
        self.intrinsics, self.H, self.W, self.bg = tuple(float(v) for v in intrinsics), int(H), int(W), float(bg)
        self.images = None if images is None else images.to(torch.float32).contiguous().view(self.poses.shape[0], H * W, -1)

    def batch(self, pix, pose_idx=None, out=None):
        return pixel_batch(self.poses, self.intrinsics, self.H, self.W, pix, pose_idx, self.images, self.bg, out)
