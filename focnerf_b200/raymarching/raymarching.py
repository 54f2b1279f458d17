"""Operator surface of the reference's `raymarching` package (raymarching/raymarching.py:19-373) over libfoc_b200.

Same public names, positional signatures, defaults, dtypes, padding rules and in-place conventions, so
nerf/renderer.py / legacy/nerf/renderer.py run unchanged.  Documented deviations:
  * B9  no torch.cuda.empty_cache() inside march_rays_train (raymarching.py:231) -- performance only;
  * B10 kernels run on the current stream and launch errors raise RuntimeError;
  * B11 march_rays_train hands out sample slots by a ray-ordered prefix scan, so `rays` is deterministic
        (the reference's atomics give an arrival-order permutation of the same slots).
"""
import torch
from torch.autograd import Function
from torch.amp import custom_bwd, custom_fwd

from .backend import _backend
from .._lib import valid_rows

_fwd32 = custom_fwd(device_type="cuda", cast_inputs=torch.float32)
_bwd = custom_bwd(device_type="cuda")


def _cuda(t):
    return t if t.is_cuda else t.cuda()


def _pad_up(m, align):
    """The reference always ADDS padding, a full `align` when m is already aligned (raymarching.py:201-202)."""
    return m + (align - m % align) if align > 0 else m


class _near_far_from_aabb(Function):
    @staticmethod
    @_fwd32
    def forward(ctx, rays_o, rays_d, aabb, min_near=0.2):
        rays_o, rays_d = _cuda(rays_o).contiguous().view(-1, 3), _cuda(rays_d).contiguous().view(-1, 3)
        aabb = _cuda(aabb).contiguous()
        N = rays_o.shape[0]
        nears = torch.empty(N, dtype=rays_o.dtype, device=rays_o.device)
        fars = torch.empty(N, dtype=rays_o.dtype, device=rays_o.device)
        _backend.near_far_from_aabb(rays_o, rays_d, aabb, N, min_near, nears, fars)
        return nears, fars


near_far_from_aabb = _near_far_from_aabb.apply


class _sph_from_ray(Function):
    @staticmethod
    @_fwd32
    def forward(ctx, rays_o, rays_d, radius):
        rays_o, rays_d = _cuda(rays_o).contiguous().view(-1, 3), _cuda(rays_d).contiguous().view(-1, 3)
        N = rays_o.shape[0]
        coords = torch.empty(N, 2, dtype=rays_o.dtype, device=rays_o.device)
        _backend.sph_from_ray(rays_o, rays_d, radius, N, coords)
        return coords


sph_from_ray = _sph_from_ray.apply


class _morton3D(Function):
    @staticmethod
    def forward(ctx, coords):
        coords = _cuda(coords)
        N = coords.shape[0]
        indices = torch.empty(N, dtype=torch.int32, device=coords.device)
        _backend.morton3D(coords.int().contiguous(), N, indices)
        return indices


morton3D = _morton3D.apply


class _morton3D_invert(Function):
    @staticmethod
    def forward(ctx, indices):
        indices = _cuda(indices)
        N = indices.shape[0]
        coords = torch.empty(N, 3, dtype=torch.int32, device=indices.device)
        _backend.morton3D_invert(indices.int().contiguous(), N, coords)
        return coords


morton3D_invert = _morton3D_invert.apply


class _packbits(Function):
    @staticmethod
    @_fwd32
    def forward(ctx, grid, thresh, bitfield=None):
        grid = _cuda(grid).contiguous()
        C, H3 = grid.shape[0], grid.shape[1]
        N = C * H3 // 8
        if bitfield is None:
            bitfield = torch.empty(N, dtype=torch.uint8, device=grid.device)
        _backend.packbits(grid, N, thresh, bitfield)          # writes into the caller's buffer when one is passed
        return bitfield


packbits = _packbits.apply


class _march_rays_train(Function):
    @staticmethod
    @_fwd32
    def forward(ctx, rays_o, rays_d, bound, density_bitfield, C, H, nears, fars, step_counter=None, mean_count=-1,
                perturb=False, align=-1, force_all_rays=False, dt_gamma=0, max_steps=1024):
        rays_o, rays_d = _cuda(rays_o).contiguous().view(-1, 3), _cuda(rays_d).contiguous().view(-1, 3)
        density_bitfield = _cuda(density_bitfield).contiguous()
        N = rays_o.shape[0]
        M = N * max_steps
        if not force_all_rays and mean_count > 0:
            M = _pad_up(mean_count, align)
        dev, dt = rays_o.device, rays_o.dtype
        # unused / padding slots must read as zero (the compositors rely on delta == 0, raymarching.cu:858): the reference
        # pre-zeroes the buffers here, our expand kernel zero-fills exactly the rows no ray owns
        xyzs = torch.empty(M, 3, dtype=dt, device=dev)
        dirs = torch.empty(M, 3, dtype=dt, device=dev)
        deltas = torch.empty(M, 2, dtype=dt, device=dev)
        rays = torch.empty(N, 3, dtype=torch.int32, device=dev)
        if step_counter is None:
            step_counter = torch.zeros(2, dtype=torch.int32, device=dev)
        noises = torch.rand(N, dtype=dt, device=dev) if perturb else torch.zeros(N, dtype=dt, device=dev)
        rows = valid_rows()
        if rows is not None and not force_all_rays and mean_count > 0 and M % 128 == 0:
            # graph-replayed step: M is a fixed allocation; the kernel publishes this step's valid row count for the consumers
            _backend.march_rays_train_rows(rays_o, rays_d, density_bitfield, bound, dt_gamma, max_steps, N, C, H, M, nears, fars,
                                           xyzs, dirs, deltas, rays, step_counter, noises, rows)
        else:
            _backend.march_rays_train(rays_o, rays_d, density_bitfield, bound, dt_gamma, max_steps, N, C, H, M, nears, fars,
                                      xyzs, dirs, deltas, rays, step_counter, noises)
        if force_all_rays or mean_count <= 0:
            m = _pad_up(int(step_counter[0].item()), align)      # D2H sync, first epochs only (raymarching.py:222)
            xyzs, dirs, deltas = xyzs[:m], dirs[:m], deltas[:m]
        return xyzs, dirs, deltas, rays


march_rays_train = _march_rays_train.apply


class _composite_rays_train(Function):
    @staticmethod
    @_fwd32
    def forward(ctx, sigmas, rgbs, deltas, rays, T_thresh=1e-4):
        sigmas, rgbs = sigmas.contiguous(), rgbs.contiguous()
        M, N = sigmas.shape[0], rays.shape[0]
        weights_sum = torch.empty(N, dtype=sigmas.dtype, device=sigmas.device)
        depth = torch.empty(N, dtype=sigmas.dtype, device=sigmas.device)
        image = torch.empty(N, 3, dtype=sigmas.dtype, device=sigmas.device)
        _backend.composite_rays_train_forward(sigmas, rgbs, deltas, rays, M, N, T_thresh, weights_sum, depth, image)
        ctx.save_for_backward(sigmas, rgbs, deltas, rays, weights_sum, depth, image)
        ctx.dims = [M, N, T_thresh]
        return weights_sum, depth, image

    @staticmethod
    @_bwd
    def backward(ctx, grad_weights_sum, grad_depth, grad_image):
        # grad_depth is ignored, as in the reference (raymarching.py:275)
        grad_weights_sum, grad_image = grad_weights_sum.contiguous(), grad_image.contiguous()
        sigmas, rgbs, deltas, rays, weights_sum, depth, image = ctx.saved_tensors
        M, N, T_thresh = ctx.dims
        grad_sigmas, grad_rgbs = torch.zeros_like(sigmas), torch.zeros_like(rgbs)
        _backend.composite_rays_train_backward(grad_weights_sum, grad_image, sigmas, rgbs, deltas, rays, weights_sum, image,
                                               M, N, T_thresh, grad_sigmas, grad_rgbs)
        return grad_sigmas, grad_rgbs, None, None, None


composite_rays_train = _composite_rays_train.apply


class _composite_rays_train_mse(Function):
    """composite_rays_train + background mix + MSE against gt, fused (csrc/composite.cu k_composite_mse_*).
    Returns (loss [], image_with_bg [N,3], weights_sum [N], depth [N]); only `loss` carries a gradient.
    `rays` must hold slices in ascending ray order (what march_rays_train of this package produces)."""

    @staticmethod
    @_fwd32
    def forward(ctx, sigmas, rgbs, deltas, rays, gt, bg_color=1.0, T_thresh=1e-4):
        sigmas, rgbs, gt = sigmas.contiguous(), rgbs.contiguous(), gt.contiguous()
        M, N = sigmas.shape[0], rays.shape[0]
        dev = sigmas.device
        weights_sum = torch.empty(N, dtype=torch.float32, device=dev)
        depth = torch.empty(N, dtype=torch.float32, device=dev)
        image = torch.empty(N, 3, dtype=torch.float32, device=dev)
        err = torch.empty(N, dtype=torch.float32, device=dev)
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        _backend.composite_mse_forward(sigmas, rgbs, deltas, rays, gt, float(bg_color), M, N, T_thresh, weights_sum, depth, image, err, loss)
        ctx.save_for_backward(sigmas, rgbs, deltas, rays, gt, weights_sum, image)
        ctx.dims = (M, N, T_thresh, float(bg_color))
        ctx.mark_non_differentiable(weights_sum, depth, image)
        ctx.set_materialize_grads(False)              # no zero tensors for the three non-differentiable outputs in backward
        return loss.view(()), image, weights_sum, depth

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, grad_loss, g_image, g_ws, g_depth):
        sigmas, rgbs, deltas, rays, gt, weights_sum, image = ctx.saved_tensors
        M, N, T_thresh, bg = ctx.dims
        grad_sigmas, grad_rgbs = torch.empty_like(sigmas), torch.empty_like(rgbs)       # fully written by the kernel (ordered = 1)
        rows = valid_rows()
        if rows is not None:                                                             # ... up to the step's valid row count
            _backend.composite_mse_backward_rows(grad_loss.float().contiguous().view(1), gt, bg, sigmas, rgbs, deltas, rays, weights_sum, image,
                                                 M, N, T_thresh, grad_sigmas, grad_rgbs, rows)
        else:
            _backend.composite_mse_backward(grad_loss.float().contiguous().view(1), gt, bg, sigmas, rgbs, deltas, rays, weights_sum, image,
                                            M, N, T_thresh, 1, grad_sigmas, grad_rgbs)
        return grad_sigmas, grad_rgbs, None, None, None, None, None


composite_rays_train_mse = _composite_rays_train_mse.apply


class _composite_rays_train_mse_hc(Function):
    """_composite_rays_train_mse with the colour network's raw output hc [M,16] f16 as the colour operand: rgb = half(sigmoid(hc[:, :3]))
    (nerf/network_ff.py:73) is formed inside the compositor and the backward returns grad_hc [M,16] f16 directly -- the sigmoid
    kernels of the field network and the rgb / grad_rgb [M,3] tensors disappear, the values do not change."""

    @staticmethod
    def forward(ctx, sigmas, hc, deltas, rays, gt, bg_color=1.0, T_thresh=1e-4):
        sigmas, gt, hc = sigmas.float().contiguous(), gt.float().contiguous(), hc.contiguous()
        M, N = sigmas.shape[0], rays.shape[0]
        dev = sigmas.device
        weights_sum = torch.empty(N, dtype=torch.float32, device=dev)
        depth = torch.empty(N, dtype=torch.float32, device=dev)
        image = torch.empty(N, 3, dtype=torch.float32, device=dev)
        err = torch.empty(N, dtype=torch.float32, device=dev)
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        _backend.composite_mse_hc_forward(sigmas, hc, deltas, rays, gt, float(bg_color), M, N, T_thresh, weights_sum, depth, image, err, loss)
        ctx.save_for_backward(sigmas, hc, deltas, rays, gt, weights_sum, image)
        ctx.dims = (M, N, T_thresh, float(bg_color))
        ctx.mark_non_differentiable(weights_sum, depth, image)
        ctx.set_materialize_grads(False)
        return loss.view(()), image, weights_sum, depth

    @staticmethod
    def backward(ctx, grad_loss, g_image, g_ws, g_depth):
        sigmas, hc, deltas, rays, gt, weights_sum, image = ctx.saved_tensors
        M, N, T_thresh, bg = ctx.dims
        grad_sigmas, grad_hc = torch.empty_like(sigmas), torch.empty_like(hc)          # fully written by the kernel
        _backend.composite_mse_hc_backward(grad_loss.float().contiguous().view(1), gt, bg, sigmas, hc, deltas, rays, weights_sum, image,
                                           M, N, T_thresh, grad_sigmas, grad_hc)
        return grad_sigmas, grad_hc, None, None, None, None, None


composite_rays_train_mse_hc = _composite_rays_train_mse_hc.apply


class _march_rays(Function):
    @staticmethod
    @_fwd32
    def forward(ctx, n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, density_bitfield, C, H, near, far,
                align=-1, perturb=False, dt_gamma=0, max_steps=1024):
        rays_o, rays_d = _cuda(rays_o).contiguous().view(-1, 3), _cuda(rays_d).contiguous().view(-1, 3)
        M = _pad_up(n_alive * n_step, align)
        dev, dt = rays_o.device, rays_o.dtype
        xyzs = torch.zeros(M, 3, dtype=dt, device=dev)
        dirs = torch.zeros(M, 3, dtype=dt, device=dev)
        deltas = torch.zeros(M, 2, dtype=dt, device=dev)
        noises = torch.rand(n_alive, dtype=dt, device=dev) if perturb else torch.zeros(n_alive, dtype=dt, device=dev)
        _backend.march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, dt_gamma, max_steps, C, H,
                            density_bitfield, near, far, xyzs, dirs, deltas, noises)
        return xyzs, dirs, deltas


march_rays = _march_rays.apply


class _composite_rays(Function):
    @staticmethod
    @_fwd32
    def forward(ctx, n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image, T_thresh=1e-2):
        # in place on rays_alive, rays_t, weights_sum, depth, image; returns an empty tuple like the reference
        _backend.composite_rays(n_alive, n_step, T_thresh, rays_alive, rays_t, sigmas.contiguous(), rgbs.contiguous(), deltas,
                                weights_sum, depth, image)
        return tuple()


composite_rays = _composite_rays.apply


@torch.no_grad()
def march_rays_all(rays_o, rays_d, bound, density_bitfield, C, H, nears, fars, align=128, dt_gamma=0, max_steps=1024):
    """Not in the reference: every sample of every ray (what march_rays_train(force_all_rays=True, perturb=False) returns),
    but with the sample buffers sized EXACTLY -- walk + scan first, one 4-byte D2H read of the total, then the expansion --
    instead of the reference's N * max_steps worst case (raymarching.py:196), which is 21 GB for an 800x800 frame.
    Returns xyzs [M,3], dirs [M,3], deltas [M,2], rays [N,3] with M the total rounded up to `align`."""
    rays_o, rays_d = _cuda(rays_o).float().contiguous().view(-1, 3), _cuda(rays_d).float().contiguous().view(-1, 3)
    density_bitfield = _cuda(density_bitfield).contiguous()
    N, dev = rays_o.shape[0], rays_o.device
    from .._lib import lib
    scratch = torch.empty(max(1, lib().foc_march_rays_train_scratch_bytes(N, max_steps)), dtype=torch.uint8, device=dev)
    rays = torch.empty(N, 3, dtype=torch.int32, device=dev)
    counter = torch.zeros(2, dtype=torch.int32, device=dev)
    noises = torch.zeros(N, dtype=torch.float32, device=dev)
    _backend.march_rays_count(rays_o, rays_d, density_bitfield, bound, dt_gamma, max_steps, N, C, H, nears, fars, rays, counter,
                              noises, scratch)
    M = _pad_up(max(int(counter[0].item()), 1), align)
    xyzs = torch.empty(M, 3, dtype=torch.float32, device=dev)          # the expand kernel zero-fills the padding rows
    dirs = torch.empty(M, 3, dtype=torch.float32, device=dev)
    deltas = torch.empty(M, 2, dtype=torch.float32, device=dev)
    _backend.march_rays_expand(rays_o, rays_d, bound, dt_gamma, max_steps, N, C, H, M, nears, rays, noises, scratch, xyzs, dirs, deltas)
    return xyzs, dirs, deltas, rays


class WindowedMarch:
    """Not in the reference: the render loop of legacy/nerf/renderer.py:337-366 (march_rays -> field -> composite_rays, <= 8 samples per
    ray and iteration, a boolean-mask compaction and a host sync each time) as a few DEPTH WINDOWS over one recorded march.

    __init__ walks every ray once (foc_march_rays_count: the parameters of all occupied steps stay in scratch).  window(start, width)
    hands out samples [start, start + width) of every ray that is still alive, ray-ordered and exactly sized (one 4-byte host read),
    composite(...) applies K10's recurrence to them, carrying weights_sum / depth / image / rays_t in place and clearing `alive` for
    terminated rays -- so the field is never evaluated behind a ray's termination point by more than one window."""

    def __init__(self, rays_o, rays_d, bound, density_bitfield, C, H, nears, fars, dt_gamma=0, max_steps=1024, align=128):
        self.rays_o, self.rays_d = _cuda(rays_o).float().contiguous().view(-1, 3), _cuda(rays_d).float().contiguous().view(-1, 3)
        self.N, dev = self.rays_o.shape[0], self.rays_o.device
        self.args = (float(bound), float(dt_gamma), int(max_steps), int(C), int(H))
        self.nears, self.align = nears, align
        from .._lib import lib
        N = self.N
        self.scratch = torch.empty(max(1, lib().foc_march_rays_train_scratch_bytes(N, max_steps)), dtype=torch.uint8, device=dev)
        self.rays = torch.empty(N, 3, dtype=torch.int32, device=dev)          # whole-ray (id, offset, count)
        self.rays_w = torch.empty(N, 3, dtype=torch.int32, device=dev)        # current window
        self.wcounts = torch.empty(N, dtype=torch.int32, device=dev)
        self.counter = torch.zeros(2, dtype=torch.int32, device=dev)
        self.noises = torch.zeros(N, dtype=torch.float32, device=dev)
        self.alive = torch.ones(N, dtype=torch.uint8, device=dev)
        self.rays_t = nears.clone()
        self.weights_sum = torch.zeros(N, dtype=torch.float32, device=dev)
        self.depth = torch.zeros(N, dtype=torch.float32, device=dev)
        self.image = torch.zeros(N, 3, dtype=torch.float32, device=dev)
        self.M = 0
        b, g, ms, C, H = self.args
        _backend.march_rays_count(self.rays_o, self.rays_d, _cuda(density_bitfield).contiguous(), b, g, ms, N, C, H, nears, fars, self.rays,
                                  self.counter, self.noises, self.scratch)

    @torch.no_grad()
    def window(self, start, width):
        """-> (xyzs [M,3], dirs [M,3], deltas [M,2]) of the window, or None when no alive ray has samples left."""
        b, g, ms, C, H = self.args
        dev = self.rays_o.device
        self.counter.zero_()
        _backend.march_window_count(self.scratch, self.N, int(start), int(width), self.alive, self.wcounts, self.rays_w, self.counter)
        total = int(self.counter[0].item())
        if total == 0:
            return None
        M = self.M = _pad_up(total, self.align)
        xyzs = torch.empty(M, 3, dtype=torch.float32, device=dev)
        dirs = torch.empty(M, 3, dtype=torch.float32, device=dev)
        deltas = torch.empty(M, 2, dtype=torch.float32, device=dev)
        _backend.march_window_expand(self.rays_o, self.rays_d, b, g, ms, self.N, C, H, M, int(start), self.nears, self.rays_w, self.noises,
                                     self.scratch, xyzs, dirs, deltas)
        return xyzs, dirs, deltas

    @torch.no_grad()
    def window_spec(self, start, width, cap, counter, rows):
        """window() without the host read: the samples go into buffers of `cap` rows (a multiple of 128, the caller's guess), `counter`
        (int32 [2], device) receives the exact total and `rows` (int32 [1], device) the number of valid rows min(cap, total rounded up
        to 128) for consumers that honour a device-side row count.  Rays whose slice would pass `cap` are dropped by the expansion and
        skipped by composite(): the caller compares counter[0] with cap afterwards and renders again when it did not fit."""
        b, g, ms, C, H = self.args
        dev = self.rays_o.device
        counter.zero_()
        _backend.march_window_count(self.scratch, self.N, int(start), int(width), self.alive, self.wcounts, self.rays_w, counter)
        rows.copy_(torch.clamp((counter[:1] + 127) // 128 * 128, min=128, max=int(cap)))
        M = self.M = int(cap)
        xyzs = torch.empty(M, 3, dtype=torch.float32, device=dev)
        dirs = torch.empty(M, 3, dtype=torch.float32, device=dev)
        deltas = torch.empty(M, 2, dtype=torch.float32, device=dev)
        _backend.march_window_expand(self.rays_o, self.rays_d, b, g, ms, self.N, C, H, M, int(start), self.nears, self.rays_w, self.noises,
                                     self.scratch, xyzs, dirs, deltas)
        return xyzs, dirs, deltas

    @torch.no_grad()
    def composite(self, sigmas, rgbs, deltas, T_thresh=1e-4):
        sigmas, rgbs = sigmas.float().contiguous(), rgbs.float().contiguous()
        _backend.composite_window(sigmas, rgbs, deltas, self.rays_w, self.M, self.N, float(T_thresh), self.rays_t, self.weights_sum,
                                  self.depth, self.image, self.alive)


def compact_alive(rays_alive, n_alive=None):
    """Device-side replacement for `rays_alive[rays_alive >= 0]` (legacy/nerf/renderer.py:363).

    Returns (compacted int32 [n], n_out int32 [1] ON THE DEVICE); the caller decides when to read n_out."""
    n = rays_alive.shape[0] if n_alive is None else n_alive
    out = torch.empty(n, dtype=torch.int32, device=rays_alive.device)
    n_out = torch.empty(1, dtype=torch.int32, device=rays_alive.device)
    _backend.compact_alive(rays_alive, n, out, n_out)
    return out, n_out
