"""`_backend` for the raymarching operators: same method names and positional arguments as the reference's pybind
module `_raymarching` (raymarching/src/bindings.cpp:5-19, raymarching.h:7-18), implemented by unwrapping tensors
to device pointers and calling libfoc_b200.so on the current CUDA stream.
"""
import torch

from .._lib import check, lib, ptr, require_cuda, stream


def _f32c(t, name):
    if t.dtype != torch.float32 or not t.is_contiguous():
        raise RuntimeError(f"{name} must be a contiguous float32 tensor")


class _Backend:
    @staticmethod
    def near_far_from_aabb(rays_o, rays_d, aabb, N, min_near, nears, fars):
        require_cuda(rays_o, rays_d, aabb, nears, fars)
        check(lib().foc_near_far_from_aabb(ptr(rays_o), ptr(rays_d), ptr(aabb), N, min_near, ptr(nears), ptr(fars), stream()),
              "near_far_from_aabb")

    @staticmethod
    def sph_from_ray(rays_o, rays_d, radius, N, coords):
        require_cuda(rays_o, rays_d, coords)
        check(lib().foc_sph_from_ray(ptr(rays_o), ptr(rays_d), radius, N, ptr(coords), stream()), "sph_from_ray")

    @staticmethod
    def morton3D(coords, N, indices):
        require_cuda(coords, indices)
        check(lib().foc_morton3D(ptr(coords), N, ptr(indices), stream()), "morton3D")

    @staticmethod
    def morton3D_invert(indices, N, coords):
        require_cuda(indices, coords)
        check(lib().foc_morton3D_invert(ptr(indices), N, ptr(coords), stream()), "morton3D_invert")

    @staticmethod
    def packbits(grid, N, density_thresh, bitfield):
        require_cuda(grid, bitfield)
        check(lib().foc_packbits(ptr(grid), N, density_thresh, ptr(bitfield), stream()), "packbits")

    @staticmethod
    def march_rays_train(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, M, nears, fars, xyzs, dirs, deltas,
                         rays, counter, noises):
        require_cuda(rays_o, rays_d, grid, nears, fars, xyzs, dirs, deltas, rays, counter, noises)
        scratch = torch.empty(max(1, lib().foc_march_rays_train_scratch_bytes(N, max_steps)), dtype=torch.uint8, device=rays_o.device)
        check(lib().foc_march_rays_train(ptr(rays_o), ptr(rays_d), ptr(grid), bound, dt_gamma, max_steps, N, C, H, M,
                                         ptr(nears), ptr(fars), ptr(xyzs), ptr(dirs), ptr(deltas), ptr(rays), ptr(counter),
                                         ptr(noises), ptr(scratch), stream()), "march_rays_train")

    @staticmethod
    def march_rays_train_rows(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, M, nears, fars, xyzs, dirs, deltas,
                              rays, counter, noises, rows_out):
        """march_rays_train that also publishes min(M, total rounded up to 128) in rows_out (int32 [1], device) and zero-fills only up to it"""
        require_cuda(rays_o, rays_d, grid, nears, fars, xyzs, dirs, deltas, rays, counter, noises, rows_out)
        scratch = torch.empty(max(1, lib().foc_march_rays_train_scratch_bytes(N, max_steps)), dtype=torch.uint8, device=rays_o.device)
        check(lib().foc_march_rays_train_rows(ptr(rays_o), ptr(rays_d), ptr(grid), bound, dt_gamma, max_steps, N, C, H, M,
                                              ptr(nears), ptr(fars), ptr(xyzs), ptr(dirs), ptr(deltas), ptr(rays), ptr(counter),
                                              ptr(noises), ptr(rows_out), ptr(scratch), stream()), "march_rays_train_rows")

    @staticmethod
    def composite_rays_train_forward(sigmas, rgbs, deltas, rays, M, N, T_thresh, weights_sum, depth, image):
        require_cuda(sigmas, rgbs, deltas, rays, weights_sum, depth, image)
        check(lib().foc_composite_rays_train_forward(ptr(sigmas), ptr(rgbs), ptr(deltas), ptr(rays), M, N, T_thresh,
                                                     ptr(weights_sum), ptr(depth), ptr(image), stream()),
              "composite_rays_train_forward")

    @staticmethod
    def composite_rays_train_backward(grad_weights_sum, grad_image, sigmas, rgbs, deltas, rays, weights_sum, image, M, N,
                                      T_thresh, grad_sigmas, grad_rgbs):
        require_cuda(grad_weights_sum, grad_image, sigmas, rgbs, deltas, rays, weights_sum, image, grad_sigmas, grad_rgbs)
        check(lib().foc_composite_rays_train_backward(ptr(grad_weights_sum), ptr(grad_image), ptr(sigmas), ptr(rgbs),
                                                      ptr(deltas), ptr(rays), ptr(weights_sum), ptr(image), M, N, T_thresh,
                                                      ptr(grad_sigmas), ptr(grad_rgbs), stream()),
              "composite_rays_train_backward")

    @staticmethod
    def march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, dt_gamma, max_steps, C, H, grid, nears, fars,
                   xyzs, dirs, deltas, noises):
        require_cuda(rays_alive, rays_t, rays_o, rays_d, grid, nears, fars, xyzs, dirs, deltas, noises)
        check(lib().foc_march_rays(n_alive, n_step, ptr(rays_alive), ptr(rays_t), ptr(rays_o), ptr(rays_d), bound, dt_gamma,
                                   max_steps, C, H, ptr(grid), ptr(nears), ptr(fars), ptr(xyzs), ptr(dirs), ptr(deltas),
                                   ptr(noises), stream()), "march_rays")

    @staticmethod
    def composite_rays(n_alive, n_step, T_thresh, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image):
        require_cuda(rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image)
        check(lib().foc_composite_rays(n_alive, n_step, T_thresh, ptr(rays_alive), ptr(rays_t), ptr(sigmas), ptr(rgbs),
                                       ptr(deltas), ptr(weights_sum), ptr(depth), ptr(image), stream()), "composite_rays")

    # ---- new entry points (no reference counterpart) ---------------------------------------------------------------
    @staticmethod
    def march_rays_count(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, nears, fars, rays, counter, noises, scratch):
        require_cuda(rays_o, rays_d, grid, nears, fars, rays, counter, noises, scratch)
        check(lib().foc_march_rays_count(ptr(rays_o), ptr(rays_d), ptr(grid), bound, dt_gamma, max_steps, N, C, H, ptr(nears), ptr(fars),
                                         ptr(rays), ptr(counter), ptr(noises), ptr(scratch), stream()), "march_rays_count")

    @staticmethod
    def march_rays_expand(rays_o, rays_d, bound, dt_gamma, max_steps, N, C, H, M, nears, rays, noises, scratch, xyzs, dirs, deltas):
        require_cuda(rays_o, rays_d, nears, rays, noises, scratch, xyzs, dirs, deltas)
        check(lib().foc_march_rays_expand(ptr(rays_o), ptr(rays_d), bound, dt_gamma, max_steps, N, C, H, M, ptr(nears), ptr(rays),
                                          ptr(noises), ptr(scratch), ptr(xyzs), ptr(dirs), ptr(deltas), stream()), "march_rays_expand")

    @staticmethod
    def march_window_count(scratch, N, start, width, alive, wcounts, rays_w, counter):
        require_cuda(scratch, wcounts, rays_w, counter)
        check(lib().foc_march_window_count(ptr(scratch), N, start, width, ptr(alive), ptr(wcounts), ptr(rays_w), ptr(counter), stream()),
              "march_window_count")

    @staticmethod
    def march_window_expand(rays_o, rays_d, bound, dt_gamma, max_steps, N, C, H, M, start, nears, rays_w, noises, scratch, xyzs, dirs, deltas):
        require_cuda(rays_o, rays_d, nears, rays_w, noises, scratch, xyzs, dirs, deltas)
        check(lib().foc_march_window_expand(ptr(rays_o), ptr(rays_d), bound, dt_gamma, max_steps, N, C, H, M, start, ptr(nears), ptr(rays_w),
                                            ptr(noises), ptr(scratch), ptr(xyzs), ptr(dirs), ptr(deltas), stream()), "march_window_expand")

    @staticmethod
    def composite_window(sigmas, rgbs, deltas, rays_w, M, N, T_thresh, rays_t, weights_sum, depth, image, alive):
        require_cuda(sigmas, rgbs, deltas, rays_w, rays_t, weights_sum, depth, image, alive)
        for t, nm in ((sigmas, "sigmas"), (rgbs, "rgbs"), (deltas, "deltas")):
            _f32c(t, nm)
        check(lib().foc_composite_window(ptr(sigmas), ptr(rgbs), ptr(deltas), ptr(rays_w), M, N, T_thresh, ptr(rays_t), ptr(weights_sum),
                                         ptr(depth), ptr(image), ptr(alive), stream()), "composite_window")

    @staticmethod
    def composite_mse_forward(sigmas, rgbs, deltas, rays, gt, bg, M, N, T_thresh, weights_sum, depth, image, err, loss):
        require_cuda(sigmas, rgbs, deltas, rays, gt, weights_sum, depth, image, err, loss)
        for t, nm in ((sigmas, "sigmas"), (rgbs, "rgbs"), (deltas, "deltas"), (gt, "gt")):
            _f32c(t, nm)
        check(lib().foc_composite_mse_forward(ptr(sigmas), ptr(rgbs), ptr(deltas), ptr(rays), ptr(gt), bg, M, N, T_thresh, ptr(weights_sum),
                                              ptr(depth), ptr(image), ptr(err), ptr(loss), stream()), "composite_mse_forward")

    @staticmethod
    def composite_mse_backward(grad_loss, gt, bg, sigmas, rgbs, deltas, rays, weights_sum, image, M, N, T_thresh, ordered, grad_sigmas,
                               grad_rgbs):
        require_cuda(grad_loss, gt, sigmas, rgbs, deltas, rays, weights_sum, image, grad_sigmas, grad_rgbs)
        check(lib().foc_composite_mse_backward(ptr(grad_loss), ptr(gt), bg, ptr(sigmas), ptr(rgbs), ptr(deltas), ptr(rays), ptr(weights_sum),
                                               ptr(image), M, N, T_thresh, int(ordered), ptr(grad_sigmas), ptr(grad_rgbs), stream()),
              "composite_mse_backward")

    @staticmethod
    def composite_mse_backward_rows(grad_loss, gt, bg, sigmas, rgbs, deltas, rays, weights_sum, image, M, N, T_thresh, grad_sigmas, grad_rgbs, n_rows):
        require_cuda(grad_loss, gt, sigmas, rgbs, deltas, rays, weights_sum, image, grad_sigmas, grad_rgbs, n_rows)
        check(lib().foc_composite_mse_backward_rows(ptr(grad_loss), ptr(gt), bg, ptr(sigmas), ptr(rgbs), ptr(deltas), ptr(rays), ptr(weights_sum),
                                                    ptr(image), M, N, T_thresh, ptr(grad_sigmas), ptr(grad_rgbs), ptr(n_rows), stream()),
              "composite_mse_backward_rows")

    @staticmethod
    def composite_mse_hc_forward(sigmas, hc, deltas, rays, gt, bg, M, N, T_thresh, weights_sum, depth, image, err, loss):
        require_cuda(sigmas, hc, deltas, rays, gt, weights_sum, depth, image, err, loss)
        for t, nm in ((sigmas, "sigmas"), (deltas, "deltas"), (gt, "gt")):
            _f32c(t, nm)
        if hc.dtype != torch.float16 or not hc.is_contiguous() or hc.shape[-1] != 16:
            raise RuntimeError("hc must be a contiguous CUDA half tensor [M, 16]")
        check(lib().foc_composite_mse_hc_forward(ptr(sigmas), ptr(hc), ptr(deltas), ptr(rays), ptr(gt), bg, M, N, T_thresh, ptr(weights_sum),
                                                 ptr(depth), ptr(image), ptr(err), ptr(loss), stream()), "composite_mse_hc_forward")

    @staticmethod
    def composite_mse_hc_backward(grad_loss, gt, bg, sigmas, hc, deltas, rays, weights_sum, image, M, N, T_thresh, grad_sigmas, grad_hc):
        require_cuda(grad_loss, gt, sigmas, hc, deltas, rays, weights_sum, image, grad_sigmas, grad_hc)
        check(lib().foc_composite_mse_hc_backward(ptr(grad_loss), ptr(gt), bg, ptr(sigmas), ptr(hc), ptr(deltas), ptr(rays), ptr(weights_sum),
                                                  ptr(image), M, N, T_thresh, ptr(grad_sigmas), ptr(grad_hc), stream()),
              "composite_mse_hc_backward")

    @staticmethod
    def compact_alive(rays_alive, n, out, n_out):
        require_cuda(rays_alive, out, n_out)
        scratch = torch.empty(max(4, lib().foc_compact_alive_scratch_bytes(n)), dtype=torch.uint8, device=rays_alive.device)
        check(lib().foc_compact_alive(ptr(rays_alive), n, ptr(out), ptr(n_out), ptr(scratch), stream()), "compact_alive")


_backend = _Backend()
