"""oracle -- CPU restatement of the FOC-NeRF per-ray hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
package; no product module (focnerf_b200/*) does.  numpy in, numpy out; every function mirrors the
reference operator it restates (see the C sources for file:line citations).

Parity pinning: the reference has no golden vectors and no CPU path.  The oracle is
pinned against tests/golden/*.npz, produced by tests/golden/make_golden.py from the reference's own CUDA
extensions (oracle/_ref, cross-compiled here from /root/reference by oracle/build_ref.py) run on a B200.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liborc.so")
_SRCS = ["orc_raymarch.c", "orc_composite.c", "orc_grid.c", "orc_mlp.c", "Makefile"]


def build(force=False):
    """gcc-compile the C restatement into oracle/_build/liborc.so (rebuilds when a source is newer)."""
    stale = force or not os.path.exists(_LIB_PATH)
    if not stale:
        t = os.path.getmtime(_LIB_PATH)
        stale = any(os.path.getmtime(os.path.join(_HERE, s)) > t for s in _SRCS)
    if stale:
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


u32, f32, ci = ctypes.c_uint32, ctypes.c_float, ctypes.c_int

# ---------------------------------------------------------------- raymarching


def near_far_from_aabb(rays_o, rays_d, aabb, min_near=0.2):
    rays_o, rays_d, aabb = _f32(rays_o).reshape(-1, 3), _f32(rays_d).reshape(-1, 3), _f32(aabb)
    N = rays_o.shape[0]
    nears, fars = np.empty(N, np.float32), np.empty(N, np.float32)
    lib().orc_near_far_from_aabb(_p(rays_o), _p(rays_d), _p(aabb), u32(N), f32(min_near), _p(nears), _p(fars))
    return nears, fars


def sph_from_ray(rays_o, rays_d, radius):
    rays_o, rays_d = _f32(rays_o).reshape(-1, 3), _f32(rays_d).reshape(-1, 3)
    N = rays_o.shape[0]
    coords = np.empty((N, 2), np.float32)
    lib().orc_sph_from_ray(_p(rays_o), _p(rays_d), f32(radius), u32(N), _p(coords))
    return coords


def morton3D(coords):
    coords = _i32(coords)
    out = np.empty(coords.shape[0], np.int32)
    lib().orc_morton3D(_p(coords), u32(coords.shape[0]), _p(out))
    return out


def morton3D_invert(indices):
    indices = _i32(indices)
    out = np.empty((indices.shape[0], 3), np.int32)
    lib().orc_morton3D_invert(_p(indices), u32(indices.shape[0]), _p(out))
    return out


def packbits(grid, thresh):
    grid = _f32(grid)
    N = grid.size // 8
    out = np.empty(N, np.uint8)
    lib().orc_packbits(_p(grid), u32(N), f32(thresh), _p(out))
    return out


def march_rays_train(rays_o, rays_d, bound, bitfield, C, H, nears, fars, noises, M, dt_gamma=0.0, max_steps=1024):
    """Returns xyzs[M,3], dirs[M,3], deltas[M,2] (zero-filled beyond the used slots), rays[N,3], counter[2]."""
    rays_o, rays_d = _f32(rays_o).reshape(-1, 3), _f32(rays_d).reshape(-1, 3)
    nears, fars, noises = _f32(nears), _f32(fars), _f32(noises)
    bitfield = np.ascontiguousarray(bitfield, dtype=np.uint8)
    N = rays_o.shape[0]
    xyzs, dirs, deltas = np.zeros((M, 3), np.float32), np.zeros((M, 3), np.float32), np.zeros((M, 2), np.float32)
    rays, counter = np.zeros((N, 3), np.int32), np.zeros(2, np.int32)
    lib().orc_march_rays_train(_p(rays_o), _p(rays_d), _p(bitfield), f32(bound), f32(dt_gamma), u32(max_steps), u32(N),
                               u32(C), u32(H), u32(M), _p(nears), _p(fars), _p(xyzs), _p(dirs), _p(deltas), _p(rays),
                               _p(counter), _p(noises))
    return xyzs, dirs, deltas, rays, counter


def composite_rays_train_forward(sigmas, rgbs, deltas, rays, T_thresh=1e-4):
    sigmas, rgbs, deltas, rays = _f32(sigmas), _f32(rgbs), _f32(deltas), _i32(rays)
    M, N = sigmas.shape[0], rays.shape[0]
    ws, depth, image = np.zeros(N, np.float32), np.zeros(N, np.float32), np.zeros((N, 3), np.float32)
    lib().orc_composite_rays_train_forward(_p(sigmas), _p(rgbs), _p(deltas), _p(rays), u32(M), u32(N), f32(T_thresh),
                                           _p(ws), _p(depth), _p(image))
    return ws, depth, image


def composite_rays_train_backward(grad_ws, grad_image, sigmas, rgbs, deltas, rays, ws, image, T_thresh=1e-4):
    sigmas, rgbs, deltas, rays = _f32(sigmas), _f32(rgbs), _f32(deltas), _i32(rays)
    grad_ws, grad_image, ws, image = _f32(grad_ws), _f32(grad_image), _f32(ws), _f32(image)
    M, N = sigmas.shape[0], rays.shape[0]
    gs, gc = np.zeros(M, np.float32), np.zeros((M, 3), np.float32)
    lib().orc_composite_rays_train_backward(_p(grad_ws), _p(grad_image), _p(sigmas), _p(rgbs), _p(deltas), _p(rays),
                                            _p(ws), _p(image), u32(M), u32(N), f32(T_thresh), _p(gs), _p(gc))
    return gs, gc


def march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, bitfield, C, H, nears, fars, noises,
               M=None, dt_gamma=0.0, max_steps=1024):
    rays_o, rays_d = _f32(rays_o).reshape(-1, 3), _f32(rays_d).reshape(-1, 3)
    rays_alive, rays_t, nears, fars, noises = _i32(rays_alive), _f32(rays_t), _f32(nears), _f32(fars), _f32(noises)
    bitfield = np.ascontiguousarray(bitfield, dtype=np.uint8)
    M = n_alive * n_step if M is None else M
    xyzs, dirs, deltas = np.zeros((M, 3), np.float32), np.zeros((M, 3), np.float32), np.zeros((M, 2), np.float32)
    lib().orc_march_rays(u32(n_alive), u32(n_step), _p(rays_alive), _p(rays_t), _p(rays_o), _p(rays_d), f32(bound),
                         f32(dt_gamma), u32(max_steps), u32(C), u32(H), _p(bitfield), _p(nears), _p(fars), _p(xyzs),
                         _p(dirs), _p(deltas), _p(noises))
    return xyzs, dirs, deltas


def composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image, T_thresh=1e-2):
    """In place on rays_alive, rays_t, weights_sum, depth, image (must be C-contiguous int32/float32 arrays)."""
    for a, dt in ((rays_alive, np.int32), (rays_t, np.float32), (weights_sum, np.float32), (depth, np.float32), (image, np.float32)):
        assert a.dtype == dt and a.flags.c_contiguous
    sigmas, rgbs, deltas = _f32(sigmas), _f32(rgbs), _f32(deltas)
    lib().orc_composite_rays(u32(n_alive), u32(n_step), f32(T_thresh), _p(rays_alive), _p(rays_t), _p(sigmas), _p(rgbs),
                             _p(deltas), _p(weights_sum), _p(depth), _p(image))


def foc_composite(sigmas, rgbs, nears, fars, bg=1.0):
    """sigmas [K,N,T], rgbs [K,N,T,3] -> image [N,4], depth [N], weights_sum [N] (COMBINED.py:141-200,247-251)."""
    sigmas, rgbs, nears, fars = _f32(sigmas), _f32(rgbs), _f32(nears), _f32(fars)
    K, N, T = sigmas.shape
    img, dep, ws = np.empty((N, 4), np.float32), np.empty(N, np.float32), np.empty(N, np.float32)
    lib().orc_foc_composite(_p(sigmas), _p(rgbs), _p(nears), _p(fars), u32(K), u32(N), u32(T), f32(bg), _p(img), _p(dep), _p(ws))
    return img, dep, ws

# ---------------------------------------------------------------- hash grid


def grid_offsets(input_dim=3, num_levels=16, level_dim=2, per_level_scale=2, base_resolution=16, log2_hashmap_size=19,
                 desired_resolution=None, align_corners=False):
    """Level offsets and per_level_scale exactly as gridencoder/grid.py:100-128 builds them."""
    if desired_resolution is not None:
        per_level_scale = np.exp2(np.log2(desired_resolution / base_resolution) / (num_levels - 1))
    offsets, offset, max_params = [], 0, 2 ** log2_hashmap_size
    for i in range(num_levels):
        resolution = int(np.ceil(base_resolution * per_level_scale ** i))
        n = min(max_params, (resolution if align_corners else resolution + 1) ** input_dim)
        n = int(np.ceil(n / 8) * 8)
        offsets.append(offset)
        offset += n
    offsets.append(offset)
    return np.array(offsets, dtype=np.int32), per_level_scale


def grid_set_level_scales(scales=None):
    """Override the per-level scales with device-computed ones (None restores the host exp2f)."""
    if scales is None:
        lib().orc_grid_set_level_scales(None, u32(0))
    else:
        s = _f32(scales)
        lib().orc_grid_set_level_scales(_p(s), u32(s.shape[0]))


def _tbl(a, half):
    return np.ascontiguousarray(a, dtype=np.float16 if half else np.float32)


def grid_encode_forward(inputs, embeddings, offsets, per_level_scale, base_resolution, calc_grad_inputs=False,
                        gridtype=0, align_corners=False, interp=0, half=True):
    """inputs [B,D] in [0,1]; returns outputs [L,B,C] (the kernel's native order) and dy_dx [B,L*D*C] or None."""
    inputs, offsets = _f32(inputs), _i32(offsets)
    emb = _tbl(embeddings, half)
    B, D = inputs.shape
    L, C = offsets.shape[0] - 1, emb.shape[1]
    S = np.float32(np.log2(per_level_scale))
    out = np.empty((L, B, C), emb.dtype)
    dy_dx = np.empty((B, L * D * C), emb.dtype) if calc_grad_inputs else None
    lib().orc_grid_encode_forward(_p(inputs), _p(emb), _p(offsets), _p(out), u32(B), u32(D), u32(C), u32(L), f32(S),
                                  u32(base_resolution), _p(dy_dx), u32(gridtype), ci(int(align_corners)), u32(interp), ci(int(half)))
    return out, dy_dx


def grid_encode_backward(grad, inputs, offsets, per_level_scale, base_resolution, C, gridtype=0, align_corners=False,
                         interp=0, half=True):
    """grad [L,B,C]; returns the exact (float64) scatter-add grad_embeddings [n_entries, C]."""
    inputs, offsets = _f32(inputs), _i32(offsets)
    grad = _tbl(grad, half)
    B, D = inputs.shape
    L = offsets.shape[0] - 1
    acc = np.zeros((int(offsets[-1]), C), np.float64)
    S = np.float32(np.log2(per_level_scale))
    lib().orc_grid_encode_backward(_p(grad), _p(inputs), _p(offsets), _p(acc), u32(B), u32(D), u32(C), u32(L), f32(S),
                                   u32(base_resolution), u32(gridtype), ci(int(align_corners)), u32(interp), ci(int(half)))
    return acc


def grid_input_backward(grad, dy_dx, B, D, C, L, half=True):
    grad, dy_dx = _tbl(grad, half), _tbl(dy_dx, half)
    out = np.empty((B, D), grad.dtype)
    lib().orc_grid_input_backward(_p(grad), _p(dy_dx), _p(out), u32(B), u32(D), u32(C), u32(L), ci(int(half)))
    return out


def grad_total_variation(inputs, embeddings, offsets, per_level_scale, base_resolution, weight=1e-7, gridtype=0,
                         align_corners=False):
    inputs, emb, offsets = _f32(inputs), _f32(embeddings), _i32(offsets)
    B, D = inputs.shape
    L, C = offsets.shape[0] - 1, emb.shape[1]
    acc = np.zeros(emb.shape, np.float64)
    S = np.float32(np.log2(per_level_scale))
    lib().orc_grad_total_variation(_p(inputs), _p(emb), _p(acc), _p(offsets), f32(weight), u32(B), u32(D), u32(C), u32(L),
                                   f32(S), u32(base_resolution), u32(gridtype), ci(int(align_corners)))
    return acc

# ---------------------------------------------------------------- fused MLP, encoders, activations


def ffmlp_num_params(input_dim, output_dim, hidden_dim, num_layers):
    return hidden_dim * (input_dim + hidden_dim * (num_layers - 1) + 16 * ((output_dim + 15) // 16))


def ffmlp_forward(inputs, weights, input_dim, output_dim, hidden_dim, num_layers, activation=0, output_activation=6,
                  training=True, acc_mode=0):
    """inputs [B,in] f16, weights flat f16 -> outputs [B,output_dim] f16, forward_buffer [num_layers,B,hidden] or None."""
    x, w = _tbl(inputs, True), _tbl(weights, True)
    B = x.shape[0]
    out = np.empty((B, output_dim), np.float16)
    fb = np.empty((num_layers, B, hidden_dim), np.float16) if training else None
    lib().orc_ffmlp_forward(_p(x), _p(w), u32(B), u32(input_dim), u32(output_dim), u32(hidden_dim), u32(num_layers),
                            u32(activation), u32(output_activation), _p(fb), _p(out), ci(acc_mode))
    return out, fb


def ffmlp_backward(grad, inputs, weights, forward_buffer, input_dim, output_dim, hidden_dim, num_layers, activation=0,
                   calc_grad_inputs=False):
    """Returns grad_weights (float64, flat), backward_buffer [num_layers,B,hidden] f16, grad_inputs [B,in] f16 or None."""
    g, x, w, fb = _tbl(grad, True), _tbl(inputs, True), _tbl(weights, True), _tbl(forward_buffer, True)
    B = x.shape[0]
    bb = np.zeros((num_layers, B, hidden_dim), np.float16)
    gi = np.zeros((B, input_dim), np.float16) if calc_grad_inputs else None
    gw = np.zeros(w.shape[0], np.float64)
    lib().orc_ffmlp_backward(_p(g), _p(x), _p(w), _p(fb), u32(B), u32(input_dim), u32(output_dim), u32(hidden_dim),
                             u32(num_layers), u32(activation), _p(bb), _p(gi), _p(gw), ci(0))
    return gw, bb, gi


def freq_encode_forward(inputs, degree):
    x = _f32(inputs)
    B, D = x.shape
    C = D + 2 * D * degree
    out = np.empty((B, C), np.float32)
    lib().orc_freq_encode_forward(_p(x), u32(B), u32(D), u32(degree), u32(C), _p(out))
    return out


def freq_encode_backward(grad, outputs, D, degree):
    g, o = _f32(grad), _f32(outputs)
    B, C = g.shape
    gi = np.empty((B, D), np.float32)
    lib().orc_freq_encode_backward(_p(g), _p(o), u32(B), u32(D), u32(degree), u32(C), _p(gi))
    return gi


def sh4_encode(dirs):
    d = _f32(dirs).reshape(-1, 3)
    out = np.empty((d.shape[0], 16), np.float32)
    lib().orc_sh4_encode(_p(d), u32(d.shape[0]), _p(out))
    return out


def trunc_exp(x):
    """activation.py:5-18 forward."""
    return np.exp(np.asarray(x, np.float32))


def trunc_exp_backward(g, x):
    return np.asarray(g, np.float32) * np.exp(np.clip(np.asarray(x, np.float32), -15, 15))
