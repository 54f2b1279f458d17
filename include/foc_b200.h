/*
 * foc_b200.h -- C ABI of libfoc_b200.so: the B200-native (sm_100a) implementation of FOC-NeRF's
 * per-ray rendering hot path.  Plain pointers and sizes only; no torch / ATen / pybind types.
 *
 * Every entry point replaces one function of the reference's pybind11 modules (paths relative to the
 * reference checkout).  Conventions that differ from the reference, on purpose:
 *   - all pointers are DEVICE pointers owned by the caller; nothing is allocated or cached inside;
 *   - `stream` is a cudaStream_t passed as void* (the reference launches on the legacy default stream);
 *   - the return value is a cudaError_t as int (0 = success; the reference never checks launches);
 *     FOC_ERR_UNSUPPORTED (-1) is returned for shapes the kernels do not implement.
 *   - entry points are re-entrant and thread-safe (no globals).
 * `half` buffers are IEEE binary16 stored as uint16_t.
 */
#ifndef FOC_B200_H
#define FOC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FOC_ERR_UNSUPPORTED (-1)
#define FOC_ERR_TOO_LARGE (-2)   /* a size beyond the stated maximum of an entry point with built-in (static) scratch */
#define FOC_ABI_VERSION 1

int foc_abi_version(void);

/* Measurement aid, no reference counterpart: n CUDA events; foc_timer_mark stamps slot `slot` on `stream` -- as an EXTERNAL
 * event-record node when the stream is being captured, so that every replay of the captured graph stamps it again;
 * foc_timer_elapsed = milliseconds between two stamped slots (after a synchronize).  bench.py brackets every library call of the
 * graph-replayed train step with these to report per-entry-point device times. */
int foc_timer_create(int n, void **handle);
int foc_timer_mark(void *handle, int slot, void *stream);
int foc_timer_elapsed(void *handle, int a, int b, float *ms);
int foc_timer_destroy(void *handle);

/* ---------------------------------------------------------------- raymarching (raymarching/src/raymarching.h:7-18) */

/* replaces near_far_from_aabb, raymarching.h:7 (kernel raymarching.cu:92-145). rays_* [N,3], aabb [6], nears/fars [N] */
int foc_near_far_from_aabb(const float *rays_o, const float *rays_d, const float *aabb, uint32_t N, float min_near,
                           float *nears, float *fars, void *stream);
/* replaces sph_from_ray, raymarching.h:8 (raymarching.cu:163-198). coords [N,2] */
int foc_sph_from_ray(const float *rays_o, const float *rays_d, float radius, uint32_t N, float *coords, void *stream);
/* replaces morton3D / morton3D_invert, raymarching.h:9-10 (raymarching.cu:214-254). coords [N,3] i32, indices [N] i32 */
int foc_morton3D(const int32_t *coords, uint32_t N, int32_t *indices, void *stream);
int foc_morton3D_invert(const int32_t *indices, uint32_t N, int32_t *coords, void *stream);
/* replaces packbits, raymarching.h:11 (raymarching.cu:268-289). grid [8N] f32 -> bitfield [N] u8 */
int foc_packbits(const float *grid, uint32_t N, float density_thresh, uint8_t *bitfield, void *stream);

/* replaces march_rays_train, raymarching.h:13 (raymarching.cu:312-480).
 * Same outputs; slots are handed out by an exclusive prefix scan in ray order instead of arrival-order
 * atomics, so rays[n] = (n, offset, count) is deterministic.  counter[0] += total samples, counter[1] += N.
 * scratch: foc_march_rays_train_scratch_bytes(N, max_steps) bytes of device memory (contents irrelevant): per-ray counts plus
 * the recorded sample parameters of the single serial pass, which a second, data-parallel kernel expands. */
size_t foc_march_rays_train_scratch_bytes(uint32_t N, uint32_t max_steps);
int foc_march_rays_train(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound, float dt_gamma,
                         uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, uint32_t M, const float *nears,
                         const float *fars, float *xyzs, float *dirs, float *deltas, int32_t *rays, int32_t *counter,
                         const float *noises, void *scratch, void *stream);
/* The two halves of foc_march_rays_train, for callers that size the sample buffers exactly (no reference counterpart):
 * count  = walk + exclusive scan: rays [N,3] = (ray, offset, count), counter[0] += total samples, counter[1] += N;
 * expand = write xyzs/dirs/deltas [M,*] from the parameters recorded in `scratch` by the count call (same N, max_steps). */
int foc_march_rays_count(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound, float dt_gamma,
                         uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, const float *nears, const float *fars,
                         int32_t *rays, int32_t *counter, const float *noises, void *scratch, void *stream);
int foc_march_rays_expand(const float *rays_o, const float *rays_d, float bound, float dt_gamma, uint32_t max_steps, uint32_t N,
                          uint32_t C, uint32_t H, uint32_t M, const float *nears, const int32_t *rays, const float *noises,
                          const void *scratch, float *xyzs, float *dirs, float *deltas, void *stream);

/* Depth-windowed render (no reference counterpart: replaces the host-driven march_rays / composite_rays loop of
 * legacy/nerf/renderer.py:337-366 by a few windows over the samples recorded by foc_march_rays_count, same scratch):
 * window_count    : rays_w[n] = (n, offset, min(count[n] - start, width)) for rays with alive[n] != 0 (alive == NULL: every ray), count 0
 *                   otherwise; counter[0] += total, counter[1] += N.  wcounts [N] u32 is caller-owned scratch.
 * window_expand   : foc_march_rays_expand for the records [start, start + count_w) of every ray.
 * composite_window: K10's recurrence (composite_rays, raymarching.h:18, raymarching.cu:819-905) over each ray's window slice, state
 *                   carried in place in rays_t / weights_sum / depth / image [N,*]; alive[n] = 0 once the ray terminated. */
int foc_march_window_count(void *scratch, uint32_t N, uint32_t start, uint32_t width, const uint8_t *alive, uint32_t *wcounts,
                           int32_t *rays_w, int32_t *counter, void *stream);
int foc_march_window_expand(const float *rays_o, const float *rays_d, float bound, float dt_gamma, uint32_t max_steps, uint32_t N,
                            uint32_t C, uint32_t H, uint32_t M, uint32_t start, const float *nears, const int32_t *rays_w,
                            const float *noises, const void *scratch, float *xyzs, float *dirs, float *deltas, void *stream);
int foc_composite_window(const float *sigmas, const float *rgbs, const float *deltas, const int32_t *rays_w, uint32_t M, uint32_t N,
                         float T_thresh, float *rays_t, float *weights_sum, float *depth, float *image, uint8_t *alive, void *stream);

/* replaces composite_rays_train_forward / _backward, raymarching.h:14-15 (raymarching.cu:501-577, 602-682) */
int foc_composite_rays_train_forward(const float *sigmas, const float *rgbs, const float *deltas, const int32_t *rays,
                                     uint32_t M, uint32_t N, float T_thresh, float *weights_sum, float *depth,
                                     float *image, void *stream);
int foc_composite_rays_train_backward(const float *grad_weights_sum, const float *grad_image, const float *sigmas,
                                      const float *rgbs, const float *deltas, const int32_t *rays,
                                      const float *weights_sum, const float *image, uint32_t M, uint32_t N,
                                      float T_thresh, float *grad_sigmas, float *grad_rgbs, void *stream);

/* replaces march_rays / composite_rays, raymarching.h:17-18 (raymarching.cu:701-805, 819-905) */
int foc_march_rays(uint32_t n_alive, uint32_t n_step, const int32_t *rays_alive, const float *rays_t,
                   const float *rays_o, const float *rays_d, float bound, float dt_gamma, uint32_t max_steps,
                   uint32_t C, uint32_t H, const uint8_t *grid, const float *nears, const float *fars, float *xyzs,
                   float *dirs, float *deltas, const float *noises, void *stream);
int foc_composite_rays(uint32_t n_alive, uint32_t n_step, float T_thresh, int32_t *rays_alive, float *rays_t,
                       const float *sigmas, const float *rgbs, const float *deltas, float *weights_sum, float *depth,
                       float *image, void *stream);

/* New (no reference counterpart; replaces `rays_alive = rays_alive[rays_alive >= 0]`,
 * legacy/nerf/renderer.py:363): stable compaction of the non-negative ids; *n_out (device) receives the count. */
size_t foc_compact_alive_scratch_bytes(uint32_t n);
int foc_compact_alive(const int32_t *rays_alive, uint32_t n, int32_t *out, int32_t *n_out, void *scratch, void *stream);

/* FOC-NeRF per-object uniform sampling (COMBINED.py:451-474 `run`; editable.py:443-471 `modify_rays_for_object`):
 * o' = rays_o + offset (offset [3] or NULL), d' = normalize(rays_d) when normalize_d, near/far of (o', d') against aabb
 * (raymarching.cu:92-145), then xyzs[n,i] = clip(o' + d' * (near + (far-near) * linspace(0,1,T)[i]), aabb).
 * out_o/out_d [N,3], nears/fars [N], xyzs [N,T,3] (NULL or T == 0: rays only). */
int foc_sample_uniform(const float *rays_o, const float *rays_d, const float *offset, int normalize_d, const float *aabb,
                       float min_near, uint32_t N, uint32_t T, float *out_o, float *out_d, float *nears, float *fars,
                       float *xyzs, void *stream);

/* FOC-NeRF multi-object composite (COMBINED.py:141-200, 247-251, 515-526), one fused kernel:
 * sigmas [K,N,T], rgbs [K,N,T,3] f32; per-sample argmax-sigma selection (earlier object wins ties), then a
 * front-to-back composite over T uniform samples in [near,far] of the UNSHIFTED rays.  obj_nears/obj_fars [K,N]
 * (both or neither): when given, object k's colour counts only where its OWN weight along its own (possibly shifted)
 * ray exceeds mask_thresh (1e-10 in the reference); NULL = rgbs already masked.  T_thresh = 0 is the reference's
 * behaviour (no early termination); > 0 stops a ray once its transmittance falls below it.
 * image [N,4] (rgb + sum w*sigma, + (1-sum w)*bg, clamped to [0,1]), depth [N], weights_sum [N]. */
int foc_composite_multi(const float *sigmas, const float *rgbs, const float *nears, const float *fars,
                        const float *obj_nears, const float *obj_fars, uint32_t K, uint32_t N, uint32_t T, float bg,
                        float mask_thresh, float T_thresh, float *image, float *depth, float *weights_sum, void *stream);

/* ---------------------------------------------------------------- gridencoder (gridencoder/src/gridencoder.h:12-15) */

/* dtype codes for table-typed buffers */
#define FOC_F32 0
#define FOC_F16 1

/* output / incoming-gradient layouts of the grid encoder: FOC_LBC is the reference kernel's [L,B,C];
 * FOC_BLC is [B,L,C] = the [B, L*C] tensor the network consumes (saves the wrapper's permute+reshape copy, grid.py:57) */
#define FOC_LBC 0
#define FOC_BLC 1

/* replaces grid_encode_forward, gridencoder.h:12 (gridencoder.cu:87-245). inputs [B,D] f32 in [0,1];
 * embeddings [n,C]; offsets [L+1] i32; outputs [L,B,C] or [B,L,C] (layout); dy_dx [B,L*D*C] or NULL.
 * D in {2,3}, C in {1,2,4,8}. */
int foc_grid_encode_forward(const float *inputs, const void *embeddings, const int32_t *offsets, void *outputs,
                            uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, void *dy_dx,
                            uint32_t gridtype, int align_corners, uint32_t interp, int dtype, int layout, void *stream);
/* New (diagnostic): scales[l] = exp2f(l * S) * H - 1 as the device evaluates it inside the encoder kernels. */
int foc_grid_level_scales(float S, uint32_t H, uint32_t L, float *scales, void *stream);
/* replaces grid_encode_backward, gridencoder.h:13 (gridencoder.cu:248-369). grad [L,B,C]; grad_embeddings [n,C]
 * (accumulated into; caller zeroes); dy_dx/grad_inputs NULL or [B,L*D*C]/[B,D] (table dtype). */
int foc_grid_encode_backward(const void *grad, const float *inputs, const void *embeddings, const int32_t *offsets,
                             void *grad_embeddings, uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S,
                             uint32_t H, const void *dy_dx, void *grad_inputs, uint32_t gridtype, int align_corners,
                             uint32_t interp, int dtype, int layout, void *stream);
/* replaces grad_total_variation, gridencoder.h:15 (gridencoder.cu:506-645). fp32 only (grid.py:164). */
int foc_grad_total_variation(const float *inputs, const float *embeddings, float *grad, const int32_t *offsets,
                             float weight, uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H,
                             uint32_t gridtype, int align_corners, void *stream);

/* The batch of a train step from pixel indices, one launch.  Replaces, on the device, get_rays
 * (nerf/utils.py:56-157: ~15 torch launches) and the loader's ground-truth gather (nerf/provider.py collate `images[index, inds]` plus
 * train_step's RGBA blend, legacy/nerf/utils.py:468-500): ray n looks at flat pixel pix[n] (NULL: n) of camera pose_idx[n] (NULL: 0);
 * poses [n_poses,4,4] f32 row-major camera-to-world, intrinsics as the reference passes them.  images (optional, with gt): resident
 * [n_poses, H*W, channels] f32, channels 3 or 4 (4: rgb * a + bg * (1 - a)) -> gt [N,3].  rays_o / rays_d [N,3] f32. */
int foc_pixel_batch(const float *poses, uint32_t n_poses, float fx, float fy, float cx, float cy, uint32_t H, uint32_t W,
                    const int32_t *pose_idx, const int32_t *pix, uint32_t N, float *rays_o, float *rays_d, const float *images,
                    uint32_t channels, float bg, float *gt, void *stream);

/* ---------------------------------------------------------------- ffmlp (ffmlp/src/ffmlp.h:7-13) */

/* replaces ffmlp_forward / ffmlp_inference, ffmlp.h:8-9 (ffmlp.cu:331-407, 525-709).  tcgen05/TMEM kernel.
 * inputs [B,input_dim] f16, weights flat f16 ([hidden,in] + (num_layers-1)x[hidden,hidden] + [16,hidden]),
 * forward_buffer [num_layers,B,hidden] f16 (NULL = inference), outputs [B,16] f16.  B % 128 == 0,
 * hidden_dim == 64, input_dim in {32,64}, output_dim <= 16, activation relu(0)/none(6), output activation none. */
int foc_ffmlp_forward(const uint16_t *inputs, const uint16_t *weights, uint32_t B, uint32_t input_dim,
                      uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers, uint32_t activation,
                      uint32_t output_activation, uint16_t *forward_buffer, uint16_t *outputs, void *stream);
/* replaces ffmlp_backward, ffmlp.h:11 (ffmlp.cu:410-518, 749-895).  grad [B,16] f16; backward_buffer
 * [num_layers,B,hidden] f16; grad_inputs [B,input_dim] f16 or NULL; grad_weights flat f16 (overwritten).
 * scratch: foc_ffmlp_backward_scratch_bytes(...) bytes (fp32 partial weight gradients). */
size_t foc_ffmlp_backward_scratch_bytes(uint32_t B, uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers);
int foc_ffmlp_backward(const uint16_t *grad, const uint16_t *inputs, const uint16_t *weights,
                       const uint16_t *forward_buffer, uint32_t B, uint32_t input_dim, uint32_t output_dim,
                       uint32_t hidden_dim, uint32_t num_layers, uint32_t activation, uint32_t output_activation,
                       int calc_grad_inputs, uint16_t *backward_buffer, uint16_t *grad_inputs, uint16_t *grad_weights,
                       void *scratch, void *stream);
/* grad_weights == NULL in foc_ffmlp_backward leaves the per-CTA fp32 partial weight gradients in `scratch`; this reduces them (fixed
 * order, fp16 result) -- the reference's split-K GEMMs ran on side streams too (ffmlp.cu:801-812, 892-894); here the caller owns the stream. */
int foc_ffmlp_wgrad_finish(const void *scratch, uint32_t B, uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers,
                           uint16_t *grad_weights, void *stream);
/* same reduction, ADDED to what grad_weights holds (`.grad` semantics for a persistent gradient buffer zeroed by the optimiser sweep) */
int foc_ffmlp_wgrad_accumulate(const void *scratch, uint32_t B, uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers,
                               uint16_t *grad_weights, void *stream);
/* ---- device-side row counts (no reference counterpart).  A train step replayed as a CUDA graph has fixed buffer sizes, while the number
 * of samples the march produces changes every step.  foc_march_rays_train_rows publishes rows_out[0] = min(M, total rounded up to 128) in
 * device memory and zero-fills only [total, rows_out[0]); the *_rows consumers (and the optional n_rows argument of foc_field_forward /
 * foc_ffmlp_backward_head) process the first *n_rows rows only.  NULL n_rows = all B rows.  M and *n_rows are multiples of 128. */
int foc_march_rays_train_rows(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound, float dt_gamma,
                              uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, uint32_t M, const float *nears,
                              const float *fars, float *xyzs, float *dirs, float *deltas, int32_t *rays, int32_t *counter,
                              const float *noises, int32_t *rows_out, void *scratch, void *stream);
int foc_grid_encode_forward_rows(const float *inputs, const void *embeddings, const int32_t *offsets, void *outputs, uint32_t B,
                                 uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, uint32_t gridtype, int align_corners,
                                 uint32_t interp, int dtype, const int32_t *n_rows, void *stream);
int foc_grid_encode_backward_rows(const void *grad, const float *inputs, const int32_t *offsets, void *grad_embeddings, uint32_t B,
                                  uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, uint32_t gridtype, int align_corners,
                                  uint32_t interp, int dtype, const int32_t *n_rows, void *stream);
int foc_composite_mse_backward_rows(const float *grad_loss, const float *gt, float bg, const float *sigmas, const float *rgbs,
                                    const float *deltas, const int32_t *rays, const float *weights_sum, const float *image, uint32_t M,
                                    uint32_t N, float T_thresh, float *grad_sigmas, float *grad_rgbs, const int32_t *n_rows, void *stream);

/* The FOC field behind the hash grid in one kernel (nerf/network_ff.py:51-74: sigma_net -> trunc_exp / SH-4 / cat -> color_net ->
 * sigmoid; two ffmlp_forward calls (ffmlp.h:7) + ~10 elementwise torch kernels in the reference).  enc [B,32] f16, dirs [B,3] f32,
 * fp16 weight blobs in the ffmlp.cu:632 layout, hidden width 64, B % 128 == 0 -> sigma [B] f32, rgb [B,3] f32.  Training: pass
 * forward_buffer_sigma [Ls,B,64], forward_buffer_color [Lc,B,64] and cin [B,32] (all f16), which foc_ffmlp_backward_head consumes;
 * NULL for all three = inference.  Bit-identical to foc_ffmlp_forward x 2 + foc_field_mid_forward + foc_rgb_sigmoid_forward. */
int foc_field_forward(const uint16_t *enc, const float *dirs, const uint16_t *weights_sigma, const uint16_t *weights_color, uint32_t B,
                      uint32_t num_layers_sigma, uint32_t num_layers_color, uint16_t *forward_buffer_sigma,
                      uint16_t *forward_buffer_color, uint16_t *cin, float *sigma, float *rgb, const int32_t *n_rows, void *stream);
/* foc_ffmlp_backward for the two nets of the FOC field with the output gradient formed in the kernel's prologue (see csrc/mlp_bwd.cu):
 * head 1 = colour net from grad_rgb / rgb [B,3] f32; head 2 = sigma net from grad_sigma (or NULL) / sigma [B] f32 and the colour net's
 * grad_inputs grad_cin [B,32] f16.  scratch: foc_ffmlp_backward_scratch_bytes.  accumulate != 0 adds to grad_weights. */
int foc_ffmlp_backward_head(int head, const float *head_grad, const float *head_fwd, const uint16_t *grad_cin, const uint16_t *inputs,
                            const uint16_t *weights, const uint16_t *forward_buffer, uint32_t B, uint32_t input_dim, uint32_t hidden_dim,
                            uint32_t num_layers, int calc_grad_inputs, uint16_t *grad_inputs, uint16_t *grad_weights, int accumulate,
                            const int32_t *n_rows, void *scratch, void *stream);
/* kept for API compatibility with ffmlp.h:12-13; no-ops (no side streams, no workspaces) */
int foc_allocate_splitk(size_t size);
int foc_free_splitk(void);

/* ---------------------------------------------------------------- freqencoder (freqencoder/src/freqencoder.h:7-10) */
int foc_freq_encode_forward(const float *inputs, uint32_t B, uint32_t D, uint32_t deg, uint32_t C, float *outputs, void *stream);
int foc_freq_encode_backward(const float *grad, const float *outputs, uint32_t B, uint32_t D, uint32_t deg, uint32_t C,
                             float *grad_inputs, void *stream);

/* ---------------------------------------------------------------- field-network wiring (nerf/network_ff.py:51-74), fused
 * Elementwise glue between the encoders and the two MLPs, one pass per direction instead of ~25 PyTorch kernels per step.
 * forward : h [B,16] f16 (sigma-net output), dirs [B,3] f32 -> sigma [B] f32 = exp(float(h[:,0])) (activation.py:5-18),
 *           cin [B,32] f16 = [SH4(dirs) | h[:,1:16] | 0]  (network_ff.py:63-68).
 * backward: grad_h[:,0] = grad_sigma * exp(clamp(h0,-15,15)); grad_h[:,1:16] = grad_cin[:,16:31]; grad_sigma may be NULL. */
int foc_field_mid_forward(const float *dirs, const uint16_t *h, uint32_t B, float *sigma, uint16_t *cin, void *stream);
int foc_field_mid_backward(const float *grad_sigma, const uint16_t *grad_cin, const uint16_t *h, uint32_t B, uint16_t *grad_h,
                           void *stream);
/* rgb [B,3] f32 = sigmoid(hc[:, :3]) rounded through f16 as torch.sigmoid on the half tensor does (network_ff.py:72);
 * backward: grad_hc [B,16] f16 = grad * y * (1 - y) in columns 0..2, zero elsewhere. */
int foc_rgb_sigmoid_forward(const uint16_t *hc, uint32_t B, float *rgb, void *stream);
int foc_rgb_sigmoid_backward(const float *grad_rgb, const float *rgb, uint32_t B, uint16_t *grad_hc, void *stream);

/* degree <= 4 real spherical-harmonics direction encoding (`shencoder`, imported by nerf/network_ff.py:5 but ABSENT
 * from the reference tree: closed form, parity unpinned).  dirs [B,3] -> outputs [B, degree^2] f32. */
int foc_sh_encode_forward(const float *dirs, uint32_t B, uint32_t degree, float *outputs, void *stream);

/* ---------------------------------------------------------------- density-grid refresh
 * NeRFRenderer.update_extra_state, legacy/nerf/renderer.py:445-531, without its Python cell loops and host reads.
 * cell_positions: Morton cell index i (or indices[i]) of an H^3 cascade with half-extent bound_c -> jittered position
 *                 xyz = (2 c / (H-1) - 1) (bound_c - hgs) + (2 noise - 1) hgs, hgs = bound_c / H (l.470-479); noise [n,3] in [0,1).
 * grid_update   : grid = max(grid * decay, tmp) where grid >= 0 and tmp >= 0 (l.522-523); mean[0] = mean(max(grid, 0));
 *                 bitfield = packbits(grid, min(mean, density_thresh)) with the threshold read on the device (l.524-530). */
int foc_density_cell_positions(const int32_t *indices, uint32_t n, uint32_t H, float bound_c, const float *noise, float *xyzs,
                               void *stream);
size_t foc_density_grid_update_scratch_bytes(uint32_t n);
int foc_density_grid_update(float *grid, const float *tmp, uint32_t n, float decay, float density_thresh, float *mean,
                            uint8_t *bitfield, void *scratch, void *stream);
/* NeRFRenderer.mark_untrained_grid (legacy/nerf/renderer.py:380-443; FOC copy nerf/renderer.py:357-417): density_grid [cascade, H^3] cells
 * that no pose (c2w [n_poses,4,4] row-major, pinhole fx fy cx cy) sees := -1.  n_marked: optional device counter (zeroed by the caller). */
int foc_mark_untrained_grid(const float *poses, uint32_t n_poses, float fx, float fy, float cx, float cy, uint32_t H, uint32_t cascade,
                            float bound, float *density_grid, uint32_t *n_marked, void *stream);
/* foc_density_grid_update WITHOUT a caller-supplied scratch buffer and without any allocation.  No counterpart in the reference:
 * added for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  The per-CTA partial sums of the mean live in a file-scope
 * `static __device__ __align__(256)` array inside the library: FOC_NOALLOC_MAX_CELLS / 2048 floats = 32 KB, held by every GPU that
 * loads the library until the process exits.  Same kernels, same fixed summation order, hence the same bits as
 * foc_density_grid_update.  Limits and rules:
 *   - n <= FOC_NOALLOC_MAX_CELLS (8 cascades of 128^3 cells); a larger n returns FOC_ERR_TOO_LARGE and launches nothing;
 *   - calls no cudaMalloc / cudaFree / allocating library routine; the array's device address is looked up once per device
 *     (cudaGetSymbolAddress, with that device current) on the first call made on that device -- make that first call outside a
 *     graph capture;
 *   - every call on a device shares the one array and it keeps the previous call's contents; each call overwrites every slot it
 *     later reads, so nothing needs clearing, but two calls that OVERLAP on the GPU (different streams) race: serialise them. */
#define FOC_NOALLOC_MAX_CELLS (8u * 128u * 128u * 128u)
int foc_density_grid_update_noalloc(float *grid, const float *tmp, uint32_t n, float decay, float density_thresh, float *mean,
                                    uint8_t *bitfield, void *stream);

/* ---------------------------------------------------------------- loss head fused into the training compositor
 * (no reference kernel: legacy/nerf/renderer.py:313-316 + nerf/utils.py:535-545 as ~14 elementwise / reduction launches)
 * forward : composite_rays_train_forward, plus pred = image + (1 - weights_sum) * bg, err[ray] = sum_c (pred_c - gt_c)^2 and
 *           loss[0] = sum(err) / (3N) (fixed summation order).  `image` holds the un-mixed composite (the backward needs it).
 * backward: composite_rays_train_backward with g_image = grad_loss[0] * 2 (pred - gt) / (3N), g_ws = -bg * sum_c g_image_c
 *           formed on the fly.  ordered != 0 (slices in ascending ray order, as foc_march_rays_train lays them out): every
 *           gradient row without a value is zero-filled by the kernel, the caller need not pre-zero. */
int foc_composite_mse_forward(const float *sigmas, const float *rgbs, const float *deltas, const int32_t *rays, const float *gt,
                              float bg, uint32_t M, uint32_t N, float T_thresh, float *weights_sum, float *depth, float *image,
                              float *err, float *loss, void *stream);
int foc_composite_mse_backward(const float *grad_loss, const float *gt, float bg, const float *sigmas, const float *rgbs,
                               const float *deltas, const int32_t *rays, const float *weights_sum, const float *image, uint32_t M,
                               uint32_t N, float T_thresh, int ordered, float *grad_sigmas, float *grad_rgbs, void *stream);
/* The same pair with the colour network's raw output as the colour operand: hc [M,16] f16, rgb = half(sigmoid(hc[0..2])) formed in the
 * compositor (replaces torch.sigmoid of nerf/network_ff.py:73 + its backward); grad_hc [M,16] f16 is written for every row (rows
 * without a gradient and columns 3..15 are zero; slices must be in ascending ray order, as foc_march_rays_train produces). */
int foc_composite_mse_hc_forward(const float *sigmas, const uint16_t *hc, const float *deltas, const int32_t *rays, const float *gt,
                                 float bg, uint32_t M, uint32_t N, float T_thresh, float *weights_sum, float *depth, float *image,
                                 float *err, float *loss, void *stream);
int foc_composite_mse_hc_backward(const float *grad_loss, const float *gt, float bg, const float *sigmas, const uint16_t *hc,
                                  const float *deltas, const int32_t *rays, const float *weights_sum, const float *image, uint32_t M,
                                  uint32_t N, float T_thresh, float *grad_sigmas, uint16_t *grad_hc, void *stream);

/* ---------------------------------------------------------------- optimiser step of the path
 * Dense Adam + dynamic loss scaling, replacing torch.optim.Adam(betas, eps) + torch.amp.GradScaler as the reference's
 * train step uses them (main_nerf.py:211,472; nerf/utils.py:866-868), with torch's arithmetic (torch/optim/adam.py,
 * amsgrad = False, weight_decay = 0) and GradScaler.update()'s scale rule.  One sweep over the parameters also writes
 * the fp16 copy the next forward reads (`shadow16`) and zeroes the gradient in place (`zero_grad`).
 * All scalars that change between steps (lr, scale, found_inf, step, growth_tracker) are DEVICE scalars: no host sync,
 * CUDA-graph capturable.  Up to 8 tensors per call; pointers 16-byte aligned. */
typedef struct FocAdamTensor {
    float *param;        /* [n] fp32 master parameters */
    float *exp_avg;      /* [n] first moment */
    float *exp_avg_sq;   /* [n] second moment */
    void *grad;          /* [n] gradient of (loss * scale), f32 or f16 per grad_dtype; NULL = tensor is skipped */
    uint16_t *shadow16;  /* [n] optional fp16 copy of the updated parameters, NULL = none */
    uint64_t n;
    int32_t grad_dtype;  /* 0 = f32, 1 = f16 */
    int32_t zero_grad;   /* != 0: the gradient is zeroed after use (also on a skipped step) */
} FocAdamTensor;
/* *found_inf = 1.0f when any gradient element is inf / nan (GradScaler._unscale_grads_' check); never cleared here */
int foc_adam_nonfinite(const FocAdamTensor *tensors, int count, float *found_inf, void *stream);
/* t = *step + 1; skipped entirely when *found_inf != 0.  scale / found_inf may be NULL (no loss scaling). */
int foc_adam_step(const FocAdamTensor *tensors, int count, const float *lr, float beta1, float beta2, float eps,
                  const float *scale, const float *found_inf, const int32_t *step, void *stream);
/* GradScaler.update(): found_inf -> scale *= backoff, tracker = 0; else tracker += 1 and every growth_interval clean
 * steps scale *= growth.  Also *step += 1 on a clean step, and *found_inf = 0.  scale may be NULL (step counter only). */
int foc_scaler_update(float *scale, int32_t *growth_tracker, float *found_inf, int32_t *step, float growth_factor,
                      float backoff_factor, int32_t growth_interval, void *stream);
/* foc_scaler_update + the reference's per-iteration learning-rate schedule (main_nerf.py:477: LambdaLR(optimizer, lambda iter:
 * 0.1 ** min(iter / opt.iters, 1)) with scheduler_update_every_step=True, nerf/utils.py:869-870), on the device so that a
 * CUDA-graph-replayed step needs no host write of lr: *sched_iter += 1 (skipped steps count, as scheduler.step() does);
 * *lr = lr0 * 0.1 ** min(*sched_iter / decay_iters, 1). */
int foc_scaler_update_sched(float *scale, int32_t *growth_tracker, float *found_inf, int32_t *step, float growth_factor,
                            float backoff_factor, int32_t growth_interval, float *lr, int32_t *sched_iter, float lr0,
                            float decay_iters, void *stream);

/* ---------------------------------------------------------------- data-parallel training over NVLink peer memory
 * No reference counterpart (the reference trains on one GPU; torch DDP would all-reduce every gradient over NCCL and
 * repeat the whole Adam sweep on every rank).  One process per GPU; every rank owns 1/world of the hash table's fp32
 * master parameters and Adam moments.  Per step: barrier -> foc_dp_reduce (pull + average this rank's shard of the fp16
 * table gradients from all peers, deterministic rank order) -> barrier exchanging the non-finite flags ->
 * foc_adam_step_peers (sweep over the shard, fp16 parameter copy pushed into every peer's table) -> barrier.
 * See csrc/peer.cu.  All of it is plain kernel launches: CUDA-graph capturable, no host synchronisation. */
#define FOC_MAX_PEERS 8
#define FOC_PEER_HANDLE_BYTES 64
/* peer-visible device memory: cudaMalloc'd (zero-filled), exported as a 64-byte IPC handle that another process of the
 * same node opens (peer access is enabled on first open).  A process must not open its own handles. */
int foc_peer_alloc(uint64_t bytes, void **ptr);
int foc_peer_free(void *ptr);
int foc_peer_export(void *ptr, uint8_t *handle64);
int foc_peer_open(const uint8_t *handle64, void **ptr);
int foc_peer_close(void *ptr);
typedef struct FocPeerGroup {
    int32_t world, rank;
    uint32_t *flags[FOC_MAX_PEERS]; /* every rank's signal pad (>= 24 uint32, zero-initialised), [rank] = the local one */
    uint32_t *state;                /* local: [0] epoch, [1] non-finite flag of foc_dp_reduce, [2] sticky status (1 = barrier timeout) */
} FocPeerGroup;
/* All ranks of the group meet; found_inf != NULL additionally ORs the ranks' flags (state[1], or a *found_inf that is already non-zero
 * on entry: the result of a local foc_adam_nonfinite) into *found_inf (sets 1.0f, never clears) and resets state[1].  Gives up after ~2 s without progress and sets state[2] instead of hanging. */
int foc_peer_barrier(const FocPeerGroup *group, float *found_inf, void *stream);
typedef struct FocDpReduce {
    int32_t world, rank;
    const uint16_t *grad16[FOC_MAX_PEERS]; /* every rank's fp16 gradient table (whole table; [rank] = local) */
    uint64_t shard_start, shard_n;         /* this rank's element range of the table; multiples of 8 */
    float *reduced;                        /* [shard_n] local: mean over ranks, fp32 */
    const float *tail[FOC_MAX_PEERS];      /* every rank's flat fp32 gradients of the replicated small tensors */
    uint64_t tail_n;                       /* multiple of 4 */
    float *reduced_tail;                   /* [tail_n] local: mean over ranks (computed in full by every rank) */
    uint32_t *inf_flag;                    /* local: set to 1 when a reduced value is inf / nan (FocPeerGroup.state + 1) */
    uint32_t shard_blocks;                 /* filled in by the callee */
} FocDpReduce;
int foc_dp_reduce(const FocDpReduce *args, void *stream);
/* foc_adam_step, and additionally tensor `peer_tensor`'s fp16 parameter copy is stored into peer_shadow16[r] (r < world,
 * already offset to this rank's shard: same indexing as tensors[peer_tensor].shadow16, which may be NULL) */
int foc_adam_step_peers(const FocAdamTensor *tensors, int count, const float *lr, float beta1, float beta2, float eps,
                        const float *scale, const float *found_inf, const int32_t *step, int peer_tensor, int world,
                        uint16_t *const *peer_shadow16, void *stream);
/* The exchange INSIDE the sweep (no reference counterpart; replaces foc_dp_reduce + foc_adam_step_peers and one of the three barriers):
 * tensor `peer_tensor` (this rank's shard of the table: param / exp_avg / exp_avg_sq point at the shard, n = its length, a multiple of
 * 8) takes as gradient the mean over the ranks of grads->grad16[r][i] (fp16, every pointer offset to the shard), summed in fp32 in rank
 * order; every other tensor's `grad` must point into grads->tail[rank] (fp32, 4-float aligned, rows zero-padded to 4) and takes the mean
 * of the same offset in every rank's tail.  Updated fp16 copies of the shard go to peer_shadow16[r] as in foc_adam_step_peers.
 * Protocol per step: local foc_adam_nonfinite over the rank's own gradients -> foc_peer_barrier(found_inf) [gradients complete, flags
 * exchanged: a sum of <= 8 finite fp16 values is finite, so "any rank non-finite" == "mean non-finite"] -> foc_adam_step_dp ->
 * foc_peer_barrier [shards delivered, everybody has pulled] -> zero the gradients. */
typedef struct FocDpGrads {
    int32_t world, rank;
    const uint16_t *grad16[FOC_MAX_PEERS];
    const float *tail[FOC_MAX_PEERS];
} FocDpGrads;
int foc_adam_step_dp(const FocAdamTensor *tensors, int count, const float *lr, float beta1, float beta2, float eps,
                     const float *scale, const float *found_inf, const int32_t *step, int peer_tensor, const FocDpGrads *grads,
                     uint16_t *const *peer_shadow16, void *stream);

/* Epilogue of an inference render (legacy/nerf/renderer.py:369-372), one launch: image + (1 - weights_sum) * bg, depth
 * normalised to [near, far] (depth_from_first_sample != 0: `depth` came from the training compositor, which measures from
 * the first sample -- near * weights_sum is added first), packed into rows [r, g, b, depth, weights_sum].  The N rays are
 * this rank's share of an n_total-ray frame dealt in 128-ray blocks (block b -> rank b % world; world = 1: the whole
 * frame): every block's rows are stored into frame[r] (an [n_total, 5] f32 buffer of rank r, peer-visible) for all r.
 * Follow with foc_peer_barrier before any rank reads its frame. */
typedef struct FocFrameOut {
    int32_t world, rank;
    uint32_t n_total;
    float *frame[FOC_MAX_PEERS];
} FocFrameOut;
int foc_render_finish(const float *image, const float *depth, const float *weights_sum, const float *nears, const float *fars,
                      uint32_t N, float bg, int depth_from_first_sample, const FocFrameOut *out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* FOC_B200_H */