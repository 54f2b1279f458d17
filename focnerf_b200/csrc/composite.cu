// composite.cu -- front-to-back alpha compositing for sm_100a.
//
// Replaces the reference kernels K7, K8, K10 (raymarching/src/raymarching.cu:501-577, 602-682, 819-905), the
// Python boolean-mask compaction of the render loop (legacy/nerf/renderer.py:363) and FOC-NeRF's multi-object
// composition (COMBINED.py:141-200, 247-251, 515-526; editable.py:443-471) behind the C ABI of include/foc_b200.h.
//
// Training compositors: ONE WARP PER RAY.  A ray's samples are a contiguous slice [offset, offset+count) of the
// sample arrays, so the 32 lanes read sigma / rgb / delta with coalesced 128 / 384 / 256-byte requests (the
// reference gives each ray to one thread, which walks its slice serially with 4/12/8-byte strided loads).  The
// recurrences T_i = prod(1-alpha), t_i = sum(delta1), acc_i = sum(w c) become warp-shuffle inclusive scans over
// 32-sample chunks with the carry kept in registers; early termination (T < T_thresh, tested AFTER the sample
// is accumulated, raymarching.cu:566) is a ballot + find-first-set.
// Float results agree with the reference to rounding (different association of the same products/sums), which
// BASELINE.json bounds at 1e-3 relative; nothing here produces integers.
#include "common.cuh"

using namespace foc;

namespace {

__device__ __forceinline__ float excl_from_incl_prod(float incl, int lane) {
    const float up = __shfl_up_sync(0xffffffffu, incl, 1);
    return lane == 0 ? 1.0f : up;
}
__device__ __forceinline__ float bcast_last(float v) { return __shfl_sync(0xffffffffu, v, 31); }

constexpr int kCompWarpsPerCta = 4;   // 128 threads, 4 rays per CTA

}  // namespace

// K7 (raymarching.cu:501-577): weights_sum / depth / image are indexed by the ray id stored in rays[n,0].
__global__ void __launch_bounds__(32 * kCompWarpsPerCta)
k_composite_train_fwd(const float* __restrict__ sigmas, const float* __restrict__ rgbs, const float* __restrict__ deltas,
                      const int32_t* __restrict__ rays, uint32_t M, uint32_t N, float T_thresh,
                      float* __restrict__ weights_sum, float* __restrict__ depth, float* __restrict__ image) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const uint32_t n = blockIdx.x * kCompWarpsPerCta + (threadIdx.x >> 5);
    if (n >= N) return;
    const uint32_t id = (uint32_t)__ldg(rays + 3 * (size_t)n), off = (uint32_t)__ldg(rays + 3 * (size_t)n + 1),
                   cnt = (uint32_t)__ldg(rays + 3 * (size_t)n + 2);
    float r = 0.f, g = 0.f, b = 0.f, d = 0.f, ws = 0.f;
    if (cnt != 0 && (uint64_t)off + cnt <= M) {            // raymarching.cu:521-528
        float T = 1.0f, t_base = 0.0f;
        for (uint32_t base = 0; base < cnt; base += 32) {
            const uint32_t s = base + lane;
            const bool valid = s < cnt;
            const size_t i = (size_t)off + (valid ? s : 0);
            float alpha = 0.f, d1 = 0.f, cr = 0.f, cg = 0.f, cb = 0.f;
            if (valid) {
                const float2 dl = __ldg(reinterpret_cast<const float2*>(deltas) + i);
                alpha = 1.0f - __expf(-__ldg(sigmas + i) * dl.x);
                d1 = dl.y;
                cr = __ldg(rgbs + 3 * i); cg = __ldg(rgbs + 3 * i + 1); cb = __ldg(rgbs + 3 * i + 2);
            }
            const float incl = warp_incl_prod(1.0f - alpha, lane);
            const float w = alpha * (T * excl_from_incl_prod(incl, lane));
            const float T_after = T * incl;
            const float t = t_base + warp_incl_sum(d1, lane);
            const uint32_t stop = __ballot_sync(0xffffffffu, valid && T_after < T_thresh);
            const int last = stop ? (__ffs(stop) - 1) : 31;
            if (valid && lane <= last) {
                r = fmaf(w, cr, r); g = fmaf(w, cg, g); b = fmaf(w, cb, b);
                d = fmaf(w, t, d); ws += w;
            }
            if (stop) break;
            T = bcast_last(T_after); t_base = bcast_last(t);
        }
        r = warp_sum(r); g = warp_sum(g); b = warp_sum(b); d = warp_sum(d); ws = warp_sum(ws);
    }
    if (lane == 0) {
        weights_sum[id] = ws; depth[id] = d;
        image[3 * (size_t)id] = r; image[3 * (size_t)id + 1] = g; image[3 * (size_t)id + 2] = b;
    }
}

// K8 (raymarching.cu:602-682).  grad_sigma_i = delta0_i * ( sum_c g_c (T_after_i c_i - (final_c - acc_c,i))
//                                                     + g_ws (1 - ws_final) ),  acc INCLUDING sample i.
// Samples after the early-termination sample are not written (the caller pre-zeroes, raymarching.py:280-281).
__global__ void __launch_bounds__(32 * kCompWarpsPerCta)
k_composite_train_bwd(const float* __restrict__ grad_weights_sum, const float* __restrict__ grad_image,
                      const float* __restrict__ sigmas, const float* __restrict__ rgbs, const float* __restrict__ deltas,
                      const int32_t* __restrict__ rays, const float* __restrict__ weights_sum, const float* __restrict__ image,
                      uint32_t M, uint32_t N, float T_thresh, float* __restrict__ grad_sigmas, float* __restrict__ grad_rgbs) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const uint32_t n = blockIdx.x * kCompWarpsPerCta + (threadIdx.x >> 5);
    if (n >= N) return;
    const uint32_t id = (uint32_t)__ldg(rays + 3 * (size_t)n), off = (uint32_t)__ldg(rays + 3 * (size_t)n + 1),
                   cnt = (uint32_t)__ldg(rays + 3 * (size_t)n + 2);
    if (cnt == 0 || (uint64_t)off + cnt > M) return;
    const float gw = __ldg(grad_weights_sum + id);
    const float gr = __ldg(grad_image + 3 * (size_t)id), gg = __ldg(grad_image + 3 * (size_t)id + 1), gb = __ldg(grad_image + 3 * (size_t)id + 2);
    const float rf = __ldg(image + 3 * (size_t)id), gf = __ldg(image + 3 * (size_t)id + 1), bf = __ldg(image + 3 * (size_t)id + 2);
    const float ws_term = gw * (1.0f - __ldg(weights_sum + id));
    float T = 1.0f, ar = 0.f, ag = 0.f, ab = 0.f;       // carries: transmittance and accumulated colour
    for (uint32_t base = 0; base < cnt; base += 32) {
        const uint32_t s = base + lane;
        const bool valid = s < cnt;
        const size_t i = (size_t)off + (valid ? s : 0);
        float alpha = 0.f, d0 = 0.f, cr = 0.f, cg = 0.f, cb = 0.f;
        if (valid) {
            d0 = __ldg(deltas + 2 * i);
            alpha = 1.0f - __expf(-__ldg(sigmas + i) * d0);
            cr = __ldg(rgbs + 3 * i); cg = __ldg(rgbs + 3 * i + 1); cb = __ldg(rgbs + 3 * i + 2);
        }
        const float incl = warp_incl_prod(1.0f - alpha, lane);
        const float w = alpha * (T * excl_from_incl_prod(incl, lane));
        const float T_after = T * incl;
        const float pr = ar + warp_incl_sum(w * cr, lane), pg = ag + warp_incl_sum(w * cg, lane), pb = ab + warp_incl_sum(w * cb, lane);
        const uint32_t stop = __ballot_sync(0xffffffffu, valid && T_after < T_thresh);
        const int last = stop ? (__ffs(stop) - 1) : 31;
        if (valid && lane <= last) {
            grad_rgbs[3 * i] = gr * w; grad_rgbs[3 * i + 1] = gg * w; grad_rgbs[3 * i + 2] = gb * w;
            grad_sigmas[i] = d0 * (gr * (T_after * cr - (rf - pr)) + gg * (T_after * cg - (gf - pg)) +
                                   gb * (T_after * cb - (bf - pb)) + ws_term);
        }
        if (stop) break;
        T = bcast_last(T_after); ar = bcast_last(pr); ag = bcast_last(pg); ab = bcast_last(pb);
    }
}

// ---- loss head fused into the compositor (no reference kernel: legacy/nerf/renderer.py:313-316 + nerf/utils.py:535-545
// run ~14 elementwise / reduction kernels for  pred = image + (1 - weights_sum) * bg;  loss = mean((pred - gt)^2) )
//
// forward : the K7 scan, then lane 0 also writes err[id] = sum_c (pred_c - gt_c)^2; k_mse_finish sums err in a fixed order.
// backward: the K8 scan with  g_image = grad_loss * 2 (pred - gt) / (3N)  and  g_ws = -bg * sum_c g_image_c  formed in
//           registers.  With `ordered` (slices in ascending ray order, what foc_march_rays_train produces) the kernel also
//           zero-fills every gradient row it does not own a value for -- past the termination sample, the rows of dropped
//           rays and the padding rows [total, M) -- so the caller need not pre-zero the buffers.
// HC = true: the colour operand is the colour network's raw output hc [M,16] f16 and rgb = half(sigmoid(hc[0..2])) is formed here
// with k_rgb_fwd's operations (field_glue.cu) -- the rgb [M,3] tensor and its launch disappear; the backward then emits
// grad_hc [M,16] f16 directly (k_rgb_bwd's operations).  Same bits as the three-kernel sequence.
template <bool HC>
__device__ __forceinline__ void load_colour(const void* __restrict__ colours, size_t i, float& cr, float& cg, float& cb) {
    if constexpr (HC) {
        const uint2 v = __ldg(reinterpret_cast<const uint2*>(static_cast<const __half*>(colours) + i * 16));
        const __half* hv = reinterpret_cast<const __half*>(&v);
        cr = __half2float(__float2half_rn(1.0f / (1.0f + expf(-__half2float(hv[0])))));
        cg = __half2float(__float2half_rn(1.0f / (1.0f + expf(-__half2float(hv[1])))));
        cb = __half2float(__float2half_rn(1.0f / (1.0f + expf(-__half2float(hv[2])))));
    } else {
        const float* rgbs = static_cast<const float*>(colours);
        cr = __ldg(rgbs + 3 * i); cg = __ldg(rgbs + 3 * i + 1); cb = __ldg(rgbs + 3 * i + 2);
    }
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) { const __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<const uint32_t*>(&h); }

template <bool HC>
__global__ void __launch_bounds__(32 * kCompWarpsPerCta)
k_composite_mse_fwd(const float* __restrict__ sigmas, const void* __restrict__ rgbs, const float* __restrict__ deltas,
                    const int32_t* __restrict__ rays, const float* __restrict__ gt, float bg, uint32_t M, uint32_t N, float T_thresh,
                    float* __restrict__ weights_sum, float* __restrict__ depth, float* __restrict__ image, float* __restrict__ err) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const uint32_t n = blockIdx.x * kCompWarpsPerCta + (threadIdx.x >> 5);
    if (n >= N) return;
    const uint32_t id = (uint32_t)__ldg(rays + 3 * (size_t)n), off = (uint32_t)__ldg(rays + 3 * (size_t)n + 1),
                   cnt = (uint32_t)__ldg(rays + 3 * (size_t)n + 2);
    float r = 0.f, g = 0.f, b = 0.f, d = 0.f, ws = 0.f;
    if (cnt != 0 && (uint64_t)off + cnt <= M) {
        float T = 1.0f, t_base = 0.0f;
        for (uint32_t base = 0; base < cnt; base += 32) {
            const uint32_t s = base + lane;
            const bool valid = s < cnt;
            const size_t i = (size_t)off + (valid ? s : 0);
            float alpha = 0.f, d1 = 0.f, cr = 0.f, cg = 0.f, cb = 0.f;
            if (valid) {
                const float2 dl = __ldg(reinterpret_cast<const float2*>(deltas) + i);
                alpha = 1.0f - __expf(-__ldg(sigmas + i) * dl.x);
                d1 = dl.y;
                load_colour<HC>(rgbs, i, cr, cg, cb);
            }
            const float incl = warp_incl_prod(1.0f - alpha, lane);
            const float w = alpha * (T * excl_from_incl_prod(incl, lane));
            const float T_after = T * incl;
            const float t = t_base + warp_incl_sum(d1, lane);
            const uint32_t stop = __ballot_sync(0xffffffffu, valid && T_after < T_thresh);
            const int last = stop ? (__ffs(stop) - 1) : 31;
            if (valid && lane <= last) {
                r = fmaf(w, cr, r); g = fmaf(w, cg, g); b = fmaf(w, cb, b);
                d = fmaf(w, t, d); ws += w;
            }
            if (stop) break;
            T = bcast_last(T_after); t_base = bcast_last(t);
        }
        r = warp_sum(r); g = warp_sum(g); b = warp_sum(b); d = warp_sum(d); ws = warp_sum(ws);
    }
    if (lane == 0) {
        weights_sum[id] = ws; depth[id] = d;
        image[3 * (size_t)id] = r; image[3 * (size_t)id + 1] = g; image[3 * (size_t)id + 2] = b;
        const float mix = (1.0f - ws) * bg;
        const float e0 = (r + mix) - __ldg(gt + 3 * (size_t)id), e1 = (g + mix) - __ldg(gt + 3 * (size_t)id + 1), e2 = (b + mix) - __ldg(gt + 3 * (size_t)id + 2);
        err[id] = e0 * e0 + e1 * e1 + e2 * e2;
    }
}

// loss[0] = sum(err[0..N)) / (3N), one CTA, fixed summation order
__global__ void __launch_bounds__(1024) k_mse_finish(const float* __restrict__ err, uint32_t N, float* __restrict__ loss) {
    pdl_sync();
    __shared__ float part[32];
    float acc = 0.f;
    for (uint32_t i = threadIdx.x; i < N; i += 1024) acc += __ldg(err + i);
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        const float t = warp_sum(part[threadIdx.x]);
        if (threadIdx.x == 0) loss[0] = t / (3.0f * (float)N);
    }
}

template <bool HC>
__global__ void __launch_bounds__(32 * kCompWarpsPerCta)
k_composite_mse_bwd(const float* __restrict__ grad_loss, const float* __restrict__ gt, float bg,
                    const float* __restrict__ sigmas, const void* __restrict__ rgbs, const float* __restrict__ deltas,
                    const int32_t* __restrict__ rays, const float* __restrict__ weights_sum, const float* __restrict__ image,
                    uint32_t M, uint32_t N, float T_thresh, int ordered, float* __restrict__ grad_sigmas, void* __restrict__ grad_colours,
                    const int32_t* __restrict__ n_rows) {
    float* grad_rgbs = static_cast<float*>(grad_colours);            // HC = false: [M,3] f32
    uint4* grad_hc = static_cast<uint4*>(grad_colours);              // HC = true : [M,16] f16 = two 16-byte words per row
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const uint32_t n = blockIdx.x * kCompWarpsPerCta + (threadIdx.x >> 5);
    if (n >= N) return;
    const uint32_t id = (uint32_t)__ldg(rays + 3 * (size_t)n), off = (uint32_t)__ldg(rays + 3 * (size_t)n + 1),
                   cnt = (uint32_t)__ldg(rays + 3 * (size_t)n + 2);
    auto zero_rows = [&](uint64_t lo, uint64_t hi) {                  // rows [lo, hi) of both gradient arrays
        for (uint64_t i = lo + lane; i < hi; i += 32) {
            grad_sigmas[i] = 0.f;
            if constexpr (HC) { grad_hc[2 * i] = make_uint4(0u, 0u, 0u, 0u); grad_hc[2 * i + 1] = make_uint4(0u, 0u, 0u, 0u); }
            else { grad_rgbs[3 * i] = 0.f; grad_rgbs[3 * i + 1] = 0.f; grad_rgbs[3 * i + 2] = 0.f; }
        }
    };
    const bool kept = (uint64_t)off + cnt <= M;
    if (ordered) {
        if (!kept && off < M) zero_rows(off, M);                       // the FIRST dropped ray (off <= M < off + cnt) owns the rest (< max_steps rows)
        // nothing dropped: the padding rows [total, M) -- up to a whole buffer quantum (16 k rows) when the step is replayed as a graph --
        // are dealt to ALL warps in 32-row pieces (one warp clearing them alone was the kernel's tail)
        const uint64_t total = (uint64_t)(uint32_t)__ldg(rays + 3 * (size_t)(N - 1) + 1) + (uint32_t)__ldg(rays + 3 * (size_t)(N - 1) + 2);
        // n_rows (optional, device memory): the consumers of the gradients only read the first *n_rows rows -- no need to clear beyond
        uint64_t hi = M;
        if (n_rows != nullptr) { const uint64_t v = (uint64_t)max(__ldg(n_rows), 0); hi = v < hi ? v : hi; }
        for (uint64_t i = total + (uint64_t)n * 32u + lane; i < hi; i += (uint64_t)N * 32u) {
            grad_sigmas[i] = 0.f;
            if constexpr (HC) { grad_hc[2 * i] = make_uint4(0u, 0u, 0u, 0u); grad_hc[2 * i + 1] = make_uint4(0u, 0u, 0u, 0u); }
            else { grad_rgbs[3 * i] = 0.f; grad_rgbs[3 * i + 1] = 0.f; grad_rgbs[3 * i + 2] = 0.f; }
        }
    }
    if (cnt == 0 || !kept) return;
    const float rf = __ldg(image + 3 * (size_t)id), gf = __ldg(image + 3 * (size_t)id + 1), bf = __ldg(image + 3 * (size_t)id + 2);
    const float wsf = __ldg(weights_sum + id);
    const float mix = (1.0f - wsf) * bg, k = __ldg(grad_loss) * (2.0f / (3.0f * (float)N));
    const float gr = k * ((rf + mix) - __ldg(gt + 3 * (size_t)id)), gg = k * ((gf + mix) - __ldg(gt + 3 * (size_t)id + 1)),
                gb = k * ((bf + mix) - __ldg(gt + 3 * (size_t)id + 2));
    const float gw = -bg * (gr + gg + gb);
    const float ws_term = gw * (1.0f - wsf);
    float T = 1.0f, ar = 0.f, ag = 0.f, ab = 0.f;
    uint32_t done = cnt;                                               // first row of this ray that holds no gradient
    for (uint32_t base = 0; base < cnt; base += 32) {
        const uint32_t s = base + lane;
        const bool valid = s < cnt;
        const size_t i = (size_t)off + (valid ? s : 0);
        float alpha = 0.f, d0 = 0.f, cr = 0.f, cg = 0.f, cb = 0.f;
        if (valid) {
            d0 = __ldg(deltas + 2 * i);
            alpha = 1.0f - __expf(-__ldg(sigmas + i) * d0);
            load_colour<HC>(rgbs, i, cr, cg, cb);
        }
        const float incl = warp_incl_prod(1.0f - alpha, lane);
        const float w = alpha * (T * excl_from_incl_prod(incl, lane));
        const float T_after = T * incl;
        const float pr = ar + warp_incl_sum(w * cr, lane), pg = ag + warp_incl_sum(w * cg, lane), pb = ab + warp_incl_sum(w * cb, lane);
        const uint32_t stop = __ballot_sync(0xffffffffu, valid && T_after < T_thresh);
        const int last = stop ? (__ffs(stop) - 1) : 31;
        if (valid && lane <= last) {
            if constexpr (HC) {                                        // k_rgb_bwd: the gradient of a half tensor is half; d sigmoid = y (1 - y)
                const float g0 = __half2float(__float2half_rn(gr * w)) * (1.0f - cr) * cr, g1 = __half2float(__float2half_rn(gg * w)) * (1.0f - cg) * cg,
                            g2 = __half2float(__float2half_rn(gb * w)) * (1.0f - cb) * cb;
                grad_hc[2 * i] = make_uint4(pack_h2(g0, g1), pack_h2(g2, 0.f), 0u, 0u); grad_hc[2 * i + 1] = make_uint4(0u, 0u, 0u, 0u);
            } else { grad_rgbs[3 * i] = gr * w; grad_rgbs[3 * i + 1] = gg * w; grad_rgbs[3 * i + 2] = gb * w; }
            grad_sigmas[i] = d0 * (gr * (T_after * cr - (rf - pr)) + gg * (T_after * cg - (gf - pg)) +
                                   gb * (T_after * cb - (bf - pb)) + ws_term);
        }
        if (stop) { done = base + (uint32_t)last + 1; break; }
        T = bcast_last(T_after); ar = bcast_last(pr); ag = bcast_last(pg); ab = bcast_last(pb);
    }
    if (ordered) zero_rows((uint64_t)off + done, (uint64_t)off + cnt);
}

// ------------------------------------------------------------------------------------------------ C ABI
static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

FOC_API int foc_composite_rays_train_forward(const float* sigmas, const float* rgbs, const float* deltas, const int32_t* rays,
                                             uint32_t M, uint32_t N, float T_thresh, float* weights_sum, float* depth,
                                             float* image, void* stream) {
    if (N == 0) return 0;
    foc::launch(k_composite_train_fwd, ceil_div(N, kCompWarpsPerCta), 32 * kCompWarpsPerCta, 0, S(stream), 
        sigmas, rgbs, deltas, rays, M, N, T_thresh, weights_sum, depth, image);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_composite_rays_train_backward(const float* grad_weights_sum, const float* grad_image, const float* sigmas,
                                              const float* rgbs, const float* deltas, const int32_t* rays,
                                              const float* weights_sum, const float* image, uint32_t M, uint32_t N,
                                              float T_thresh, float* grad_sigmas, float* grad_rgbs, void* stream) {
    if (N == 0) return 0;
    foc::launch(k_composite_train_bwd, ceil_div(N, kCompWarpsPerCta), 32 * kCompWarpsPerCta, 0, S(stream), 
        grad_weights_sum, grad_image, sigmas, rgbs, deltas, rays, weights_sum, image, M, N, T_thresh, grad_sigmas, grad_rgbs);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_composite_mse_forward(const float* sigmas, const float* rgbs, const float* deltas, const int32_t* rays, const float* gt,
                                      float bg, uint32_t M, uint32_t N, float T_thresh, float* weights_sum, float* depth, float* image,
                                      float* err, float* loss, void* stream) {
    if (N == 0) return 0;
    foc::launch(k_composite_mse_fwd<false>, ceil_div(N, kCompWarpsPerCta), 32 * kCompWarpsPerCta, 0, S(stream),
        sigmas, (const void*)rgbs, deltas, rays, gt, bg, M, N, T_thresh, weights_sum, depth, image, err);
    FOC_LAUNCH_CHECK();
    foc::launch(k_mse_finish, 1, 1024, 0, S(stream), err, N, loss);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_composite_mse_backward(const float* grad_loss, const float* gt, float bg, const float* sigmas, const float* rgbs,
                                       const float* deltas, const int32_t* rays, const float* weights_sum, const float* image, uint32_t M,
                                       uint32_t N, float T_thresh, int ordered, float* grad_sigmas, float* grad_rgbs, void* stream) {
    if (N == 0) return 0;
    foc::launch(k_composite_mse_bwd<false>, ceil_div(N, kCompWarpsPerCta), 32 * kCompWarpsPerCta, 0, S(stream),
        grad_loss, gt, bg, sigmas, (const void*)rgbs, deltas, rays, weights_sum, image, M, N, T_thresh, ordered, grad_sigmas, (void*)grad_rgbs,
        (const int32_t*)nullptr);
    FOC_LAUNCH_CHECK(); return 0;
}

// foc_composite_mse_backward(ordered = 1) that clears the padding rows only up to *n_rows (device memory): see foc_march_rays_train_rows
FOC_API int foc_composite_mse_backward_rows(const float* grad_loss, const float* gt, float bg, const float* sigmas, const float* rgbs,
                                            const float* deltas, const int32_t* rays, const float* weights_sum, const float* image, uint32_t M,
                                            uint32_t N, float T_thresh, float* grad_sigmas, float* grad_rgbs, const int32_t* n_rows, void* stream) {
    if (N == 0) return 0;
    if (n_rows == nullptr) return FOC_ERR_UNSUPPORTED;
    foc::launch(k_composite_mse_bwd<false>, ceil_div(N, kCompWarpsPerCta), 32 * kCompWarpsPerCta, 0, S(stream),
        grad_loss, gt, bg, sigmas, (const void*)rgbs, deltas, rays, weights_sum, image, M, N, T_thresh, 1, grad_sigmas, (void*)grad_rgbs, n_rows);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_composite_mse_hc_forward(const float* sigmas, const uint16_t* hc, const float* deltas, const int32_t* rays, const float* gt,
                                         float bg, uint32_t M, uint32_t N, float T_thresh, float* weights_sum, float* depth, float* image,
                                         float* err, float* loss, void* stream) {
    if (N == 0) return 0;
    foc::launch(k_composite_mse_fwd<true>, ceil_div(N, kCompWarpsPerCta), 32 * kCompWarpsPerCta, 0, S(stream),
        sigmas, (const void*)hc, deltas, rays, gt, bg, M, N, T_thresh, weights_sum, depth, image, err);
    FOC_LAUNCH_CHECK();
    foc::launch(k_mse_finish, 1, 1024, 0, S(stream), err, N, loss);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_composite_mse_hc_backward(const float* grad_loss, const float* gt, float bg, const float* sigmas, const uint16_t* hc,
                                          const float* deltas, const int32_t* rays, const float* weights_sum, const float* image, uint32_t M,
                                          uint32_t N, float T_thresh, float* grad_sigmas, uint16_t* grad_hc, void* stream) {
    if (N == 0) return 0;
    if ((reinterpret_cast<uintptr_t>(hc) & 7u) || (reinterpret_cast<uintptr_t>(grad_hc) & 15u)) return FOC_ERR_UNSUPPORTED;
    foc::launch(k_composite_mse_bwd<true>, ceil_div(N, kCompWarpsPerCta), 32 * kCompWarpsPerCta, 0, S(stream),
        grad_loss, gt, bg, sigmas, (const void*)hc, deltas, rays, weights_sum, image, M, N, T_thresh, 1, grad_sigmas, (void*)grad_hc, (const int32_t*)nullptr);
    FOC_LAUNCH_CHECK(); return 0;
}
