// composite_infer.cu -- inference-loop compositor (K10) and the alive-list compaction that follows it.
//
// K10 (raymarching/src/raymarching.cu:819-905) accumulates at most n_step (<= 8 in the caller,
// legacy/nerf/renderer.py:349) samples per alive ray into the per-ray running sums, in place.  n_step is tiny, so
// one thread per alive ray is the right shape: with n_step == 1 (most of a render) the sample loads are fully
// coalesced; the per-ray state (28 B read + written) is gathered through rays_alive[].
//   T is 1 - sum(w) here (raymarching.cu:868), NOT the running product of the training compositor.
//
// foc_compact_alive replaces `rays_alive = rays_alive[rays_alive >= 0]` (legacy/nerf/renderer.py:363; torch:
// nonzero + index_select + a D2H sync for the new length) by a stable two-launch ballot/prefix compaction whose
// count stays on the device.
#include "common.cuh"

using namespace foc;

__global__ void __launch_bounds__(128)
k_composite_rays(uint32_t n_alive, uint32_t n_step, float T_thresh, int32_t* __restrict__ rays_alive,
                 float* __restrict__ rays_t, const float* __restrict__ sigmas, const float* __restrict__ rgbs,
                 const float* __restrict__ deltas, float* __restrict__ weights_sum, float* __restrict__ depth,
                 float* __restrict__ image) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_alive) return;
    const int32_t id = rays_alive[n];
    if (id < 0) return;                                   // defensive: the reference would index out of bounds
    float t = rays_t[id], ws = weights_sum[id], d = depth[id];
    float r = image[3 * (size_t)id], g = image[3 * (size_t)id + 1], b = image[3 * (size_t)id + 2];
    const size_t base = (size_t)n * n_step;
    uint32_t s = 0;
    for (; s < n_step; ++s) {
        const size_t i = base + s;
        const float2 dl = __ldg(reinterpret_cast<const float2*>(deltas) + i);
        if (dl.x == 0.0f) break;                          // zero-filled slot: the ray left the box (raymarching.cu:858)
        const float alpha = 1.0f - __expf(-__ldg(sigmas + i) * dl.x);
        const float T = 1.0f - ws;
        const float w = alpha * T;
        ws += w;
        t += dl.y;
        d = fmaf(w, t, d);
        r = fmaf(w, __ldg(rgbs + 3 * i), r); g = fmaf(w, __ldg(rgbs + 3 * i + 1), g); b = fmaf(w, __ldg(rgbs + 3 * i + 2), b);
        if (T < T_thresh) break;                          // T BEFORE this sample (raymarching.cu:885)
    }
    if (s < n_step) rays_alive[n] = -1; else rays_t[id] = t;
    weights_sum[id] = ws; depth[id] = d;
    image[3 * (size_t)id] = r; image[3 * (size_t)id + 1] = g; image[3 * (size_t)id + 2] = b;
}

// K10's recurrence over a DEPTH WINDOW of ray-ordered samples (render without the host-driven loop: a frame is a few windows
// over the records of foc_march_rays_count).  One warp per ray: the window's slice [off, off + cnt) is read with coalesced
// requests, the carried state (weights_sum, depth, image, rays_t) is updated in place, alive[n] is cleared when the ray
// terminated -- K10's rule: stop AFTER accumulating the first sample whose transmittance BEFORE it is below T_thresh
// (raymarching.cu:885).  T = 1 - weights_sum is carried as K10 does; inside a window it is the running product.
__global__ void __launch_bounds__(128)
k_composite_window(const float* __restrict__ sigmas, const float* __restrict__ rgbs, const float* __restrict__ deltas,
                   const int32_t* __restrict__ rays_w, uint32_t M, uint32_t N, float T_thresh, float* __restrict__ rays_t,
                   float* __restrict__ weights_sum, float* __restrict__ depth, float* __restrict__ image, uint8_t* __restrict__ alive) {
    const int lane = threadIdx.x & 31;
    const uint32_t n = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (n >= N) return;
    const uint32_t off = (uint32_t)__ldg(rays_w + 3 * (size_t)n + 1), cnt = (uint32_t)__ldg(rays_w + 3 * (size_t)n + 2);
    if (cnt == 0 || (uint64_t)off + cnt > M) return;
    const float ws0 = weights_sum[n];
    float T = 1.0f - ws0, t_base = rays_t[n];
    float r = 0.f, g = 0.f, b = 0.f, d = 0.f, ws = 0.f;
    bool terminated = false;
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
        const float up = __shfl_up_sync(0xffffffffu, incl, 1);
        const float T_before = T * (lane == 0 ? 1.0f : up);
        const float w = alpha * T_before;
        const float t = t_base + warp_incl_sum(d1, lane);
        const uint32_t stop = __ballot_sync(0xffffffffu, valid && T_before < T_thresh);
        const int last = stop ? (__ffs(stop) - 1) : 31;
        if (valid && lane <= last) {
            r = fmaf(w, cr, r); g = fmaf(w, cg, g); b = fmaf(w, cb, b);
            d = fmaf(w, t, d); ws += w;
        }
        if (stop) { terminated = true; break; }
        T *= __shfl_sync(0xffffffffu, incl, 31); t_base = __shfl_sync(0xffffffffu, t, 31);
    }
    r = warp_sum(r); g = warp_sum(g); b = warp_sum(b); d = warp_sum(d); ws = warp_sum(ws);
    if (lane == 0) {
        weights_sum[n] = ws0 + ws; depth[n] += d;
        image[3 * (size_t)n] += r; image[3 * (size_t)n + 1] += g; image[3 * (size_t)n + 2] += b;
        if (terminated) alive[n] = 0; else rays_t[n] = t_base;
    }
}

// ------------------------------------------------------------------------------------------------ compaction
namespace {
constexpr uint32_t kCompactThreads = 256;
constexpr uint32_t kCompactItems = 8;                                // ids per thread
constexpr uint32_t kCompactTile = kCompactThreads * kCompactItems;   // 2048 ids per CTA
}

// pass 1: survivors per 2048-id tile
__global__ void __launch_bounds__(kCompactThreads)
k_compact_count(const int32_t* __restrict__ in, uint32_t n, uint32_t* __restrict__ tile_counts) {
    __shared__ uint32_t s_warp[kCompactThreads / 32];
    const uint32_t t0 = blockIdx.x * kCompactTile + threadIdx.x * kCompactItems;
    uint32_t c = 0;
    #pragma unroll
    for (uint32_t k = 0; k < kCompactItems; ++k) { const uint32_t i = t0 + k; c += (i < n && __ldg(in + i) >= 0); }
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tot = 0;
        #pragma unroll
        for (uint32_t w = 0; w < kCompactThreads / 32; ++w) tot += s_warp[w];
        tile_counts[blockIdx.x] = tot;
    }
}

// pass 2: every CTA sums the counts of the tiles before it (<= a few hundred values), then scatters its own
// survivors in order.  The last CTA also publishes the total.
__global__ void __launch_bounds__(kCompactThreads)
k_compact_scatter(const int32_t* __restrict__ in, uint32_t n, const uint32_t* __restrict__ tile_counts,
                  int32_t* __restrict__ out, int32_t* __restrict__ n_out) {
    __shared__ uint32_t s_warp[kCompactThreads / 32];
    __shared__ uint32_t s_base;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t part = 0;
    for (uint32_t b = threadIdx.x; b < blockIdx.x; b += kCompactThreads) part += __ldg(tile_counts + b);
    part = __reduce_add_sync(0xffffffffu, part);
    if (lane == 0) s_warp[wid] = part;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tot = 0;
        #pragma unroll
        for (uint32_t w = 0; w < kCompactThreads / 32; ++w) tot += s_warp[w];
        s_base = tot;
    }
    __syncthreads();
    const uint32_t tile_base = s_base;
    __syncthreads();

    const uint32_t t0 = blockIdx.x * kCompactTile + threadIdx.x * kCompactItems;
    int32_t v[kCompactItems];
    uint32_t c = 0;
    #pragma unroll
    for (uint32_t k = 0; k < kCompactItems; ++k) {
        const uint32_t i = t0 + k;
        v[k] = i < n ? __ldg(in + i) : -1;
        c += (v[k] >= 0);
    }
    const uint32_t incl = warp_incl_sum_u32(c, lane);
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    uint32_t woff = 0;
    #pragma unroll
    for (uint32_t w = 0; w < kCompactThreads / 32; ++w) woff += (w < (uint32_t)wid) ? s_warp[w] : 0u;
    uint32_t pos = tile_base + woff + incl - c;
    #pragma unroll
    for (uint32_t k = 0; k < kCompactItems; ++k) if (v[k] >= 0) out[pos++] = v[k];
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == kCompactThreads - 1) *n_out = (int32_t)pos;
}

// ------------------------------------------------------------------------------------------------ C ABI
static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

FOC_API int foc_composite_rays(uint32_t n_alive, uint32_t n_step, float T_thresh, int32_t* rays_alive, float* rays_t,
                               const float* sigmas, const float* rgbs, const float* deltas, float* weights_sum,
                               float* depth, float* image, void* stream) {
    if (n_alive == 0) return 0;
    k_composite_rays<<<ceil_div(n_alive, 128), 128, 0, S(stream)>>>(n_alive, n_step, T_thresh, rays_alive, rays_t, sigmas,
                                                                  rgbs, deltas, weights_sum, depth, image);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_composite_window(const float* sigmas, const float* rgbs, const float* deltas, const int32_t* rays_w, uint32_t M, uint32_t N,
                                 float T_thresh, float* rays_t, float* weights_sum, float* depth, float* image, uint8_t* alive, void* stream) {
    if (N == 0) return 0;
    k_composite_window<<<ceil_div(N, 4), 128, 0, S(stream)>>>(sigmas, rgbs, deltas, rays_w, M, N, T_thresh, rays_t, weights_sum, depth, image, alive);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API size_t foc_compact_alive_scratch_bytes(uint32_t n) { return (size_t)(ceil_div(n, kCompactTile) + 1) * sizeof(uint32_t); }

FOC_API int foc_compact_alive(const int32_t* rays_alive, uint32_t n, int32_t* out, int32_t* n_out, void* scratch, void* stream) {
    if (n == 0) return (int)cudaMemsetAsync(n_out, 0, sizeof(int32_t), S(stream));
    const uint32_t tiles = ceil_div(n, kCompactTile);
    uint32_t* tile_counts = static_cast<uint32_t*>(scratch);
    k_compact_count<<<tiles, kCompactThreads, 0, S(stream)>>>(rays_alive, n, tile_counts);
    FOC_LAUNCH_CHECK();
    k_compact_scatter<<<tiles, kCompactThreads, 0, S(stream)>>>(rays_alive, n, tile_counts, out, n_out);
    FOC_LAUNCH_CHECK(); return 0;
}
