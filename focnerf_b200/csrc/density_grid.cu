// density_grid.cu -- the density-grid refresh that feeds the marcher's bitfield.
//
// Reference: NeRFRenderer.update_extra_state, legacy/nerf/renderer.py:445-531 -- Python loops over 128^3 cells in S^3 blocks
// x cascades (meshgrid, morton3D, jitter, density(), scatter into tmp_grid), then a masked EMA-max, `.item()` of the mean
// and packbits with a host threshold.  Here the cells are generated directly in Morton order on the device, and the
// EMA-max, the mean and the thresholding stay on the device (no host read of the mean):
//   k_cell_positions : cell index (Morton) -> jittered sample position of that cell in its cascade     (l.470-479, 503-509)
//   k_grid_ema       : grid = max(grid * decay, tmp) where both are valid (>= 0); per-CTA sums of max(grid, 0)  (l.522-524)
//   k_grid_mean      : fixed-order sum of the per-CTA partials -> mean                                    (l.524)
//   k_packbits_dev   : bit = grid > min(mean, density_thresh), threshold read from device memory          (l.529-530)
#include <cstdio>
#include <mutex>

#include "common.cuh"
#include "../../include/foc_b200.h"

using namespace foc;

// Built-in scratch of foc_density_grid_update_noalloc (file scope): one partial sum per k_grid_ema CTA of 2048 cells, 8192 floats =
// 32 KB on every GPU that loads the library.  In HOST code this name is only a stand-in for the symbol, not a device address: the
// address comes from cudaGetSymbolAddress (noalloc_partials() below), once per device.
constexpr uint32_t kNoallocPartials = FOC_NOALLOC_MAX_CELLS / 2048u;
static __device__ __align__(256) float g_noalloc_partials[kNoallocPartials];

namespace {

constexpr int kEmaThreads = 256, kEmaPerThread = 8;

// indices == nullptr: cell i of the cascade IS Morton index i (full update); else indices[i] (partial update)
__global__ void __launch_bounds__(256)
k_cell_positions(const int32_t* __restrict__ indices, uint32_t n, uint32_t H, float bound_c, const float* __restrict__ noise,
                 float* __restrict__ xyzs) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t m = indices ? (uint32_t)__ldg(indices + i) : i;
    const uint32_t c[3] = {compact3(m), compact3(m >> 1), compact3(m >> 2)};
    const float hgs = bound_c / (float)H;                              // half a cell (renderer.py:475)
    #pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float u = 2.0f * (float)c[d] / (float)(H - 1) - 1.0f;    // [-1, 1]
        xyzs[3 * (size_t)i + d] = u * (bound_c - hgs) + (__ldg(noise + 3 * (size_t)i + d) * 2.0f - 1.0f) * hgs;
    }
}

// tmp values are scattered to tmp_grid[indices] by the caller (or are the full grid when indices == nullptr)
__global__ void __launch_bounds__(kEmaThreads)
k_grid_ema(float* __restrict__ grid, const float* __restrict__ tmp, uint32_t n, float decay, float* __restrict__ partial) {
    __shared__ float s[kEmaThreads / 32];
    float acc = 0.f;
    const uint32_t base = blockIdx.x * (kEmaThreads * kEmaPerThread);
    #pragma unroll
    for (int k = 0; k < kEmaPerThread; ++k) {
        const uint32_t i = base + k * kEmaThreads + threadIdx.x;
        if (i < n) {
            float g = grid[i];
            const float t = __ldg(tmp + i);
            if (g >= 0.f && t >= 0.f) { g = fmaxf(g * decay, t); grid[i] = g; }
            acc += fmaxf(g, 0.f);                                      // -1 regions count as zero density (l.524)
        }
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        #pragma unroll
        for (int w = 0; w < kEmaThreads / 32; ++w) t += s[w];
        partial[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(1024) k_grid_mean(const float* __restrict__ partial, uint32_t n_partial, uint32_t n, float* __restrict__ mean) {
    __shared__ float s[32];
    float acc = 0.f;
    for (uint32_t i = threadIdx.x; i < n_partial; i += 1024) acc += __ldg(partial + i);
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        const float t = warp_sum(s[threadIdx.x]);
        if (threadIdx.x == 0) mean[0] = t / (float)n;
    }
}

__global__ void __launch_bounds__(256)
k_packbits_dev(const float* __restrict__ grid, uint32_t N, const float* __restrict__ mean, float density_thresh, uint8_t* __restrict__ bitfield) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float thresh = fminf(__ldg(mean), density_thresh);
    const float4 a = __ldg(reinterpret_cast<const float4*>(grid) + (size_t)n * 2), b = __ldg(reinterpret_cast<const float4*>(grid) + (size_t)n * 2 + 1);
    bitfield[n] = (uint8_t)((a.x > thresh) | (a.y > thresh) << 1 | (a.z > thresh) << 2 | (a.w > thresh) << 3 |
                            (b.x > thresh) << 4 | (b.y > thresh) << 5 | (b.z > thresh) << 6 | (b.w > thresh) << 7);
}

static_assert(kNoallocPartials == FOC_NOALLOC_MAX_CELLS / (uint32_t)(kEmaThreads * kEmaPerThread), "one slot per k_grid_ema CTA");

constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_addr[kMaxDevices] = {};                              // per device, filled on the first call made on that device

// device address of g_noalloc_partials on the CURRENT device (looked up once per device); nullptr + *rc on failure
float* noalloc_partials(int* rc) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { *rc = (int)e; return nullptr; }
    if (dev < 0 || dev >= kMaxDevices) { *rc = FOC_ERR_UNSUPPORTED; return nullptr; }
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_addr[dev] == nullptr) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_partials);
        if (e != cudaSuccess || p == nullptr) { *rc = e != cudaSuccess ? (int)e : (int)cudaErrorInvalidSymbol; return nullptr; }
        g_noalloc_addr[dev] = static_cast<float*>(p);
    }
    *rc = 0;
    return g_noalloc_addr[dev];
}

}  // namespace

// NeRFRenderer.mark_untrained_grid (legacy/nerf/renderer.py:380-443): a cell no training camera sees gets density -1 and is left out of
// the EMA and the bitfield for good.  The reference walks the grid in 64^3 blocks x cascades x pose batches of 64 from Python (meshgrid,
// morton3D, a batched matmul and three masks per block); here one thread owns one cell (Morton index m of cascade `cas`) and tests it
// against the poses (c2w [B,4,4] row-major, staged through shared memory) until one sees it:
//     x_world = (2 c / (H-1) - 1) (bound_c - hgs),  hgs = bound_c / H                                  (l.406-413)
//     x_cam   = (x_world - t) @ R                    = R^T (x_world - t) for the c2w rotation R, translation t   (l.420-421)
//     seen    = z > 0  and  |x| < cx/fx z + 2 hgs  and  |y| < cy/fy z + 2 hgs                               (l.424-426)
__global__ void __launch_bounds__(256)
k_mark_untrained(const float* __restrict__ poses, uint32_t n_poses, float fx, float fy, float cx, float cy, uint32_t H, uint32_t cascade,
                 float bound, float* __restrict__ grid, uint32_t* __restrict__ n_marked) {
    extern __shared__ float s_pose[];                                   // [chunk, 12]: R row-major (9) + t (3)
    const uint32_t n = H * H * H;
    const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = gid < (uint64_t)cascade * n;
    const uint32_t cas = live ? (uint32_t)(gid / n) : 0u, m = live ? (uint32_t)(gid - (uint64_t)cas * n) : 0u;
    const float bound_c = fminf((float)(1u << cas), bound), hgs = bound_c / (float)H;
    float xw[3];
    {
        const uint32_t c[3] = {compact3(m), compact3(m >> 1), compact3(m >> 2)};
        #pragma unroll
        for (int d = 0; d < 3; ++d) xw[d] = (2.0f * (float)c[d] / (float)(H - 1) - 1.0f) * (bound_c - hgs);
    }
    const float kx = cx / fx, ky = cy / fy, slack = hgs * 2.0f;
    bool seen = false;
    constexpr uint32_t kChunk = 64;                                      // the reference's pose batch (S = 64)
    for (uint32_t p0 = 0; p0 < n_poses; p0 += kChunk) {
        const uint32_t np = min(kChunk, n_poses - p0);
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < np * 12; i += blockDim.x) {
            const uint32_t p = i / 12, e = i - p * 12;
            s_pose[i] = e < 9 ? __ldg(poses + (size_t)(p0 + p) * 16 + (e / 3) * 4 + (e % 3)) : __ldg(poses + (size_t)(p0 + p) * 16 + (e - 9) * 4 + 3);
        }
        __syncthreads();
        if (live && !seen) {
            for (uint32_t p = 0; p < np; ++p) {
                const float* P = s_pose + p * 12;
                const float dx = xw[0] - P[9], dy = xw[1] - P[10], dz = xw[2] - P[11];
                const float z = dx * P[2] + dy * P[5] + dz * P[8];
                if (!(z > 0.0f)) continue;
                const float x = dx * P[0] + dy * P[3] + dz * P[6], y = dx * P[1] + dy * P[4] + dz * P[7];
                if (fabsf(x) < kx * z + slack && fabsf(y) < ky * z + slack) { seen = true; break; }
            }
        }
    }
    if (live && !seen) {
        grid[gid] = -1.0f;
        if (n_marked) atomicAdd(n_marked, 1u);
    }
}

static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// density_grid [cascade, H^3] (Morton order): cells outside every camera frustum := -1.  n_marked (optional, device, zeroed by the caller)
// receives how many.
FOC_API int foc_mark_untrained_grid(const float* poses, uint32_t n_poses, float fx, float fy, float cx, float cy, uint32_t H, uint32_t cascade,
                                    float bound, float* density_grid, uint32_t* n_marked, void* stream) {
    if (H < 2 || H > 1024 || cascade == 0 || cascade > 8 || !(fx > 0.f) || !(fy > 0.f)) return FOC_ERR_UNSUPPORTED;
    const uint64_t total = (uint64_t)cascade * H * H * H;
    k_mark_untrained<<<(uint32_t)((total + 255) / 256), 256, 64 * 12 * sizeof(float), S(stream)>>>(poses, n_poses, fx, fy, cx, cy, H, cascade, bound,
                                                                                                   density_grid, n_marked);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_density_cell_positions(const int32_t* indices, uint32_t n, uint32_t H, float bound_c, const float* noise, float* xyzs,
                                       void* stream) {
    if (n == 0) return 0;
    if (H < 2 || H > 1024) return FOC_ERR_UNSUPPORTED;
    k_cell_positions<<<ceil_div(n, 256), 256, 0, S(stream)>>>(indices, n, H, bound_c, noise, xyzs);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API size_t foc_density_grid_update_scratch_bytes(uint32_t n) {
    return (size_t)ceil_div(n, (uint32_t)(kEmaThreads * kEmaPerThread)) * sizeof(float);
}

// grid [n] (all cascades), tmp [n] (-1 where not sampled); writes mean[0] and bitfield [n / 8]
FOC_API int foc_density_grid_update(float* grid, const float* tmp, uint32_t n, float decay, float density_thresh, float* mean,
                                    uint8_t* bitfield, void* scratch, void* stream) {
    if (n == 0) return 0;
    if (n % 8 != 0 || (reinterpret_cast<uintptr_t>(grid) & 15u)) return FOC_ERR_UNSUPPORTED;
    const uint32_t blocks = ceil_div(n, (uint32_t)(kEmaThreads * kEmaPerThread));
    float* partial = static_cast<float*>(scratch);
    k_grid_ema<<<blocks, kEmaThreads, 0, S(stream)>>>(grid, tmp, n, decay, partial);
    FOC_LAUNCH_CHECK();
    k_grid_mean<<<1, 1024, 0, S(stream)>>>(partial, blocks, n, mean);
    FOC_LAUNCH_CHECK();
    k_packbits_dev<<<ceil_div(n / 8, 256), 256, 0, S(stream)>>>(grid, n / 8, mean, density_thresh, bitfield);
    FOC_LAUNCH_CHECK(); return 0;
}

// foc_density_grid_update with the partial sums in the library's own static device array: allocates nothing (see include/foc_b200.h).
// k_grid_ema writes partial[b] for every CTA b < blocks and k_grid_mean reads exactly those, so the array's stale contents from an
// earlier (possibly larger) call are never read; overlapping calls on one device would race on it.
FOC_API int foc_density_grid_update_noalloc(float* grid, const float* tmp, uint32_t n, float decay, float density_thresh, float* mean,
                                            uint8_t* bitfield, void* stream) {
    if (n == 0) return 0;
    if (n > FOC_NOALLOC_MAX_CELLS) {
        fprintf(stderr, "foc_density_grid_update_noalloc: n = %u cells exceeds the built-in scratch (FOC_NOALLOC_MAX_CELLS = %u); "
                        "use foc_density_grid_update with a caller-owned scratch buffer\n", n, (unsigned)FOC_NOALLOC_MAX_CELLS);
        return FOC_ERR_TOO_LARGE;
    }
    if (n % 8 != 0 || (reinterpret_cast<uintptr_t>(grid) & 15u)) return FOC_ERR_UNSUPPORTED;
    int rc = 0;
    float* partial = noalloc_partials(&rc);
    if (partial == nullptr) return rc;
    const uint32_t blocks = ceil_div(n, (uint32_t)(kEmaThreads * kEmaPerThread));
    k_grid_ema<<<blocks, kEmaThreads, 0, S(stream)>>>(grid, tmp, n, decay, partial);
    FOC_LAUNCH_CHECK();
    k_grid_mean<<<1, 1024, 0, S(stream)>>>(partial, blocks, n, mean);
    FOC_LAUNCH_CHECK();
    k_packbits_dev<<<ceil_div(n / 8, 256), 256, 0, S(stream)>>>(grid, n / 8, mean, density_thresh, bitfield);
    FOC_LAUNCH_CHECK(); return 0;
}
