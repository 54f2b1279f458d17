// hashgrid_bwd.cu -- hash-grid backward (K12-K14: gridencoder/src/gridencoder.cu:248-369, 506-610).
//
// K12 is split by level: coarse levels (cells larger than the spacing of consecutive samples) aggregate per thread
// over runs of consecutive samples before touching L2 (k_grid_bwd_runs); fine levels scatter directly
// (k_grid_bwd_scatter).  Both use VECTOR reductions that return nothing (`red.global.add.noftz.f16x2` for the fp16 C=2
// table, `red.global.add.v2.f32` for fp32), so the L2 atomic units do the read-modify-write and the SM never waits.
// Order of accumulation is arbitrary, as in the reference: parity for this kernel is tolerance-based.
#include <cstdlib>
#include "hashgrid.cuh"

using namespace foc;

namespace {

// Reductions WITHOUT a return value, on the global window: the CUDA atomicAdd() overloads compile to a generic-address
// ATOM (result routed back to the SM, plus a shared-memory CAS branch); `red.global` is fire-and-forget at the L2.
__device__ __forceinline__ void red_f16x2(__half* p, float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    asm volatile("red.global.add.noftz.f16x2 [%0], %1;" :: "l"(__cvta_generic_to_global(p)), "r"(*reinterpret_cast<const uint32_t*>(&h)) : "memory");
}
__device__ __forceinline__ void red_f16x4(__half* p, float a, float b, float c, float d) {
    const __half2 h0 = __floats2half2_rn(a, b), h1 = __floats2half2_rn(c, d);
    asm volatile("red.global.add.noftz.v2.f16x2 [%0], {%1, %2};" :: "l"(__cvta_generic_to_global(p)),
                 "r"(*reinterpret_cast<const uint32_t*>(&h0)), "r"(*reinterpret_cast<const uint32_t*>(&h1)) : "memory");
}
__device__ __forceinline__ void red_f16(__half* p, float a) {
    const __half h = __float2half_rn(a);
    asm volatile("red.global.add.noftz.f16 [%0], %1;" :: "l"(__cvta_generic_to_global(p)), "h"(*reinterpret_cast<const uint16_t*>(&h)) : "memory");
}
__device__ __forceinline__ void red_f32(float* p, float a) {
    asm volatile("red.global.add.f32 [%0], %1;" :: "l"(__cvta_generic_to_global(p)), "f"(a) : "memory");
}
__device__ __forceinline__ void red_f32x2(float* p, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" :: "l"(__cvta_generic_to_global(p)), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void red_f32x4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(__cvta_generic_to_global(p)), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <typename T, int C>
__device__ __forceinline__ void red_row(T* p, const float (&v)[C]) {
    if constexpr (sizeof(T) == 2) {
        __half* h = reinterpret_cast<__half*>(p);
        if constexpr (C % 4 == 0) {
            #pragma unroll
            for (int c = 0; c < C; c += 4) red_f16x4(h + c, v[c], v[c + 1], v[c + 2], v[c + 3]);
        } else if constexpr (C % 2 == 0) {
            #pragma unroll
            for (int c = 0; c < C; c += 2) red_f16x2(h + c, v[c], v[c + 1]);
        } else {
            #pragma unroll
            for (int c = 0; c < C; ++c) red_f16(h + c, v[c]);
        }
    } else {
        float* f = reinterpret_cast<float*>(p);
        if constexpr (C % 4 == 0) {
            #pragma unroll
            for (int c = 0; c < C; c += 4) red_f32x4(f + c, v[c], v[c + 1], v[c + 2], v[c + 3]);
        } else if constexpr (C % 2 == 0) {
            #pragma unroll
            for (int c = 0; c < C; c += 2) red_f32x2(f + c, v[c], v[c + 1]);
        } else {
            #pragma unroll
            for (int c = 0; c < C; ++c) red_f32(f + c, v[c]);
        }
    }
}

// K12 for the FINE levels [level0, level0 + nl): one thread per (sample, level), consecutive threads = the levels of
// one sample ([B,L,C] gradients are then read contiguously).  Cells are smaller than the sample spacing here, every
// update is a different row: nothing to combine, the scatter goes straight to the L2 reduction units.
template <typename T, int D, int C, int LAYOUT>
__device__ __forceinline__ void scatter_body(uint64_t t, const T* __restrict__ grad, const float* __restrict__ inputs, const int32_t* __restrict__ offsets,
                   T* __restrict__ grad_table, uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners,
                   uint32_t interp, uint32_t level0, uint32_t nl) {
    if (t >= (uint64_t)B * nl) return;
    uint32_t b, level;
    split_index(t, nl, b, level);
    level += level0;
    const T* gp = LAYOUT == 0 ? grad + ((size_t)level * B + b) * C : grad + ((size_t)b * L + level) * C;
    float g[C];
    bool any = false;
    #pragma unroll
    for (int c = 0; c < C; ++c) { g[c] = Elem<T>::ld(gp + c); any |= (g[c] != 0.0f); }
    // A zero gradient adds nothing: skip its 2^D reductions.  This is where the padding rows of the sample buffer (all at
    // the origin, all zero-gradient) would otherwise pile tens of thousands of updates onto the same 2^D addresses per
    // level, which the L2 serialises.  (NaN != 0, so non-finite gradients still propagate for the GradScaler to see.)
    if (!any) return;
    float x[D];
    #pragma unroll
    for (int d = 0; d < D; ++d) x[d] = __ldg(inputs + (size_t)b * D + d);
    if (grid_oob<D>(x)) return;                                                   // gridencoder.cu:276-281
    const GridLevel lv = grid_level(offsets, level, S, H);
    const GridIndexer<D> gidx(gridtype, align_corners, lv.hashmap_size, lv.resolution);
    T* tab = grad_table + (size_t)lv.offset * C;
    float pos[D], pder[D];
    uint32_t pg[D];
    grid_locate<D>(x, lv.scale, align_corners, interp, pos, pder, pg);
    // corners come in pairs along x (idx, idx | 1).  Their rows are neighbours whenever the dense index is used, and on a
    // hashed level whenever x is even (x + 1 only flips bit 0 of x * 1, the first hash prime): an aligned neighbour pair of
    // the C = 2 fp16 table is 8 contiguous bytes, one vector reduction instead of two.
    #pragma unroll
    for (int idx = 0; idx < (1 << D); idx += 2) {
        float wt[2];
        uint32_t row[2];
        #pragma unroll
        for (int s = 0; s < 2; ++s) {
            float w = 1.0f;
            uint32_t pl[D];
            #pragma unroll
            for (int d = 0; d < D; ++d) {
                if (((idx + s) >> d) & 1) { w = __fmul_rn(w, pos[d]); pl[d] = pg[d] + 1; }
                else { w = __fmul_rn(w, __fsub_rn(1.0f, pos[d])); pl[d] = pg[d]; }
            }
            wt[s] = w; row[s] = gidx(pl);
        }
        float v0[C], v1[C];
        #pragma unroll
        for (int c = 0; c < C; ++c) { v0[c] = __fmul_rn(wt[0], g[c]); v1[c] = __fmul_rn(wt[1], g[c]); }
        if constexpr (sizeof(T) == 2 && C == 2) {
            if (row[1] == row[0] + 1 && (row[0] & 1u) == 0) { red_f16x4(reinterpret_cast<__half*>(tab) + (size_t)row[0] * 2, v0[0], v0[1], v1[0], v1[1]); continue; }
        }
        red_row<T, C>(tab + (size_t)row[0] * C, v0);
        red_row<T, C>(tab + (size_t)row[1] * C, v1);
    }
}

template <typename T, int D, int C, int LAYOUT>
__global__ void __launch_bounds__(256)
k_grid_bwd_scatter(const T* __restrict__ grad, const float* __restrict__ inputs, const int32_t* __restrict__ offsets,
                   T* __restrict__ grad_table, uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners,
                   uint32_t interp, uint32_t level0, uint32_t nl) {
    pdl_sync();
    scatter_body<T, D, C, LAYOUT>((uint64_t)blockIdx.x * blockDim.x + threadIdx.x, grad, inputs, offsets, grad_table, B, L, S, H, gridtype,
                                  align_corners, interp, level0, nl);
}

// K12 for the COARSE levels: one thread walks kRun consecutive samples of one level and keeps the 2^D corner sums of the
// current cell in registers, issuing reductions only when the cell changes.  Samples arrive in ray order (the march
// writes each ray's samples contiguously, ~0.0017 apart in the unit cube), so at resolution 16..64 a thread sees 1-3
// cells per run: ~10x fewer L2 reductions, and -- what matters -- ~10x fewer updates to the few hundred hot rows the
// object occupies on those levels, which the L2 atomic units serialise per address.
constexpr int kRun = 16;

template <typename T, int D, int C, int LAYOUT>
__device__ __forceinline__ void runs_body(uint32_t level, uint32_t run, const T* __restrict__ grad, const float* __restrict__ inputs,
                const int32_t* __restrict__ offsets, T* __restrict__ grad_table, uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype,
                bool align_corners, uint32_t interp) {
    const uint32_t b0 = run * (uint32_t)kRun;
    if (b0 >= B) return;
    const GridLevel lv = grid_level(offsets, level, S, H);
    const GridIndexer<D> gidx(gridtype, align_corners, lv.hashmap_size, lv.resolution);
    T* tab = grad_table + (size_t)lv.offset * C;
    float acc[1 << D][C];
    uint32_t cur[D];
    bool have = false;
    auto flush = [&]() {
        #pragma unroll
        for (int idx = 0; idx < (1 << D); idx += 2) {                // x-neighbour pairs, see k_grid_bwd_scatter
            uint32_t row[2];
            #pragma unroll
            for (int s = 0; s < 2; ++s) {
                uint32_t pl[D];
                #pragma unroll
                for (int d = 0; d < D; ++d) pl[d] = cur[d] + (((idx + s) >> d) & 1);
                row[s] = gidx(pl);
            }
            if constexpr (sizeof(T) == 2 && C == 2) {
                if (row[1] == row[0] + 1 && (row[0] & 1u) == 0) {
                    red_f16x4(reinterpret_cast<__half*>(tab) + (size_t)row[0] * 2, acc[idx][0], acc[idx][1], acc[idx + 1][0], acc[idx + 1][1]);
                    continue;
                }
            }
            red_row<T, C>(tab + (size_t)row[0] * C, acc[idx]);
            red_row<T, C>(tab + (size_t)row[1] * C, acc[idx + 1]);
        }
    };
    const uint32_t n = min((uint32_t)kRun, B - b0);
    #pragma unroll 1
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t b = b0 + i;
        float x[D];
        #pragma unroll
        for (int d = 0; d < D; ++d) x[d] = __ldg(inputs + (size_t)b * D + d);
        if (grid_oob<D>(x)) continue;                                             // gridencoder.cu:276-281
        const T* gp = LAYOUT == 0 ? grad + ((size_t)level * B + b) * C : grad + ((size_t)b * L + level) * C;
        float g[C];
        bool any = false;
        #pragma unroll
        for (int c = 0; c < C; ++c) { g[c] = Elem<T>::ld(gp + c); any |= (g[c] != 0.0f); }
        if (!any) continue;                                                       // see k_grid_bwd
        float pos[D], pder[D];
        uint32_t pg[D];
        grid_locate<D>(x, lv.scale, align_corners, interp, pos, pder, pg);
        bool same = have;
        #pragma unroll
        for (int d = 0; d < D; ++d) same = same && (pg[d] == cur[d]);
        if (!same) {
            if (have) flush();
            #pragma unroll
            for (int d = 0; d < D; ++d) cur[d] = pg[d];
            #pragma unroll
            for (int idx = 0; idx < (1 << D); ++idx) {
                #pragma unroll
                for (int c = 0; c < C; ++c) acc[idx][c] = 0.0f;
            }
            have = true;
        }
        #pragma unroll
        for (int idx = 0; idx < (1 << D); ++idx) {
            float wt = 1.0f;
            #pragma unroll
            for (int d = 0; d < D; ++d) wt = __fmul_rn(wt, ((idx >> d) & 1) ? pos[d] : __fsub_rn(1.0f, pos[d]));
            #pragma unroll
            for (int c = 0; c < C; ++c) acc[idx][c] = fmaf(wt, g[c], acc[idx][c]);
        }
    }
    if (have) flush();
}

// K13 (gridencoder.cu:343-369): grad_inputs[b,d] = sum_{l,c} grad[l,b,c] * dy_dx[b,l,d,c], running sum in T.
template <typename T, int LAYOUT>
__global__ void __launch_bounds__(256)
k_grid_input_bwd(const T* __restrict__ grad, const T* __restrict__ dy_dx, T* __restrict__ grad_inputs, uint32_t B,
                 uint32_t D, uint32_t C, uint32_t L) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (uint64_t)B * D) return;
    const uint32_t b = (uint32_t)(t / D), d = (uint32_t)(t % D);
    float acc = 0.f;
    for (uint32_t l = 0; l < L; ++l) {
        for (uint32_t c = 0; c < C; ++c) {
            const size_t gi = LAYOUT == 0 ? ((size_t)l * B + b) * C + c : ((size_t)b * L + l) * C + c;
            const size_t di = (((size_t)b * L + l) * D + d) * C + c;
            const float gv = Elem<T>::ld(grad + gi), dv = Elem<T>::ld(dy_dx + di);
            if constexpr (sizeof(T) == 2) acc = Elem<T>::rnd(__fadd_rn(acc, Elem<T>::rnd(__fmul_rn(gv, dv))));
            else acc = fmaf(gv, dv, acc);
        }
    }
    Elem<T>::st(grad_inputs + t, acc);
}

// K14 (gridencoder.cu:506-610): normalised total-variation gradient at the cell of each input point; fp32 only.
template <int D, int C>
__global__ void __launch_bounds__(256)
k_grad_tv(const float* __restrict__ inputs, const float* __restrict__ table, float* __restrict__ grad,
          const int32_t* __restrict__ offsets, float weight, uint32_t B, uint32_t L, float S, uint32_t H,
          uint32_t gridtype, bool align_corners) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x, level = blockIdx.y;
    if (b >= B) return;
    float x[D];
    #pragma unroll
    for (int d = 0; d < D; ++d) x[d] = __ldg(inputs + (size_t)b * D + d);
    if (grid_oob<D>(x)) return;
    const GridLevel lv = grid_level(offsets, level, S, H);
    const GridIndexer<D> gidx(gridtype, align_corners, lv.hashmap_size, lv.resolution);
    const float* tab = table + (size_t)lv.offset * C;
    uint32_t pg[D];
    #pragma unroll
    for (int d = 0; d < D; ++d) pg[d] = (uint32_t)floorf(fmaf(x[d], lv.scale, align_corners ? 0.0f : 0.5f));
    const uint32_t row0 = gidx(pg);
    float c0[C], res[C], idelta[C];
    #pragma unroll
    for (int c = 0; c < C; ++c) { c0[c] = __ldg(tab + (size_t)row0 * C + c); res[c] = 0.f; idelta[c] = 0.f; }
    const float w = weight / (float)(2 * D);
    #pragma unroll
    for (int d = 0; d < D; ++d) {
        const uint32_t cur = pg[d];
        if (cur < lv.resolution) {
            pg[d] = cur + 1;
            const uint32_t r = gidx(pg);
            #pragma unroll
            for (int c = 0; c < C; ++c) { const float gdiff = c0[c] - __ldg(tab + (size_t)r * C + c); res[c] += gdiff; idelta[c] = fmaf(gdiff, gdiff, idelta[c]); }
        }
        if (cur > 0) {
            pg[d] = cur - 1;
            const uint32_t r = gidx(pg);
            #pragma unroll
            for (int c = 0; c < C; ++c) { const float gdiff = c0[c] - __ldg(tab + (size_t)r * C + c); res[c] += gdiff; idelta[c] = fmaf(gdiff, gdiff, idelta[c]); }
        }
        pg[d] = cur;
    }
    float v[C];
    #pragma unroll
    for (int c = 0; c < C; ++c) v[c] = (w * res[c]) * rsqrtf(idelta[c] + 1e-9f);
    red_row<float, C>(grad + ((size_t)lv.offset + row0) * C, v);
}

template <typename T, int D, int C, int LAYOUT>
__global__ void __launch_bounds__(128)
k_grid_bwd_runs(const T* __restrict__ grad, const float* __restrict__ inputs, const int32_t* __restrict__ offsets,
                T* __restrict__ grad_table, uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners,
                uint32_t interp) {
    pdl_sync();
    runs_body<T, D, C, LAYOUT>(blockIdx.y, blockIdx.x * blockDim.x + threadIdx.x, grad, inputs, offsets, grad_table, B, L, S, H, gridtype,
                               align_corners, interp);
}

// Both halves in ONE launch: the first `run_ctas` CTAs walk the coarse levels' runs (few, latency-bound threads: 21 % occupancy
// alone), the rest scatter the fine levels (bound by the L2 reduction units).  Launched one after the other they cost 30 + 50 us;
// side by side the runs hide under the scatter.  The levels are disjoint table slices, so no ordering is needed between them.
template <typename T, int D, int C, int LAYOUT>
__global__ void __launch_bounds__(256, 4)
k_grid_bwd_fused(const T* __restrict__ grad, const float* __restrict__ inputs, const int32_t* __restrict__ offsets,
                 T* __restrict__ grad_table, uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners,
                 uint32_t interp, uint32_t n_coarse, uint32_t run_ctas_per_level, const int32_t* __restrict__ n_rows) {
    pdl_sync();
    // n_rows (optional): valid leading rows, read from device memory (see hashgrid_fwd.cu); the CTA roles below are laid out for the
    // host-side B, threads whose samples lie beyond the device-side count find nothing to do
    if (n_rows != nullptr) { const uint32_t v = (uint32_t)max(__ldg(n_rows), 0); B = v < B ? v : B; }
    const uint32_t run_ctas = run_ctas_per_level * n_coarse;
    if (blockIdx.x < run_ctas) {
        const uint32_t level = blockIdx.x / run_ctas_per_level, blk = blockIdx.x - level * run_ctas_per_level;
        runs_body<T, D, C, LAYOUT>(level, blk * 256u + threadIdx.x, grad, inputs, offsets, grad_table, B, L, S, H, gridtype, align_corners, interp);
    } else {
        scatter_body<T, D, C, LAYOUT>((uint64_t)(blockIdx.x - run_ctas) * 256u + threadIdx.x, grad, inputs, offsets, grad_table, B, L, S, H, gridtype,
                                      align_corners, interp, n_coarse, L - n_coarse);
    }
}

// levels whose resolution is <= kCoarseRes go through the per-thread runs (measured on B200, 156k ray-ordered samples:
// 162 us all-scatter -> 89 us with the split at 128; 16 -> 148, 64 -> 97, 256 -> 89, 512 -> 93, all levels -> 127)
constexpr uint32_t kCoarseRes = 128;

template <typename T, int D, int C, int LAYOUT>
int launch_bwd_layout(const void* grad, const float* inputs, const int32_t* offsets, void* grad_table, uint32_t B, uint32_t L,
                      float S, uint32_t H, uint32_t gridtype, bool ac, uint32_t interp, cudaStream_t st, const int32_t* n_rows = nullptr) {
    uint32_t n_coarse = 0;                                           // same level resolution rule as grid_level()
    while (n_coarse < L && (uint32_t)ceilf(exp2f((float)n_coarse * S) * (float)H - 1.0f) + 1u <= kCoarseRes) ++n_coarse;
    if (n_rows != nullptr && !(n_coarse > 0 && n_coarse < L)) return FOC_ERR_UNSUPPORTED;      // device-side row count: fused kernel only
    if (n_coarse > 0 && n_coarse < L && (n_rows != nullptr || std::getenv("FOC_GRID_BWD_SPLIT") == nullptr)) {
        const uint32_t rpl = ceil_div(ceil_div(B, (uint32_t)kRun), 256u);
        const uint64_t sc = ((uint64_t)B * (L - n_coarse) + 255) / 256;
        foc::launch(k_grid_bwd_fused<T, D, C, LAYOUT>, (uint32_t)(rpl * n_coarse + sc), 256, 0, st, (const T*)grad, inputs, offsets, (T*)grad_table, B, L, S, H,
                    gridtype, ac, interp, n_coarse, rpl, n_rows);
        FOC_LAUNCH_CHECK();
        return 0;
    }
    if (n_coarse > 0) {
        dim3 rgrid(ceil_div(ceil_div(B, (uint32_t)kRun), 128u), n_coarse);
        foc::launch(k_grid_bwd_runs<T, D, C, LAYOUT>, rgrid, 128, 0, st, (const T*)grad, inputs, offsets, (T*)grad_table, B, L, S, H, gridtype, ac, interp);
        FOC_LAUNCH_CHECK();
    }
    if (n_coarse < L) {
        const uint32_t nl = L - n_coarse;
        const uint64_t total = (uint64_t)B * nl;
        foc::launch(k_grid_bwd_scatter<T, D, C, LAYOUT>, (uint32_t)((total + 255) / 256), 256, 0, st, (const T*)grad, inputs, offsets, (T*)grad_table, B, L, S, H,
                                                                                          gridtype, ac, interp, n_coarse, nl);
        FOC_LAUNCH_CHECK();
    }
    return 0;
}
template <typename T, int D, int C>
int launch_bwd(const void* grad, const float* inputs, const int32_t* offsets, void* grad_table, uint32_t B, uint32_t L,
               float S, uint32_t H, uint32_t gridtype, bool ac, uint32_t interp, int layout, cudaStream_t st, const int32_t* n_rows = nullptr) {
    return layout == 0 ? launch_bwd_layout<T, D, C, 0>(grad, inputs, offsets, grad_table, B, L, S, H, gridtype, ac, interp, st, n_rows)
                       : launch_bwd_layout<T, D, C, 1>(grad, inputs, offsets, grad_table, B, L, S, H, gridtype, ac, interp, st, n_rows);
}
template <typename T, int D>
int dispatch_bwd(uint32_t C, const void* grad, const float* inputs, const int32_t* offsets, void* grad_table, uint32_t B,
                 uint32_t L, float S, uint32_t H, uint32_t gridtype, bool ac, uint32_t interp, int layout, cudaStream_t st,
                 const int32_t* n_rows = nullptr) {
    switch (C) {
        case 1: return launch_bwd<T, D, 1>(grad, inputs, offsets, grad_table, B, L, S, H, gridtype, ac, interp, layout, st, n_rows);
        case 2: return launch_bwd<T, D, 2>(grad, inputs, offsets, grad_table, B, L, S, H, gridtype, ac, interp, layout, st, n_rows);
        case 4: return launch_bwd<T, D, 4>(grad, inputs, offsets, grad_table, B, L, S, H, gridtype, ac, interp, layout, st, n_rows);
        case 8: return launch_bwd<T, D, 8>(grad, inputs, offsets, grad_table, B, L, S, H, gridtype, ac, interp, layout, st, n_rows);
        default: return FOC_ERR_UNSUPPORTED;
    }
}

}  // namespace

FOC_API int foc_grid_encode_backward(const void* grad, const float* inputs, const void* embeddings, const int32_t* offsets,
                                     void* grad_embeddings, uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S,
                                     uint32_t H, const void* dy_dx, void* grad_inputs, uint32_t gridtype, int align_corners,
                                     uint32_t interp, int dtype, int layout, void* stream) {
    (void)embeddings;
    if (B == 0 || L == 0) return 0;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const bool ac = align_corners != 0;
    int rc = FOC_ERR_UNSUPPORTED;
    if (dtype == 1) {
        if (D == 2) rc = dispatch_bwd<__half, 2>(C, grad, inputs, offsets, grad_embeddings, B, L, S, H, gridtype, ac, interp, layout, st);
        if (D == 3) rc = dispatch_bwd<__half, 3>(C, grad, inputs, offsets, grad_embeddings, B, L, S, H, gridtype, ac, interp, layout, st);
    } else if (dtype == 0) {
        if (D == 2) rc = dispatch_bwd<float, 2>(C, grad, inputs, offsets, grad_embeddings, B, L, S, H, gridtype, ac, interp, layout, st);
        if (D == 3) rc = dispatch_bwd<float, 3>(C, grad, inputs, offsets, grad_embeddings, B, L, S, H, gridtype, ac, interp, layout, st);
    }
    if (rc != 0) return rc;
    if (dy_dx && grad_inputs) {
        const uint64_t total = (uint64_t)B * D;
        const uint32_t blocks = (uint32_t)((total + 255) / 256);
        if (dtype == 1) {
            if (layout == 0) k_grid_input_bwd<__half, 0><<<blocks, 256, 0, st>>>((const __half*)grad, (const __half*)dy_dx, (__half*)grad_inputs, B, D, C, L);
            else k_grid_input_bwd<__half, 1><<<blocks, 256, 0, st>>>((const __half*)grad, (const __half*)dy_dx, (__half*)grad_inputs, B, D, C, L);
        } else {
            if (layout == 0) k_grid_input_bwd<float, 0><<<blocks, 256, 0, st>>>((const float*)grad, (const float*)dy_dx, (float*)grad_inputs, B, D, C, L);
            else k_grid_input_bwd<float, 1><<<blocks, 256, 0, st>>>((const float*)grad, (const float*)dy_dx, (float*)grad_inputs, B, D, C, L);
        }
        FOC_LAUNCH_CHECK();
    }
    return 0;
}

// foc_grid_encode_backward over the first *n_rows rows only (device-side count, see foc_grid_encode_forward_rows); point-major layout,
// no input gradients.
FOC_API int foc_grid_encode_backward_rows(const void* grad, const float* inputs, const int32_t* offsets, void* grad_embeddings, uint32_t B,
                                          uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, uint32_t gridtype, int align_corners,
                                          uint32_t interp, int dtype, const int32_t* n_rows, void* stream) {
    if (B == 0 || L == 0) return 0;
    if (n_rows == nullptr || D != 3) return FOC_ERR_UNSUPPORTED;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const bool ac = align_corners != 0;
    if (dtype == 1) return dispatch_bwd<__half, 3>(C, grad, inputs, offsets, grad_embeddings, B, L, S, H, gridtype, ac, interp, 1, st, n_rows);
    if (dtype == 0) return dispatch_bwd<float, 3>(C, grad, inputs, offsets, grad_embeddings, B, L, S, H, gridtype, ac, interp, 1, st, n_rows);
    return FOC_ERR_UNSUPPORTED;
}

FOC_API int foc_grad_total_variation(const float* inputs, const float* embeddings, float* grad, const int32_t* offsets,
                                     float weight, uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H,
                                     uint32_t gridtype, int align_corners, void* stream) {
    if (B == 0 || L == 0) return 0;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    dim3 grid(ceil_div(B, 256), L);
    const bool ac = align_corners != 0;
#define FOC_TV(DD, CC) k_grad_tv<DD, CC><<<grid, 256, 0, st>>>(inputs, embeddings, grad, offsets, weight, B, L, S, H, gridtype, ac)
    if (D == 3 && C == 1) FOC_TV(3, 1); else if (D == 3 && C == 2) FOC_TV(3, 2); else if (D == 3 && C == 4) FOC_TV(3, 4);
    else if (D == 3 && C == 8) FOC_TV(3, 8); else if (D == 2 && C == 1) FOC_TV(2, 1); else if (D == 2 && C == 2) FOC_TV(2, 2);
    else if (D == 2 && C == 4) FOC_TV(2, 4); else if (D == 2 && C == 8) FOC_TV(2, 8); else return FOC_ERR_UNSUPPORTED;
#undef FOC_TV
    FOC_LAUNCH_CHECK(); return 0;
}
