// hashgrid_bwd.cu -- hash-grid backward (K12-K14: gridencoder/src/gridencoder.cu:248-369, 506-610).
//
// K12 scatter: one thread per (point, level) adds w * grad to the 2^D corner rows with VECTOR reductions that return
// nothing (`red.global.add.noftz.f16x2` for the fp16 C=2 table, `red.global.add.v2.f32` for fp32), so the L2 atomic
// units do the read-modify-write and the SM never waits for a result.  Threads of a warp are consecutive samples of
// the same level (layout 0) -- along a ray these fall into the same coarse cell, so equal target rows inside a warp
// are first combined with __match_any_sync and one leader lane issues the reduction (warp-aggregated atomics).
// Order of accumulation is arbitrary, as in the reference: parity for this kernel is tolerance-based.
#include "hashgrid.cuh"

using namespace foc;

namespace {

template <typename T, int C>
__device__ __forceinline__ void red_row(T* p, const float (&v)[C]) {
    if constexpr (sizeof(T) == 2) {
        if constexpr (C % 2 == 0) {
            #pragma unroll
            for (int c = 0; c < C; c += 2) atomicAdd(reinterpret_cast<__half2*>(p + c), __floats2half2_rn(v[c], v[c + 1]));
        } else {
            #pragma unroll
            for (int c = 0; c < C; ++c) atomicAdd(p + c, __float2half_rn(v[c]));
        }
    } else {
        if constexpr (C % 2 == 0) {
            #pragma unroll
            for (int c = 0; c < C; c += 2) atomicAdd(reinterpret_cast<float2*>(p + c), make_float2(v[c], v[c + 1]));
        } else {
            #pragma unroll
            for (int c = 0; c < C; ++c) atomicAdd(p + c, v[c]);
        }
    }
}

// Combine the contributions of lanes that target the same row; returns true on the lane that must issue the add.
template <int C>
__device__ __forceinline__ bool warp_aggregate(uint32_t active, uint32_t row, float (&v)[C]) {
    const uint32_t peers = __match_any_sync(active, row);
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(peers) - 1;
    if (peers == (1u << lane)) return true;                       // common case on fine levels: no duplicate
    // tree-free gather: the leader pulls every peer's value (peer sets are small except on the coarsest levels)
    uint32_t rest = peers & ~(1u << leader);
    float acc[C];
    #pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = v[c];
    // all lanes of a peer group walk the same `rest` bits, so the shuffles below are convergent within the group
    while (rest) {
        const int src = __ffs(rest) - 1;
        rest &= rest - 1;
        #pragma unroll
        for (int c = 0; c < C; ++c) acc[c] += __shfl_sync(peers, v[c], src);
    }
    #pragma unroll
    for (int c = 0; c < C; ++c) v[c] = acc[c];
    return lane == leader;
}

template <typename T, int D, int C, int LAYOUT, bool AGG>
__global__ void __launch_bounds__(256)
k_grid_bwd(const T* __restrict__ grad, const float* __restrict__ inputs, const int32_t* __restrict__ offsets,
           T* __restrict__ grad_table, uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners,
           uint32_t interp) {
    uint32_t b, level;
    bool live = true;
    if constexpr (LAYOUT == 0) { b = blockIdx.x * blockDim.x + threadIdx.x; level = blockIdx.y; live = b < B; }
    else {
        const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
        live = t < (uint64_t)B * L;
        b = (uint32_t)(t / L); level = (uint32_t)(t % L);
    }
    float x[D];
    if (live) {
        #pragma unroll
        for (int d = 0; d < D; ++d) x[d] = __ldg(inputs + (size_t)b * D + d);
        live = !grid_oob<D>(x);                                                   // gridencoder.cu:276-281
    }
    const uint32_t active = __ballot_sync(0xffffffffu, live);
    if (!live) return;
    const GridLevel lv = grid_level(offsets, level, S, H);
    T* tab = grad_table + (size_t)lv.offset * C;
    float pos[D], pder[D];
    uint32_t pg[D];
    grid_locate<D>(x, lv.scale, align_corners, interp, pos, pder, pg);
    const T* gp = LAYOUT == 0 ? grad + ((size_t)level * B + b) * C : grad + ((size_t)b * L + level) * C;
    float g[C];
    #pragma unroll
    for (int c = 0; c < C; ++c) g[c] = Elem<T>::ld(gp + c);
    #pragma unroll
    for (int idx = 0; idx < (1 << D); ++idx) {
        float wt = 1.0f;
        uint32_t pl[D];
        #pragma unroll
        for (int d = 0; d < D; ++d) {
            if ((idx >> d) & 1) { wt = __fmul_rn(wt, pos[d]); pl[d] = pg[d] + 1; }
            else { wt = __fmul_rn(wt, __fsub_rn(1.0f, pos[d])); pl[d] = pg[d]; }
        }
        const uint32_t row = grid_index<D>(gridtype, align_corners, lv.hashmap_size, lv.resolution, pl);
        float v[C];
        #pragma unroll
        for (int c = 0; c < C; ++c) v[c] = __fmul_rn(wt, g[c]);
        if constexpr (AGG) {
            if (warp_aggregate<C>(active, row, v)) red_row<T, C>(tab + (size_t)row * C, v);
        } else {
            red_row<T, C>(tab + (size_t)row * C, v);
        }
    }
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
    const float* tab = table + (size_t)lv.offset * C;
    uint32_t pg[D];
    #pragma unroll
    for (int d = 0; d < D; ++d) pg[d] = (uint32_t)floorf(fmaf(x[d], lv.scale, align_corners ? 0.0f : 0.5f));
    const uint32_t row0 = grid_index<D>(gridtype, align_corners, lv.hashmap_size, lv.resolution, pg);
    float c0[C], res[C], idelta[C];
    #pragma unroll
    for (int c = 0; c < C; ++c) { c0[c] = __ldg(tab + (size_t)row0 * C + c); res[c] = 0.f; idelta[c] = 0.f; }
    const float w = weight / (float)(2 * D);
    #pragma unroll
    for (int d = 0; d < D; ++d) {
        const uint32_t cur = pg[d];
        if (cur < lv.resolution) {
            pg[d] = cur + 1;
            const uint32_t r = grid_index<D>(gridtype, align_corners, lv.hashmap_size, lv.resolution, pg);
            #pragma unroll
            for (int c = 0; c < C; ++c) { const float gdiff = c0[c] - __ldg(tab + (size_t)r * C + c); res[c] += gdiff; idelta[c] = fmaf(gdiff, gdiff, idelta[c]); }
        }
        if (cur > 0) {
            pg[d] = cur - 1;
            const uint32_t r = grid_index<D>(gridtype, align_corners, lv.hashmap_size, lv.resolution, pg);
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

template <typename T, int D, int C>
int launch_bwd(const void* grad, const float* inputs, const int32_t* offsets, void* grad_table, uint32_t B, uint32_t L,
               float S, uint32_t H, uint32_t gridtype, bool ac, uint32_t interp, int layout, cudaStream_t st) {
    if (layout == 0) {
        dim3 grid(ceil_div(B, 256), L);
        k_grid_bwd<T, D, C, 0, true><<<grid, 256, 0, st>>>((const T*)grad, inputs, offsets, (T*)grad_table, B, L, S, H, gridtype, ac, interp);
    } else {
        const uint64_t total = (uint64_t)B * L;
        k_grid_bwd<T, D, C, 1, false><<<(uint32_t)((total + 255) / 256), 256, 0, st>>>((const T*)grad, inputs, offsets, (T*)grad_table, B, L, S, H, gridtype, ac, interp);
    }
    FOC_LAUNCH_CHECK(); return 0;
}
template <typename T, int D>
int dispatch_bwd(uint32_t C, const void* grad, const float* inputs, const int32_t* offsets, void* grad_table, uint32_t B,
                 uint32_t L, float S, uint32_t H, uint32_t gridtype, bool ac, uint32_t interp, int layout, cudaStream_t st) {
    switch (C) {
        case 1: return launch_bwd<T, D, 1>(grad, inputs, offsets, grad_table, B, L, S, H, gridtype, ac, interp, layout, st);
        case 2: return launch_bwd<T, D, 2>(grad, inputs, offsets, grad_table, B, L, S, H, gridtype, ac, interp, layout, st);
        case 4: return launch_bwd<T, D, 4>(grad, inputs, offsets, grad_table, B, L, S, H, gridtype, ac, interp, layout, st);
        case 8: return launch_bwd<T, D, 8>(grad, inputs, offsets, grad_table, B, L, S, H, gridtype, ac, interp, layout, st);
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
