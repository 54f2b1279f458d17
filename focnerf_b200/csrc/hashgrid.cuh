// hashgrid.cuh -- shared device helpers of the multiresolution hash / tiled grid encoder (gridencoder.cu:50-159).
#pragma once
#include "common.cuh"

namespace foc {

// element <-> float with the reference's rounding points: c10::Half arithmetic is "convert both to float, operate,
// narrow the result to half", and `half += float` first narrows the float (implicit Half(float) conversion).
template <typename T> struct Elem;
template <> struct Elem<float> {
    static __device__ __forceinline__ float ld(const float* p) { return __ldg(p); }
    static __device__ __forceinline__ float rnd(float v) { return v; }
    static __device__ __forceinline__ void st(float* p, float v) { *p = v; }
    // results += w * g   (nvcc contracts the reference's fp32 expression into one FMA)
    static __device__ __forceinline__ float acc(float res, float w, float g) { return fmaf(w, g, res); }
};
template <> struct Elem<__half> {
    static __device__ __forceinline__ float ld(const __half* p) { return __half2float(__ldg(p)); }
    static __device__ __forceinline__ float rnd(float v) { return __half2float(__float2half_rn(v)); }
    static __device__ __forceinline__ void st(__half* p, float v) { *p = __float2half_rn(v); }
    static __device__ __forceinline__ float acc(float res, float w, float g) { return rnd(__fadd_rn(res, rnd(__fmul_rn(w, g)))); }
};

// gridencoder.cu:50-84: dense index while the running stride still fits the level, else xor-of-primes hash, then
// `% hashmap_size`.  Everything that does not depend on the corner is resolved ONCE per (thread, level): which dims enter
// the dense index, whether the level is hashed, and whether the final modulo is the identity (dense level: index <
// (res+1)^D <= size), a mask (power-of-two table, the 2^19 of every hashed level here) or a real division (tiled grids,
// odd sizes).  The per-corner cost drops from ~45 instructions (a 32-bit division each) to ~8; the indices are identical.
template <int D>
struct GridIndexer {
    uint32_t st[D];          // stride of dim d in the dense index, 0 when the reference's loop skips the dim
    uint32_t size, and_mask;
    bool hashed, need_mod;
    __device__ __forceinline__ GridIndexer(uint32_t gridtype, bool align_corners, uint32_t hashmap_size, uint32_t resolution) {
        uint32_t stride = 1;
        bool wrapped = false;                   // the reference's 32-bit stride product may wrap at very fine levels
        #pragma unroll
        for (int d = 0; d < D; ++d) {
            if (stride <= hashmap_size) {
                st[d] = stride;
                const uint64_t wide = (uint64_t)stride * (align_corners ? resolution : (resolution + 1));
                wrapped |= (wide >> 32) != 0;
                stride = (uint32_t)wide;
            } else st[d] = 0;
        }
        size = hashmap_size;
        const bool pow2 = (hashmap_size & (hashmap_size - 1)) == 0;
        hashed = gridtype == 0 && stride > hashmap_size;
        // the final `% size`: identity for a dense level (index < (res+1)^D <= size, unless the product wrapped), a mask
        // for a power-of-two hash table, a division otherwise
        need_mod = hashed ? !pow2 : !(stride <= hashmap_size && !wrapped);
        and_mask = (hashed && pow2) ? hashmap_size - 1 : 0xffffffffu;
    }
    // branch-free in the common configurations (lanes of a warp sit on different levels under the sample-major mapping)
    __device__ __forceinline__ uint32_t operator()(const uint32_t (&pg)[D]) const {
        constexpr uint32_t primes[7] = {1u, 2654435761u, 805459861u, 3674653429u, 2097192037u, 1434869437u, 2165219737u};
        uint32_t h = 0, dense = 0;
        #pragma unroll
        for (int d = 0; d < D; ++d) { h ^= pg[d] * primes[d]; dense += pg[d] * st[d]; }
        uint32_t index = hashed ? h : dense;
        if (need_mod) index %= size; else index &= and_mask;
        return index;
    }
};

// t -> (t / L, t % L) without a 64-bit division in the common cases (L a power of two; t < 2^32)
__device__ __forceinline__ void split_index(uint64_t t, uint32_t L, uint32_t& q, uint32_t& r) {
    if ((L & (L - 1)) == 0) { q = (uint32_t)(t >> (__ffs((int)L) - 1)); r = (uint32_t)t & (L - 1); }
    else if ((t >> 32) == 0) { const uint32_t t32 = (uint32_t)t; q = t32 / L; r = t32 - q * L; }
    else { q = (uint32_t)(t / L); r = (uint32_t)(t % L); }
}

struct GridLevel {
    float scale;
    uint32_t resolution, hashmap_size, offset;
};
// gridencoder.cu:137-139.  exp2f() is the same device routine the reference calls, so non-integer per-level scales
// round identically; `x * H - 1` is one FMA in the reference's SASS.
__device__ __forceinline__ GridLevel grid_level(const int32_t* __restrict__ offsets, uint32_t level, float S, uint32_t H) {
    GridLevel l;
    l.offset = (uint32_t)__ldg(offsets + level);
    l.hashmap_size = (uint32_t)__ldg(offsets + level + 1) - l.offset;
    l.scale = fmaf(exp2f((float)level * S), (float)H, -1.0f);
    l.resolution = (uint32_t)ceilf(l.scale) + 1;
    return l;
}

// gridencoder.cu:146-159: cell corner + fractional position (optionally smoothstepped)
template <int D>
__device__ __forceinline__ void grid_locate(const float (&x)[D], float scale, bool align_corners, uint32_t interp,
                                            float (&pos)[D], float (&pder)[D], uint32_t (&pg)[D]) {
    #pragma unroll
    for (int d = 0; d < D; ++d) {
        float p = fmaf(x[d], scale, align_corners ? 0.0f : 0.5f);
        pg[d] = (uint32_t)floorf(p);
        p = __fsub_rn(p, (float)pg[d]);
        if (interp == 1) { pder[d] = 6.0f * p * (1.0f - p); p = p * p * (3.0f - 2.0f * p); }
        else pder[d] = 1.0f;
        pos[d] = p;
    }
}

template <int D>
__device__ __forceinline__ bool grid_oob(const float (&x)[D]) {
    bool oob = false;
    #pragma unroll
    for (int d = 0; d < D; ++d) oob |= (x[d] < 0.0f) | (x[d] > 1.0f);
    return oob;
}

}  // namespace foc
