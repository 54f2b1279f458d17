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

// gridencoder.cu:50-84: dense index while the running stride still fits the level, else xor-of-primes hash.
template <int D>
__device__ __forceinline__ uint32_t grid_index(uint32_t gridtype, bool align_corners, uint32_t hashmap_size,
                                               uint32_t resolution, const uint32_t (&pg)[D]) {
    constexpr uint32_t primes[7] = {1u, 2654435761u, 805459861u, 3674653429u, 2097192037u, 1434869437u, 2165219737u};
    uint32_t stride = 1, index = 0;
    #pragma unroll
    for (int d = 0; d < D; ++d) {
        if (stride <= hashmap_size) {
            index += pg[d] * stride;
            stride *= align_corners ? resolution : (resolution + 1);
        }
    }
    if (gridtype == 0 && stride > hashmap_size) {
        index = 0;
        #pragma unroll
        for (int d = 0; d < D; ++d) index ^= pg[d] * primes[d];
    }
    return index % hashmap_size;
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
