// common.cuh -- shared helpers for libfoc_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

#define FOC_API extern "C" __attribute__((visibility("default")))
#define FOC_ERR_UNSUPPORTED (-1)

#define FOC_LAUNCH_CHECK() do { cudaError_t e__ = cudaGetLastError(); if (e__ != cudaSuccess) return (int)e__; } while (0)

#include <cstdlib>
#include <utility>

namespace foc {

constexpr int kNumSMs = 148;   // B200: 2 dies x 74 SMs (tuning thresholds; persistent grids ask the device: num_sms())

// SM count of the CURRENT device, cached per device (no process-wide assumption: a part with SMs fused off, or MIG, reports fewer)
inline int num_sms() {
    static int cache[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return kNumSMs;
    if (cache[dev] == 0) {
        int v = 0;
        cache[dev] = (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) ? v : kNumSMs;
    }
    return cache[dev];
}

// ---- programmatic dependent launch (PDL).  The train step is ~45 back-to-back launches of 5-50 us kernels; with PDL the
// next kernel's CTAs are scheduled while the previous kernel drains and run their prologue (barrier init, TMEM allocation,
// tensor-map prefetch, parameter loads) up to pdl_wait(), which returns once the previous grid has completed and its
// writes are visible.  Rules kept by every kernel in this library: no global-memory access before pdl_wait(); and
// pdl_trigger() right after it (dependents may be scheduled as soon as every CTA of this grid has started).
// Kernels launched after a kernel that never triggers (any torch kernel) simply start when it completes.
// Measured on B200 inside the CUDA-graph replay of the train step: 0.464 ms with the attribute, 0.456 ms without (the
// early-scheduled CTAs hold registers / shared memory / TMEM while they wait, and graph replay already has ~1 us
// kernel-to-kernel gaps), so the attribute is OFF by default; FOC_PDL=1 in the environment turns it on.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() { pdl_wait(); pdl_trigger(); }

inline bool pdl_enabled() {
    static const bool on = [] { const char* e = getenv("FOC_PDL"); return e && e[0] == '1'; }();
    return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

__host__ __device__ inline uint32_t ceil_div(uint32_t a, uint32_t b) { return (a + b - 1) / b; }

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is PER DEVICE: remember it per (kernel instantiation, device), not per process --
// one `DynSmemOptIn` object per kernel template instantiation (a function-local static in its launcher).  A racing second thread
// sets the same value again, which is harmless.
struct DynSmemOptIn {
    static constexpr int kMaxDevices = 64;
    bool done[kMaxDevices] = {};
    template <typename K> int ensure(K kernel, int bytes) {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return (int)e;
        if (dev < 0 || dev >= kMaxDevices) return FOC_ERR_UNSUPPORTED;
        if (!done[dev]) {
            e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
            if (e != cudaSuccess) return (int)e;
            done[dev] = true;
        }
        return 0;
    }
};

// Explicitly-rounded float ops: the marcher's integer results (sample counts, cell indices) must not
// depend on the compiler's FMA-contraction choices, so every operation states its rounding.
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(hi, fmaxf(lo, x)); }

__device__ __forceinline__ uint32_t spread3(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__device__ __forceinline__ uint32_t morton3(uint32_t x, uint32_t y, uint32_t z) {
    return spread3(x) | (spread3(y) << 1) | (spread3(z) << 2);
}
__device__ __forceinline__ uint32_t compact3(uint32_t x) {
    x &= 0x49249249u;
    x = (x | (x >> 2)) & 0xc30c30c3u;
    x = (x | (x >> 4)) & 0x0f00f00fu;
    x = (x | (x >> 8)) & 0xff0000ffu;
    x = (x | (x >> 16)) & 0x0000ffffu;
    return x;
}

// warp primitives -----------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_incl_sum(float v, int lane) {
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { float n = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += n; }
    return v;
}
__device__ __forceinline__ float warp_incl_prod(float v, int lane) {
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { float n = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v *= n; }
    return v;
}
__device__ __forceinline__ uint32_t warp_incl_sum_u32(uint32_t v, int lane) {
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += n; }
    return v;
}

// mlp_generic.cu: the FFMLP shapes outside the tensor-core kernels' (hidden 64, input 32 / 64)
bool mlp_generic_shape_ok(uint32_t B, uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers);
size_t mlp_generic_num_params(uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers);
size_t mlp_generic_scratch_bytes(uint32_t B, uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers);
int mlp_generic_forward(const uint16_t* inputs, const uint16_t* weights, uint32_t B, uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim,
                        uint32_t num_layers, uint32_t activation, uint32_t output_activation, uint16_t* forward_buffer, uint16_t* outputs, cudaStream_t st);
int mlp_generic_backward(const uint16_t* grad, const uint16_t* inputs, const uint16_t* weights, const uint16_t* forward_buffer, uint32_t B,
                         uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers, uint32_t activation, int calc_grad_inputs,
                         uint16_t* backward_buffer, uint16_t* grad_inputs, uint16_t* grad_weights, int accumulate, void* scratch, cudaStream_t st);
int mlp_generic_finish(const void* scratch, uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers, uint16_t* grad_weights, int accumulate, cudaStream_t st);
// num_layers == 1 (one hidden layer: tcnn's n_hidden_layers = 1; the reference's FFMLP module asserts >= 2) also goes to the generic kernels
inline bool mlp_tensor_shape(uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers) { return hidden_dim == 64 && (input_dim == 32 || input_dim == 64) && num_layers >= 2; }

}  // namespace foc
