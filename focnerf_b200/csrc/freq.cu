// freq.cu -- frequency (positional) encoding fwd/bwd (K19/K20: freqencoder/src/freqencoder.cu:30-129) and the
// degree<=4 real spherical-harmonics direction encoder that nerf/network_ff.py:5 imports from the `shencoder`
// package, which is ABSENT from the reference tree (parity unpinned for that one op).
//
// Both are pure streaming kernels: one thread per OUTPUT element so stores are fully coalesced; the D (=3) inputs
// of a point are re-read by its C threads out of L1.  The reference compiles freqencoder with -use_fast_math, so
// the forward uses the same __sinf(scalbnf(x, f) + phase) expression.
#include "common.cuh"

using namespace foc;

__global__ void __launch_bounds__(256)
k_freq_fwd(const float* __restrict__ inputs, uint32_t B, uint32_t D, uint32_t deg, uint32_t C, float* __restrict__ outputs) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (uint64_t)B * C) return;
    const uint32_t b = (uint32_t)(t / C), c = (uint32_t)(t - (uint64_t)b * C);
    const float* x = inputs + (size_t)b * D;
    if (c < D) { outputs[t] = __ldg(x + c); return; }
    const uint32_t col = c / D - 1, d = c % D, freq = col >> 1;
    const float phase = (float)(col & 1u) * 1.57079632679489661923f;
    outputs[t] = __sinf(scalbnf(__ldg(x + d), (int)freq) + phase);
}

// d/dx [sin(2^f x), cos(2^f x)] expressed through the stored outputs (freqencoder.cu:63-94)
__global__ void __launch_bounds__(256)
k_freq_bwd(const float* __restrict__ grad, const float* __restrict__ outputs, uint32_t B, uint32_t D, uint32_t deg, uint32_t C,
           float* __restrict__ grad_inputs) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (uint64_t)B * D) return;
    const uint32_t b = (uint32_t)(t / D), d = (uint32_t)(t - (uint64_t)b * D);
    const float* g = grad + (size_t)b * C;
    const float* o = outputs + (size_t)b * C;
    float r = __ldg(g + d);
    g += D; o += D;
    for (uint32_t f = 0; f < deg; ++f) {
        r += scalbnf(1.0f, (int)f) * (__ldg(g + d) * __ldg(o + D + d) - __ldg(g + D + d) * __ldg(o + d));
        g += 2 * D; o += 2 * D;
    }
    grad_inputs[t] = r;
}

// Real SH basis up to degree 4 (16 coefficients) of direction (x,y,z); closed form shared with torch-ngp's shencoder
// and tiny-cuda-nn's SphericalHarmonics encoding.  One thread per point: 12 B in, up to 64 B out.
__global__ void __launch_bounds__(256)
k_sh_fwd(const float* __restrict__ dirs, uint32_t B, uint32_t degree, float* __restrict__ out) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const float x = __ldg(dirs + 3 * (size_t)b), y = __ldg(dirs + 3 * (size_t)b + 1), z = __ldg(dirs + 3 * (size_t)b + 2);
    const uint32_t Cn = degree * degree;
    float* o = out + (size_t)Cn * b;
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    o[0] = 0.28209479177387814f;
    if (degree <= 1) return;
    o[1] = -0.48860251190291987f * y; o[2] = 0.48860251190291987f * z; o[3] = -0.48860251190291987f * x;
    if (degree <= 2) return;
    o[4] = 1.0925484305920792f * xy; o[5] = -1.0925484305920792f * yz;
    o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
    o[7] = -1.0925484305920792f * xz; o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
    if (degree <= 3) return;
    o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2); o[10] = 2.8906114426405538f * xy * z;
    o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2); o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
    o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2); o[14] = 1.4453057213202769f * z * (x2 - y2);
    o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
}

static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

FOC_API int foc_freq_encode_forward(const float* inputs, uint32_t B, uint32_t D, uint32_t deg, uint32_t C, float* outputs, void* stream) {
    if (B == 0) return 0;
    if (C != D + 2 * D * deg) return FOC_ERR_UNSUPPORTED;
    const uint64_t total = (uint64_t)B * C;
    k_freq_fwd<<<(uint32_t)((total + 255) / 256), 256, 0, S(stream)>>>(inputs, B, D, deg, C, outputs);
    FOC_LAUNCH_CHECK(); return 0;
}
FOC_API int foc_freq_encode_backward(const float* grad, const float* outputs, uint32_t B, uint32_t D, uint32_t deg, uint32_t C,
                                     float* grad_inputs, void* stream) {
    if (B == 0) return 0;
    if (C != D + 2 * D * deg) return FOC_ERR_UNSUPPORTED;
    const uint64_t total = (uint64_t)B * D;
    k_freq_bwd<<<(uint32_t)((total + 255) / 256), 256, 0, S(stream)>>>(grad, outputs, B, D, deg, C, grad_inputs);
    FOC_LAUNCH_CHECK(); return 0;
}
FOC_API int foc_sh_encode_forward(const float* dirs, uint32_t B, uint32_t degree, float* outputs, void* stream) {
    if (B == 0) return 0;
    if (degree < 1 || degree > 4) return FOC_ERR_UNSUPPORTED;
    k_sh_fwd<<<ceil_div(B, 256), 256, 0, S(stream)>>>(dirs, B, degree, outputs);
    FOC_LAUNCH_CHECK(); return 0;
}
