// field_glue.cu -- the elementwise wiring BETWEEN the encoders and the two fused MLPs of the field network, fused.
//
// Reference (nerf/network_ff.py:51-74; activation.py:5-18), all separate PyTorch kernels per step:
//     sigma    = trunc_exp(h[..., 0])                      slice + cast to fp32 + exp
//     geo_feat = h[..., 1:]                                slice
//     d        = encoder_dir(d)                            SH degree 4 (the `shencoder` extension, absent from the reference tree)
//     h2       = cat([d, geo_feat, zeros_like(geo_feat[..., :1])])      zeros + cat (+ FFMLP's own zero-row padding cat)
//     rgb      = sigmoid(color_net(h2))                    sigmoid in fp16, then the compositor casts to fp32
// and the mirror image of each in backward (sigmoid_backward, exp(clamp(x,-15,15)), slice-backward zero fills, adds).
// Here: one thread per sample, one pass each way.  Rounding points follow the reference: geo features stay fp16; SH is
// evaluated in fp32 and rounded to fp16 when it enters the colour network; sigma = exp(float(h0)) in fp32; rgb is rounded
// through fp16 (torch.sigmoid on a half tensor) before the fp32 compositor sees it; gradients of half tensors are half.
#include "common.cuh"

using namespace foc;

namespace {

__device__ __forceinline__ void sh4(float x, float y, float z, float (&o)[16]) {
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    o[0] = 0.28209479177387814f;
    o[1] = -0.48860251190291987f * y; o[2] = 0.48860251190291987f * z; o[3] = -0.48860251190291987f * x;
    o[4] = 1.0925484305920792f * xy; o[5] = -1.0925484305920792f * yz;
    o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
    o[7] = -1.0925484305920792f * xz; o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
    o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2); o[10] = 2.8906114426405538f * xy * z;
    o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2); o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
    o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2); o[14] = 1.4453057213202769f * z * (x2 - y2);
    o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
}
__device__ __forceinline__ uint32_t pack2(float a, float b) { const __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<const uint32_t*>(&h); }
__device__ __forceinline__ uint32_t pack2h(__half a, __half b) { const __half2 h = __halves2half2(a, b); return *reinterpret_cast<const uint32_t*>(&h); }

}  // namespace

// h [B,16] f16 (sigma-net output), dirs [B,3] f32  ->  sigma [B] f32,  cin [B,32] f16 = [SH4(d) | h[1:16] | 0]
__global__ void __launch_bounds__(256)
k_field_mid_fwd(const float* __restrict__ dirs, const __half* __restrict__ h, uint32_t B, float* __restrict__ sigma, __half* __restrict__ cin) {
    pdl_sync();
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const uint4 h0 = __ldg(reinterpret_cast<const uint4*>(h + (size_t)b * 16)), h1 = __ldg(reinterpret_cast<const uint4*>(h + (size_t)b * 16) + 1);
    const __half* hv0 = reinterpret_cast<const __half*>(&h0);
    const __half* hv1 = reinterpret_cast<const __half*>(&h1);
    __half hh[16];
    #pragma unroll
    for (int i = 0; i < 8; ++i) { hh[i] = hv0[i]; hh[8 + i] = hv1[i]; }
    sigma[b] = expf(__half2float(hh[0]));                                       // activation.py:10 (torch.exp on fp32)
    float s[16];
    sh4(__ldg(dirs + 3 * (size_t)b), __ldg(dirs + 3 * (size_t)b + 1), __ldg(dirs + 3 * (size_t)b + 2), s);
    uint4* out = reinterpret_cast<uint4*>(cin + (size_t)b * 32);
    out[0] = make_uint4(pack2(s[0], s[1]), pack2(s[2], s[3]), pack2(s[4], s[5]), pack2(s[6], s[7]));
    out[1] = make_uint4(pack2(s[8], s[9]), pack2(s[10], s[11]), pack2(s[12], s[13]), pack2(s[14], s[15]));
    out[2] = make_uint4(pack2h(hh[1], hh[2]), pack2h(hh[3], hh[4]), pack2h(hh[5], hh[6]), pack2h(hh[7], hh[8]));
    out[3] = make_uint4(pack2h(hh[9], hh[10]), pack2h(hh[11], hh[12]), pack2h(hh[13], hh[14]), pack2h(hh[15], __float2half_rn(0.f)));
}

// grad_sigma [B] f32 (or NULL), grad_cin [B,32] f16, h [B,16] f16  ->  grad_h [B,16] f16
//   grad_h[0] = grad_sigma * exp(clamp(h0, -15, 15))  (activation.py:16);  grad_h[1:16] = grad_cin[16:31]
__global__ void __launch_bounds__(256)
k_field_mid_bwd(const float* __restrict__ grad_sigma, const __half* __restrict__ grad_cin, const __half* __restrict__ h, uint32_t B,
                __half* __restrict__ grad_h) {
    pdl_sync();
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const uint4 g2 = __ldg(reinterpret_cast<const uint4*>(grad_cin + (size_t)b * 32) + 2), g3 = __ldg(reinterpret_cast<const uint4*>(grad_cin + (size_t)b * 32) + 3);
    const __half* a = reinterpret_cast<const __half*>(&g2);
    const __half* c = reinterpret_cast<const __half*>(&g3);
    const float h0 = __half2float(__ldg(h + (size_t)b * 16));
    const float gs = grad_sigma ? __ldg(grad_sigma + b) * expf(fminf(15.0f, fmaxf(-15.0f, h0))) : 0.0f;
    uint4* out = reinterpret_cast<uint4*>(grad_h + (size_t)b * 16);
    out[0] = make_uint4(pack2h(__float2half_rn(gs), a[0]), pack2h(a[1], a[2]), pack2h(a[3], a[4]), pack2h(a[5], a[6]));
    out[1] = make_uint4(pack2h(a[7], c[0]), pack2h(c[1], c[2]), pack2h(c[3], c[4]), pack2h(c[5], c[6]));
}

// hc [B,16] f16 (colour-net output)  ->  rgb [B,3] f32 = float(half(sigmoid(float(hc[0:3]))))
__global__ void __launch_bounds__(256)
k_rgb_fwd(const __half* __restrict__ hc, uint32_t B, float* __restrict__ rgb) {
    pdl_sync();
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(hc + (size_t)b * 16));
    const __half* hv = reinterpret_cast<const __half*>(&v);
    #pragma unroll
    for (int c = 0; c < 3; ++c)
        rgb[3 * (size_t)b + c] = __half2float(__float2half_rn(1.0f / (1.0f + expf(-__half2float(hv[c])))));
}

// grad_rgb [B,3] f32, rgb [B,3] f32 (the forward's output)  ->  grad_hc [B,16] f16 (columns 3..15 zero)
__global__ void __launch_bounds__(256)
k_rgb_bwd(const float* __restrict__ grad_rgb, const float* __restrict__ rgb, uint32_t B, __half* __restrict__ grad_hc) {
    pdl_sync();
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    float g[3];
    #pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float y = __ldg(rgb + 3 * (size_t)b + c);
        const float go = __half2float(__float2half_rn(__ldg(grad_rgb + 3 * (size_t)b + c)));   // the gradient of a half tensor is half
        g[c] = go * (1.0f - y) * y;
    }
    uint4* out = reinterpret_cast<uint4*>(grad_hc + (size_t)b * 16);
    out[0] = make_uint4(pack2(g[0], g[1]), pack2(g[2], 0.f), 0u, 0u);
    out[1] = make_uint4(0u, 0u, 0u, 0u);
}

static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

FOC_API int foc_field_mid_forward(const float* dirs, const uint16_t* h, uint32_t B, float* sigma, uint16_t* cin, void* stream) {
    if (B == 0) return 0;
    foc::launch(k_field_mid_fwd, ceil_div(B, 256), 256, 0, S(stream), dirs, reinterpret_cast<const __half*>(h), B, sigma, reinterpret_cast<__half*>(cin));
    FOC_LAUNCH_CHECK(); return 0;
}
FOC_API int foc_field_mid_backward(const float* grad_sigma, const uint16_t* grad_cin, const uint16_t* h, uint32_t B, uint16_t* grad_h, void* stream) {
    if (B == 0) return 0;
    foc::launch(k_field_mid_bwd, ceil_div(B, 256), 256, 0, S(stream), grad_sigma, reinterpret_cast<const __half*>(grad_cin), reinterpret_cast<const __half*>(h), B,
                                                            reinterpret_cast<__half*>(grad_h));
    FOC_LAUNCH_CHECK(); return 0;
}
FOC_API int foc_rgb_sigmoid_forward(const uint16_t* hc, uint32_t B, float* rgb, void* stream) {
    if (B == 0) return 0;
    foc::launch(k_rgb_fwd, ceil_div(B, 256), 256, 0, S(stream), reinterpret_cast<const __half*>(hc), B, rgb);
    FOC_LAUNCH_CHECK(); return 0;
}
FOC_API int foc_rgb_sigmoid_backward(const float* grad_rgb, const float* rgb, uint32_t B, uint16_t* grad_hc, void* stream) {
    if (B == 0) return 0;
    foc::launch(k_rgb_bwd, ceil_div(B, 256), 256, 0, S(stream), grad_rgb, rgb, B, reinterpret_cast<__half*>(grad_hc));
    FOC_LAUNCH_CHECK(); return 0;
}
