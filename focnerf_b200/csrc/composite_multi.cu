// composite_multi.cu -- FOC-NeRF multi-object composition as fused kernels.
//
// Reference math (all PyTorch, nothing fused): COMBINED.py:451-534 `run` (uniform samples, per-object weights and colour
// mask), COMBINED.py:247-251 `best_densities_and_colors_v3` (per-sample arg-max density, strict '>' so the EARLIER object
// wins ties), COMBINED.py:141-200 `image_depth_generation` (front-to-back composite, 4th channel = sum w*sigma),
// editable.py:443-471 `modify_rays_for_object` (per-object ray-origin offset + re-normalised direction).
// The reference materialises densities [K,N,T] and rgbs [K,N,T,3], a running-max copy of both, and every
// intermediate of the cumprod compositor in HBM; here:
//   foc_sample_uniform   : (optional offset) -> near/far -> T uniform samples clipped to the AABB, one launch;
//   foc_composite_multi  : ONE WARP PER RAY reads the K objects' sigma (coalesced over samples), selects the winner per
//                          sample in registers, fetches only the winner's colour, optionally applies each object's own
//                          `w_k > 1e-10` colour mask (per-object transmittance scans), and composites with warp scans.
#include <math.h>
#include "common.cuh"

using namespace foc;

namespace {

constexpr int kMaxObjects = 16;
constexpr int kWarpsPerCta = 4;

// torch.linspace(0, 1, T) in fp32: start + step*i below the midpoint, end - step*(T-1-i) above it
__device__ __forceinline__ float linspace01(uint32_t i, uint32_t T) {
    if (T <= 1) return 0.0f;
    const float step = __fdiv_rn(1.0f, (float)(T - 1));
    return (i < T / 2) ? __fmul_rn(step, (float)i) : __fsub_rn(1.0f, __fmul_rn(step, (float)(T - 1 - i)));
}
// z_i = near + (far - near) * u_i : separate mul and add, as the eager torch ops round (COMBINED.py:464-466)
__device__ __forceinline__ float z_at(float near, float span, uint32_t i, uint32_t T) {
    return __fadd_rn(near, __fmul_rn(span, linspace01(i, T)));
}
__device__ __forceinline__ float delta_at(float near, float span, uint32_t i, uint32_t T) {
    return (i + 1 < T) ? __fsub_rn(z_at(near, span, i + 1, T), z_at(near, span, i, T)) : __fdiv_rn(span, (float)T);
}

}  // namespace

// editable.py:443-471 + raymarching.cu:92-145 + COMBINED.py:464-474.  One thread per ray prepares (o', d', near, far);
// a second grid-stride phase of the same kernel would need a grid sync, so sampling is a separate launch.
__global__ void __launch_bounds__(256)
k_prepare_rays(const float* __restrict__ rays_o, const float* __restrict__ rays_d, const float* __restrict__ offset,
               int normalize_d, const float* __restrict__ aabb, float min_near, uint32_t N,
               float* __restrict__ out_o, float* __restrict__ out_d, float* __restrict__ nears, float* __restrict__ fars) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    float o[3], d[3];
    #pragma unroll
    for (int a = 0; a < 3; ++a) { o[a] = __ldg(rays_o + 3 * (size_t)n + a); d[a] = __ldg(rays_d + 3 * (size_t)n + a); }
    if (offset) {
        #pragma unroll
        for (int a = 0; a < 3; ++a) o[a] = __fadd_rn(o[a], __ldg(offset + a));
    }
    if (normalize_d) {                               // F.normalize: d / max(||d||, 1e-12)
        const float nrm = fmaxf(sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]), 1e-12f);
        #pragma unroll
        for (int a = 0; a < 3; ++a) d[a] = __fdiv_rn(d[a], nrm);
    }
    float tn = 0.f, tf = 0.f;
    bool miss = false;
    #pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float r = __frcp_rn(d[a]);
        float lo = __fmul_rn(__fsub_rn(__ldg(aabb + a), o[a]), r), hi = __fmul_rn(__fsub_rn(__ldg(aabb + a + 3), o[a]), r);
        if (lo > hi) { const float s = lo; lo = hi; hi = s; }
        if (a == 0) { tn = lo; tf = hi; }
        else if (!miss) {
            if (tn > hi || lo > tf) miss = true;
            else { if (lo > tn) tn = lo; if (hi < tf) tf = hi; }
        }
    }
    if (miss) { tn = tf = 3.402823466e+38f; }
    else if (tn < min_near) tn = min_near;
    #pragma unroll
    for (int a = 0; a < 3; ++a) { out_o[3 * (size_t)n + a] = o[a]; out_d[3 * (size_t)n + a] = d[a]; }
    nears[n] = tn; fars[n] = tf;
}

// xyzs[n, i, :] = clip(o + d * z_i, aabb): one thread per (ray, sample) -> coalesced 12-byte rows
__global__ void __launch_bounds__(256)
k_sample_uniform(const float* __restrict__ o, const float* __restrict__ d, const float* __restrict__ nears,
                 const float* __restrict__ fars, const float* __restrict__ aabb, uint32_t N, uint32_t T, float* __restrict__ xyzs) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (uint64_t)N * T) return;
    const uint32_t n = (uint32_t)(t / T), i = (uint32_t)(t - (uint64_t)n * T);
    const float near = __ldg(nears + n), span = __fsub_rn(__ldg(fars + n), near);
    const float z = z_at(near, span, i, T);
    #pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float v = __fadd_rn(__ldg(o + 3 * (size_t)n + a), __fmul_rn(__ldg(d + 3 * (size_t)n + a), z));
        xyzs[3 * t + a] = fminf(fmaxf(v, __ldg(aabb + a)), __ldg(aabb + a + 3));
    }
}

__global__ void __launch_bounds__(32 * kWarpsPerCta)
k_composite_multi(const float* __restrict__ sigmas, const float* __restrict__ rgbs, const float* __restrict__ nears,
                  const float* __restrict__ fars, const float* __restrict__ obj_nears, const float* __restrict__ obj_fars,
                  uint32_t K, uint32_t N, uint32_t T, float bg, float mask_thresh, float T_thresh,
                  float* __restrict__ image, float* __restrict__ depth, float* __restrict__ weights_sum) {
    const int lane = threadIdx.x & 31;
    const uint32_t n = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
    if (n >= N) return;
    const float near = __ldg(nears + n), far = __ldg(fars + n), span = __fsub_rn(far, near);
    const bool masked = obj_nears != nullptr;
    float Tk[kMaxObjects], nk[kMaxObjects], sk[kMaxObjects];        // per-object transmittance carry, near, span
    #pragma unroll
    for (int k = 0; k < kMaxObjects; ++k) {
        Tk[k] = 1.0f; nk[k] = near; sk[k] = span;
        if (masked && (uint32_t)k < K) { nk[k] = __ldg(obj_nears + (size_t)k * N + n); sk[k] = __fsub_rn(__ldg(obj_fars + (size_t)k * N + n), nk[k]); }
    }
    float Tr = 1.0f, a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, dsum = 0.f, ws = 0.f;
    for (uint32_t base = 0; base < T; base += 32) {
        const uint32_t i = base + lane;
        const bool valid = i < T;
        float best = 0.0f;
        int best_k = 0;
        bool keep = true;                                             // winner's colour survives its own mask
        #pragma unroll
        for (int k = 0; k < kMaxObjects; ++k) {
            if ((uint32_t)k >= K) break;
            const float s = valid ? __ldg(sigmas + ((size_t)k * N + n) * T + i) : 0.0f;
            bool pass = true;
            if (masked) {                                             // w_k = alpha_k * prod_{j<i}(1 - alpha_k + 1e-15)  (COMBINED.py:515-525)
                const float ak = valid ? 1.0f - expf(-delta_at(nk[k], sk[k], i, T) * s) : 0.0f;
                const float f = valid ? __fadd_rn(__fsub_rn(1.0f, ak), 1e-15f) : 1.0f;
                const float incl = warp_incl_prod(f, lane);
                const float up = __shfl_up_sync(0xffffffffu, incl, 1);
                pass = ak * (Tk[k] * (lane == 0 ? 1.0f : up)) > mask_thresh;
                Tk[k] *= __shfl_sync(0xffffffffu, incl, 31);
            }
            if (k == 0 || s > best) { best = s; best_k = k; keep = pass; }
        }
        float c0 = 0.f, c1 = 0.f, c2 = 0.f;
        if (valid && keep) {
            const float* c = rgbs + 3 * (((size_t)best_k * N + n) * T + i);
            c0 = __ldg(c); c1 = __ldg(c + 1); c2 = __ldg(c + 2);
        }
        const float alpha = valid ? 1.0f - expf(-delta_at(near, span, i, T) * best) : 0.0f;
        const float f = valid ? __fadd_rn(__fsub_rn(1.0f, alpha), 1e-15f) : 1.0f;
        const float incl = warp_incl_prod(f, lane);
        const float up = __shfl_up_sync(0xffffffffu, incl, 1);
        const float w = alpha * (Tr * (lane == 0 ? 1.0f : up));
        if (valid) {
            a0 = fmaf(w, c0, a0); a1 = fmaf(w, c1, a1); a2 = fmaf(w, c2, a2); a3 = fmaf(w, best, a3);
            const float od = fminf(1.0f, fmaxf(0.0f, __fdiv_rn(__fsub_rn(z_at(near, span, i, T), near), span)));
            dsum = fmaf(w, od, dsum); ws += w;
        }
        Tr *= __shfl_sync(0xffffffffu, incl, 31);
        if (Tr < T_thresh) break;                                     // T_thresh = 0 reproduces the reference exactly
    }
    a0 = warp_sum(a0); a1 = warp_sum(a1); a2 = warp_sum(a2); a3 = warp_sum(a3); dsum = warp_sum(dsum); ws = warp_sum(ws);
    if (lane == 0) {
        const float r = (1.0f - ws) * bg;
        float4 o;
        o.x = fminf(1.0f, fmaxf(0.0f, a0 + r)); o.y = fminf(1.0f, fmaxf(0.0f, a1 + r));
        o.z = fminf(1.0f, fmaxf(0.0f, a2 + r)); o.w = fminf(1.0f, fmaxf(0.0f, a3 + r));
        *reinterpret_cast<float4*>(image + 4 * (size_t)n) = o;
        depth[n] = dsum; weights_sum[n] = ws;
    }
}

static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

FOC_API int foc_sample_uniform(const float* rays_o, const float* rays_d, const float* offset, int normalize_d, const float* aabb,
                               float min_near, uint32_t N, uint32_t T, float* out_o, float* out_d, float* nears, float* fars,
                               float* xyzs, void* stream) {
    if (N == 0) return 0;
    k_prepare_rays<<<ceil_div(N, 256), 256, 0, S(stream)>>>(rays_o, rays_d, offset, normalize_d, aabb, min_near, N, out_o, out_d, nears, fars);
    FOC_LAUNCH_CHECK();
    if (T == 0 || !xyzs) return 0;
    const uint64_t total = (uint64_t)N * T;
    k_sample_uniform<<<(uint32_t)((total + 255) / 256), 256, 0, S(stream)>>>(out_o, out_d, nears, fars, aabb, N, T, xyzs);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_composite_multi(const float* sigmas, const float* rgbs, const float* nears, const float* fars,
                                const float* obj_nears, const float* obj_fars, uint32_t K, uint32_t N, uint32_t T, float bg,
                                float mask_thresh, float T_thresh, float* image, float* depth, float* weights_sum, void* stream) {
    if (N == 0) return 0;
    if (K == 0 || K > (uint32_t)kMaxObjects || T == 0 || (obj_nears == nullptr) != (obj_fars == nullptr)) return FOC_ERR_UNSUPPORTED;
    k_composite_multi<<<ceil_div(N, kWarpsPerCta), 32 * kWarpsPerCta, 0, S(stream)>>>(
        sigmas, rgbs, nears, fars, obj_nears, obj_fars, K, N, T, bg, mask_thresh, T_thresh, image, depth, weights_sum);
    FOC_LAUNCH_CHECK(); return 0;
}
