// mlp_generic.cu -- the shapes of the reference's FFMLP operator that the FOC field does NOT use (ffmlp/ffmlp.py:99-123 allows
// hidden_dim in {16, 32, 64, 128, 256}, any input_dim = 16 m, num_layers >= 2, every activation of utils.h:424-470).
//
// The field's own networks (hidden 64, input 32 / 64) run on the tcgen05 kernels of mlp_fwd.cu / mlp_bwd.cu / field_fwd.cu; everything
// else the reference module accepts -- e.g. the 144 -> 16 -> 16 feature encoder and the 48-wide colour input of nerf/network_tcnn.py:504-546,
// or the 128-wide 5-layer variant its comments mention -- comes here so that the operator is not a narrower one than the one it replaces.
// These are cold paths (a few thousand rows per call in the reference's callers): plain CUDA cores, fp32 accumulation in k order, one CTA
// per 32 rows with the activations of the current layer in shared memory.  Same storage roundings as the tensor-core path and the oracle
// (oracle/orc_mlp.c): every layer's activations and every gradient buffer are rounded once to binary16.
//
// Weight gradients: each CTA adds its 32-row partial sums into an fp32 buffer with atomicAdd (the caller's scratch, zeroed by the entry
// point); the arrival order is not fixed, so the last fp32 bit of a sum can change between runs (the tensor-core path reduces in a fixed
// order).  k_mlp_generic_finish rounds to fp16 (or adds to a persistent fp16 buffer), like foc_ffmlp_wgrad_finish.
#include "common.cuh"
#include "../../include/foc_b200.h"

using namespace foc;

namespace {

constexpr int kRows = 32, kThreads = 256, kOutPad = 16;

// ffmlp/src/utils.h:424-470 and 537-582 (through the stored POST-activation value)
__device__ __forceinline__ float g_act(uint32_t a, float v) {
    switch (a) {
        case 0: return v > 0.f ? v : 0.f;
        case 1: return __expf(v);
        case 2: return __sinf(v);
        case 3: return 1.0f / (1.0f + __expf(-v));
        case 4: { const float x = v * 10.0f; return 0.5f * (x + sqrtf(x * x + 4.0f)) / 10.0f; }
        case 5: return __logf(__expf(v * 10.0f) + 1.0f) / 10.0f;
        default: return v;
    }
}
__device__ __forceinline__ float g_act_bwd(uint32_t a, float g, float fwd) {
    switch (a) {
        case 0: return fwd > 0.f ? g : 0.f;
        case 1: return g * fwd;
        case 3: return g * (fwd * (1.0f - fwd));
        case 4: { const float y = fwd * 10.0f; return g * (y * y / (y * y + 1.0f)); }
        case 5: return g * (1.0f - __expf(-fwd * 10.0f));
        default: return g;
    }
}
__device__ __forceinline__ float rh(float v) { return __half2float(__float2half_rn(v)); }

struct GenParams {
    const __half* inputs; const __half* weights; __half* forward_buffer; __half* outputs;
    const __half* grad; const __half* fwd_in; __half* backward_buffer; __half* grad_inputs; float* wsum;
    uint32_t B, in, hidden, layers, out_dim, act, out_act;
    int calc_grad_inputs;
};

// y[r][c] = act(sum_k x[r][k] * W[c*K + k]); lanes of a warp share c (weight loads broadcast) and differ in r
__global__ void __launch_bounds__(kThreads) k_mlp_generic_fwd(const GenParams p) {
    pdl_sync();
    extern __shared__ __align__(16) uint8_t g_smem[];
    const uint32_t H = p.hidden, ld = H + 2;
    __half* bufA = reinterpret_cast<__half*>(g_smem);
    __half* bufB = bufA + kRows * ld;
    const uint32_t r = threadIdx.x & 31u, c0 = threadIdx.x >> 5;
    const size_t row = (size_t)blockIdx.x * kRows + r;
    // layer 0 straight from global memory (input_dim is not bounded by the shared-memory budget)
    {
        const __half2* x = reinterpret_cast<const __half2*>(p.inputs + row * p.in);
        for (uint32_t c = c0; c < H; c += kThreads / 32) {
            const __half2* w = reinterpret_cast<const __half2*>(p.weights + (size_t)c * p.in);
            float acc = 0.f;
            for (uint32_t k = 0; k < p.in / 2; ++k) { const float2 a = __half22float2(x[k]), b = __half22float2(__ldg(w + k)); acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc); }
            const __half h = __float2half_rn(g_act(p.act, rh(acc)));
            bufA[r * ld + c] = h;
            if (p.forward_buffer) p.forward_buffer[row * H + c] = h;
        }
    }
    __syncthreads();
    __half* cur = bufA; __half* nxt = bufB;
    const __half* W = p.weights + (size_t)H * p.in;
    for (uint32_t l = 1; l < p.layers; ++l, W += (size_t)H * H) {
        for (uint32_t c = c0; c < H; c += kThreads / 32) {
            const __half* w = W + (size_t)c * H;
            float acc = 0.f;
            for (uint32_t k = 0; k < H; ++k) acc = fmaf(__half2float(cur[r * ld + k]), __half2float(__ldg(w + k)), acc);
            const __half h = __float2half_rn(g_act(p.act, rh(acc)));
            nxt[r * ld + c] = h;
            if (p.forward_buffer) p.forward_buffer[((size_t)l * p.B + row) * H + c] = h;
        }
        __syncthreads();
        __half* t = cur; cur = nxt; nxt = t;
    }
    for (uint32_t c = c0; c < (uint32_t)kOutPad; c += kThreads / 32) {          // rows >= out_dim of W_out exist (zero-padded parameter rows, ffmlp.cu:632)
        const __half* w = W + (size_t)c * H;
        float acc = 0.f;
        for (uint32_t k = 0; k < H; ++k) acc = fmaf(__half2float(cur[r * ld + k]), __half2float(__ldg(w + k)), acc);
        p.outputs[row * kOutPad + c] = __float2half_rn(g_act(p.out_act, rh(acc)));
    }
}

// one CTA: the 32-row slice of the whole backward pass (gradient chain, optional input gradient, partial weight gradients)
__global__ void __launch_bounds__(kThreads) k_mlp_generic_bwd(const GenParams p) {
    pdl_sync();
    extern __shared__ __align__(16) uint8_t g_smem[];
    const uint32_t H = p.hidden, ld = H + 2, L = p.layers, B = p.B;
    __half* bufA = reinterpret_cast<__half*>(g_smem);
    __half* bufB = bufA + kRows * ld;
    __shared__ __half s_g[kRows][kOutPad + 2];
    const uint32_t r = threadIdx.x & 31u, c0 = threadIdx.x >> 5;
    const size_t row0 = (size_t)blockIdx.x * kRows, row = row0 + r;
    const size_t in_sz = (size_t)H * p.in, hh = (size_t)H * H;
    const __half* W_out = p.weights + in_sz + (size_t)(L - 1) * hh;
    for (uint32_t i = threadIdx.x; i < kRows * kOutPad; i += kThreads) s_g[i / kOutPad][i % kOutPad] = p.grad[row0 * kOutPad + i];
    __syncthreads();
    // slot 0: gradient at the last hidden layer's pre-activation
    for (uint32_t j = c0; j < H; j += kThreads / 32) {
        float acc = 0.f;
        #pragma unroll
        for (uint32_t o = 0; o < (uint32_t)kOutPad; ++o) acc = fmaf(__half2float(s_g[r][o]), __half2float(__ldg(W_out + (size_t)o * H + j)), acc);
        const __half h = __float2half_rn(g_act_bwd(p.act, rh(acc), __half2float(p.fwd_in[((size_t)(L - 1) * B + row) * H + j])));
        bufA[r * ld + j] = h;
        p.backward_buffer[row * H + j] = h;
    }
    __syncthreads();
    __half* cur = bufA; __half* nxt = bufB;
    for (uint32_t k = 1; k < L; ++k) {
        const uint32_t m = L - 1 - k;
        const __half* W = p.weights + in_sz + (size_t)m * hh;
        for (uint32_t j = c0; j < H; j += kThreads / 32) {
            float acc = 0.f;
            for (uint32_t o = 0; o < H; ++o) acc = fmaf(__half2float(cur[r * ld + o]), __half2float(__ldg(W + (size_t)o * H + j)), acc);
            const __half h = __float2half_rn(g_act_bwd(p.act, rh(acc), __half2float(p.fwd_in[((size_t)m * B + row) * H + j])));
            nxt[r * ld + j] = h;
            p.backward_buffer[((size_t)k * B + row) * H + j] = h;
        }
        __syncthreads();
        __half* t = cur; cur = nxt; nxt = t;
    }
    if (p.calc_grad_inputs) {
        for (uint32_t i = c0; i < p.in; i += kThreads / 32) {
            float acc = 0.f;
            for (uint32_t o = 0; o < H; ++o) acc = fmaf(__half2float(cur[r * ld + o]), __half2float(__ldg(p.weights + (size_t)o * p.in + i)), acc);
            p.grad_inputs[row * p.in + i] = __float2half_rn(acc);
        }
    }
    if (p.wsum == nullptr) return;
    __syncthreads();                                                    // this CTA's backward_buffer rows are complete (and visible to it)
    // partial weight gradients of the 32 rows: dW[o][j] += d[b][o] * a[b][j]
    const size_t n_hidden = in_sz + (size_t)(L - 1) * hh;
    for (size_t e = threadIdx.x; e < (size_t)kOutPad * H; e += kThreads) {         // output matrix: d = grad, a = last hidden activations
        const uint32_t o = (uint32_t)(e / H), j = (uint32_t)(e % H);
        float acc = 0.f;
        for (uint32_t b = 0; b < (uint32_t)kRows; ++b) acc = fmaf(__half2float(s_g[b][o]), __half2float(p.fwd_in[((size_t)(L - 1) * B + row0 + b) * H + j]), acc);
        if (acc != 0.f) atomicAdd(p.wsum + n_hidden + e, acc);
    }
    for (uint32_t m = 0; m + 1 < L; ++m) {                                        // hidden matrices: d = slot L-2-m, a = activations of layer m
        const __half* d = p.backward_buffer + ((size_t)(L - 2 - m) * B + row0) * H;
        const __half* a = p.fwd_in + ((size_t)m * B + row0) * H;
        for (size_t e = threadIdx.x; e < hh; e += kThreads) {
            const uint32_t o = (uint32_t)(e / H), j = (uint32_t)(e % H);
            float acc = 0.f;
            for (uint32_t b = 0; b < (uint32_t)kRows; ++b) acc = fmaf(__half2float(d[(size_t)b * H + o]), __half2float(a[(size_t)b * H + j]), acc);
            if (acc != 0.f) atomicAdd(p.wsum + in_sz + (size_t)m * hh + e, acc);
        }
    }
    {                                                                              // input matrix: d = slot L-1, a = inputs
        const __half* d = p.backward_buffer + ((size_t)(L - 1) * B + row0) * H;
        const __half* a = p.inputs + row0 * p.in;
        for (size_t e = threadIdx.x; e < in_sz; e += kThreads) {
            const uint32_t o = (uint32_t)(e / p.in), i = (uint32_t)(e % p.in);
            float acc = 0.f;
            for (uint32_t b = 0; b < (uint32_t)kRows; ++b) acc = fmaf(__half2float(d[(size_t)b * H + o]), __half2float(a[(size_t)b * p.in + i]), acc);
            if (acc != 0.f) atomicAdd(p.wsum + e, acc);
        }
    }
}

__global__ void __launch_bounds__(256) k_mlp_generic_finish(const float* __restrict__ wsum, uint32_t n, __half* __restrict__ out, int accumulate) {
    pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float v = wsum[i];
    out[i] = __float2half_rn(accumulate ? __half2float(out[i]) + v : v);
}

}  // namespace

namespace foc {

bool mlp_generic_shape_ok(uint32_t B, uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers) {
    const bool h_ok = hidden_dim == 16 || hidden_dim == 32 || hidden_dim == 64 || hidden_dim == 128 || hidden_dim == 256;
    return h_ok && input_dim != 0 && input_dim % 16 == 0 && output_dim <= (uint32_t)kOutPad && num_layers >= 1 && num_layers <= 16 && B % 128 == 0;
}
// fp32 weight-gradient sums, then room for a backward_buffer [num_layers, B, hidden] f16 when the caller passes none
size_t mlp_generic_scratch_bytes(uint32_t B, uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers) {
    const size_t n = (size_t)hidden_dim * (input_dim + (size_t)hidden_dim * (num_layers - 1) + kOutPad);
    return ((n * sizeof(float) + 255) & ~(size_t)255) + (size_t)num_layers * B * hidden_dim * 2;
}
size_t mlp_generic_num_params(uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers) {
    return (size_t)hidden_dim * (input_dim + (size_t)hidden_dim * (num_layers - 1) + kOutPad);
}

int mlp_generic_forward(const uint16_t* inputs, const uint16_t* weights, uint32_t B, uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim,
                        uint32_t num_layers, uint32_t activation, uint32_t output_activation, uint16_t* forward_buffer, uint16_t* outputs, cudaStream_t st) {
    if (!mlp_generic_shape_ok(B, input_dim, output_dim, hidden_dim, num_layers)) return FOC_ERR_UNSUPPORTED;
    GenParams p{};
    p.inputs = reinterpret_cast<const __half*>(inputs); p.weights = reinterpret_cast<const __half*>(weights);
    p.forward_buffer = reinterpret_cast<__half*>(forward_buffer); p.outputs = reinterpret_cast<__half*>(outputs);
    p.B = B; p.in = input_dim; p.hidden = hidden_dim; p.layers = num_layers; p.out_dim = output_dim; p.act = activation; p.out_act = output_activation;
    const uint32_t smem = 2u * kRows * (hidden_dim + 2) * 2u;
    foc::launch(k_mlp_generic_fwd, B / kRows, kThreads, smem, st, p);
    FOC_LAUNCH_CHECK(); return 0;
}

int mlp_generic_backward(const uint16_t* grad, const uint16_t* inputs, const uint16_t* weights, const uint16_t* forward_buffer, uint32_t B,
                         uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers, uint32_t activation, int calc_grad_inputs,
                         uint16_t* backward_buffer, uint16_t* grad_inputs, uint16_t* grad_weights, int accumulate, void* scratch, cudaStream_t st) {
    if (!mlp_generic_shape_ok(B, input_dim, output_dim, hidden_dim, num_layers) || activation == 2 || !scratch) return FOC_ERR_UNSUPPORTED;
    const size_t n = mlp_generic_num_params(input_dim, hidden_dim, num_layers);
    if (backward_buffer == nullptr)                                      // callers that do not want dL/dact (the autograd wrapper): the tail of scratch
        backward_buffer = reinterpret_cast<uint16_t*>(static_cast<uint8_t*>(scratch) + ((n * sizeof(float) + 255) & ~(size_t)255));
    if (cudaError_t e = cudaMemsetAsync(scratch, 0, n * sizeof(float), st)) return (int)e;
    GenParams p{};
    p.inputs = reinterpret_cast<const __half*>(inputs); p.weights = reinterpret_cast<const __half*>(weights);
    p.grad = reinterpret_cast<const __half*>(grad); p.fwd_in = reinterpret_cast<const __half*>(forward_buffer);
    p.backward_buffer = reinterpret_cast<__half*>(backward_buffer); p.grad_inputs = reinterpret_cast<__half*>(grad_inputs);
    p.wsum = static_cast<float*>(scratch);
    p.B = B; p.in = input_dim; p.hidden = hidden_dim; p.layers = num_layers; p.out_dim = output_dim; p.act = activation; p.out_act = 6;
    p.calc_grad_inputs = calc_grad_inputs && grad_inputs != nullptr;
    const uint32_t smem = 2u * kRows * (hidden_dim + 2) * 2u;
    foc::launch(k_mlp_generic_bwd, B / kRows, kThreads, smem, st, p);
    FOC_LAUNCH_CHECK();
    if (grad_weights != nullptr) {
        foc::launch(k_mlp_generic_finish, ceil_div((uint32_t)n, 256u), 256, 0, st, static_cast<const float*>(scratch), (uint32_t)n, reinterpret_cast<__half*>(grad_weights), accumulate);
        FOC_LAUNCH_CHECK();
    }
    return 0;
}

int mlp_generic_finish(const void* scratch, uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers, uint16_t* grad_weights, int accumulate, cudaStream_t st) {
    const size_t n = mlp_generic_num_params(input_dim, hidden_dim, num_layers);
    foc::launch(k_mlp_generic_finish, ceil_div((uint32_t)n, 256u), 256, 0, st, static_cast<const float*>(scratch), (uint32_t)n, reinterpret_cast<__half*>(grad_weights), accumulate);
    FOC_LAUNCH_CHECK(); return 0;
}

}  // namespace foc
