// optim.cu -- dense Adam + dynamic loss scaling for the field's parameters, fused.
//
// Replaces, for this path, what the reference's train step runs after backward (nerf/utils.py:866-868 with
// main_nerf.py:211,472): GradScaler's non-finite check (4 multi-tensor launches), the half->float copy of the table
// gradient, torch's fused Adam (one launch per parameter group), the next step's `embeddings.to(half)` and
// `zeros_like(grad)` -- ~250 us of a 0.9 ms step on B200, almost all of it HBM traffic over the 12.2 M-entry table.
// Here the table is swept ONCE: read p, m, v (fp32) and the fp16 gradient, write p, m, v, the fp16 shadow the next
// forward gathers from, and zero the gradient in place: 30 B / parameter, the HBM floor for a dense Adam.
//
// The arithmetic is torch.optim.Adam's (torch/optim/adam.py, fused=True, amsgrad=False, weight_decay=0, maximize=False):
//     g = grad / scale;  m += (1-b1)(g - m);  v = b2 v + (1-b2) g^2;
//     p -= (lr / (1-b1^t)) * m / (sqrt(v) / sqrt(1-b2^t) + eps)            t = successful steps so far + 1
// and the scale follows torch.amp.GradScaler.update(): a step whose gradients hold an inf/nan is skipped entirely
// (parameters, moments and t untouched) and the scale is multiplied by `backoff`; `growth_interval` clean steps in a
// row multiply it by `growth`.  Everything lives on the device: no host synchronisation, CUDA-graph capturable.
#include <cstdlib>
#include "common.cuh"
#include "../../include/foc_b200.h"

using namespace foc;

namespace {

constexpr int kMaxTensors = 8;
constexpr int kThreads = 256, kVec = 4, kUnroll = 4;
constexpr int kPerBlock = kThreads * kVec * kUnroll;               // 4096 parameters per CTA

struct AdamList {
    FocAdamTensor t[kMaxTensors];
    uint32_t block_start[kMaxTensors + 1];                         // CTA ranges of the tensors, concatenated
    int count;
};
// data-parallel sweep (peer.cu): tensor `tensor`'s fp16 copy goes to every rank's table instead of the local one only
struct PeerShadow {
    uint16_t* base[FOC_MAX_PEERS];
    int world, tensor;
    int self;                                  // >= 0 (foc_adam_step_dp): this rank's index in base[] -- unchanged fp16 values are not pushed
};

// foc_adam_step_dp: where the gradients of a data-parallel step come from -- every rank's buffers, read over peer memory
struct PeerGrad {
    const uint16_t* g16[FOC_MAX_PEERS];        // fp16 table gradients, already offset to THIS rank's shard
    const float* tail[FOC_MAX_PEERS];          // flat fp32 gradients of the replicated tensors; tensors[i].grad points into tail[rank]
    int rank;
};
// peer data: never served from a stale L1 line
__device__ __forceinline__ uint2 ld_peer8(const void* p) {
    uint2 v; asm volatile("ld.relaxed.sys.global.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ uint4 ld_peer16(const void* p) {
    uint4 v; asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory"); return v;
}
// mean over the W ranks, summed in fp32 in rank order: the same bits on whichever rank owns the element (k_dp_reduce's arithmetic)
template <int W>
__device__ __forceinline__ void mean_table4(const uint2 (&v)[W], float (&g)[4]) {
    g[0] = g[1] = g[2] = g[3] = 0.f;
    #pragma unroll
    for (int r = 0; r < W; ++r) {
        const __half2 a = *reinterpret_cast<const __half2*>(&v[r].x), b = *reinterpret_cast<const __half2*>(&v[r].y);
        g[0] += __low2float(a); g[1] += __high2float(a); g[2] += __low2float(b); g[3] += __high2float(b);
    }
    const float inv = 1.0f / (float)W;
    #pragma unroll
    for (int k = 0; k < 4; ++k) g[k] *= inv;
}
template <int W>
__device__ __forceinline__ void pull_tail4(const PeerGrad& G, const FocAdamTensor& T, uint64_t i, float (&g)[4]) {
    const uint64_t off = (uint64_t)(static_cast<const float*>(T.grad) - G.tail[G.rank]) + i;     // rows start on 4-float boundaries, padded with zeros
    uint4 v[W];
    #pragma unroll
    for (int r = 0; r < W; ++r) v[r] = ld_peer16(G.tail[r] + off);
    g[0] = g[1] = g[2] = g[3] = 0.f;
    #pragma unroll
    for (int r = 0; r < W; ++r) { g[0] += __uint_as_float(v[r].x); g[1] += __uint_as_float(v[r].y); g[2] += __uint_as_float(v[r].z); g[3] += __uint_as_float(v[r].w); }
    const float inv = 1.0f / (float)W;
    #pragma unroll
    for (int k = 0; k < 4; ++k) g[k] *= inv;
}

__device__ __forceinline__ int find_tensor(const AdamList& L, uint32_t block) {
    int i = 0;
    #pragma unroll
    for (int k = 1; k < kMaxTensors; ++k) if (k < L.count && block >= L.block_start[k]) i = k;
    return i;
}

__device__ __forceinline__ bool finite4(float a, float b, float c, float d) {
    // |x| < inf for all four  <=>  none is inf or nan
    return (fabsf(a) <= 3.402823466e+38f) && (fabsf(b) <= 3.402823466e+38f) && (fabsf(c) <= 3.402823466e+38f) && (fabsf(d) <= 3.402823466e+38f);
}

__device__ __forceinline__ void load_grad4(const FocAdamTensor& T, uint64_t i, uint64_t n, float (&g)[4]) {
    if (T.grad_dtype == 1) {
        const __half* p = static_cast<const __half*>(T.grad) + i;
        if (i + 4 <= n) {
            const uint2 v = *reinterpret_cast<const uint2*>(p);
            const __half2 a = *reinterpret_cast<const __half2*>(&v.x), b = *reinterpret_cast<const __half2*>(&v.y);
            g[0] = __low2float(a); g[1] = __high2float(a); g[2] = __low2float(b); g[3] = __high2float(b);
        } else {
            #pragma unroll
            for (int k = 0; k < 4; ++k) g[k] = (i + k < n) ? __half2float(p[k]) : 0.0f;
        }
    } else {
        const float* p = static_cast<const float*>(T.grad) + i;
        if (i + 4 <= n) { const float4 v = *reinterpret_cast<const float4*>(p); g[0] = v.x; g[1] = v.y; g[2] = v.z; g[3] = v.w; }
        else {
            #pragma unroll
            for (int k = 0; k < 4; ++k) g[k] = (i + k < n) ? p[k] : 0.0f;
        }
    }
}

// found_inf = 1 when any gradient element is inf / nan (never cleared here).  16-byte loads: 8 halfs / 4 floats per access.
__global__ void __launch_bounds__(kThreads) k_adam_nonfinite(const __grid_constant__ AdamList L, float* __restrict__ found_inf) {
    pdl_sync();
    const int ti = find_tensor(L, blockIdx.x);
    const FocAdamTensor& T = L.t[ti];
    if (T.grad == nullptr) return;
    const uint64_t base = (uint64_t)(blockIdx.x - L.block_start[ti]) * kPerBlock;
    const uint64_t end = min(T.n, base + (uint64_t)kPerBlock);
    bool ok = true;
    if (T.grad_dtype == 1) {
        const __half* g = static_cast<const __half*>(T.grad);
        const uint64_t i0 = base + (uint64_t)threadIdx.x * 8;
        constexpr int kIt = kPerBlock / (kThreads * 8);
        // all loads of the thread first (independent, in flight together), then branch-free tests: a short-circuit `&&` per word made
        // every load wait for the previous one's verdict
        uint4 v[kIt];
        #pragma unroll
        for (int u = 0; u < kIt; ++u) {
            const uint64_t i = i0 + (uint64_t)u * kThreads * 8;
            v[u] = (i + 8 <= end) ? *reinterpret_cast<const uint4*>(g + i) : make_uint4(0u, 0u, 0u, 0u);     // default policy: the sweep reads it next
        }
        uint32_t bad = 0;
        #pragma unroll
        for (int u = 0; u < kIt; ++u) {
            const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
            #pragma unroll
            for (int k = 0; k < 4; ++k)             // exponent field all ones <=> inf / nan, per 16-bit half
                bad |= (uint32_t)((w[k] & 0x7c00u) == 0x7c00u) | (uint32_t)((w[k] & 0x7c000000u) == 0x7c000000u);
            const uint64_t i = i0 + (uint64_t)u * kThreads * 8;
            if (i < end && i + 8 > end) { for (uint64_t j = i; j < end; ++j) bad |= (uint32_t)((__half_as_ushort(g[j]) & 0x7c00u) == 0x7c00u); }
        }
        ok = bad == 0;
    } else {
        const float* g = static_cast<const float*>(T.grad);
        uint64_t i = base + (uint64_t)threadIdx.x * 4;
        #pragma unroll
        for (int u = 0; u < kPerBlock / (kThreads * 4); ++u, i += (uint64_t)kThreads * 4) {
            if (i + 4 <= end) { const float4 v = *reinterpret_cast<const float4*>(g + i); ok = ok && finite4(v.x, v.y, v.z, v.w); }
            else { for (uint64_t j = i; j < end && j < i + 4; ++j) ok = ok && (fabsf(g[j]) <= 3.402823466e+38f); }
        }
    }
    if (!ok) *found_inf = 1.0f;
}

// PULL = W > 0 (foc_adam_step_dp): the gradient exchange of a W-rank data-parallel step happens HERE -- each element's gradient is the
// mean of the W ranks' values, loaded over peer memory at the point of use, so the shard is swept once (pull -> fp32 mean -> Adam ->
// push) instead of k_dp_reduce writing an fp32 copy that this kernel reads back.
template <bool PEERS, int PULL>
__global__ void __launch_bounds__(kThreads) k_adam_step(const __grid_constant__ AdamList L, const float* __restrict__ lr_p, float beta1, float beta2,
                                                        float eps, const float* __restrict__ scale, const float* __restrict__ found_inf,
                                                        const int32_t* __restrict__ step, const __grid_constant__ PeerShadow PS,
                                                        const __grid_constant__ PeerGrad PG) {
    pdl_sync();
    const bool bad = found_inf != nullptr && *found_inf != 0.0f;
    __shared__ float s_step_size, s_rsqrt_bc2, s_inv_scale;
    // the three per-step scalars, one per warp so that the two double-precision pow() run side by side (every CTA pays this prologue)
    if (threadIdx.x == 0) {
        const double bc1 = 1.0 - pow((double)beta1, (double)(*step + 1));
        s_step_size = (float)((double)*lr_p / bc1);
    } else if (threadIdx.x == 32) {
        const double bc2 = 1.0 - pow((double)beta2, (double)(*step + 1));
        s_rsqrt_bc2 = (float)(1.0 / sqrt(bc2));
    } else if (threadIdx.x == 64) {
        s_inv_scale = scale ? (float)(1.0 / (double)*scale) : 1.0f;
    }
    __syncthreads();
    const float step_size = s_step_size, rs = s_rsqrt_bc2, inv_scale = s_inv_scale, w1 = 1.0f - beta1, w2 = 1.0f - beta2;
    // block-stride over the 4096-parameter blocks (the grid covers them all by default; a smaller persistent grid was
    // measured slower: 2 CTAs per SM do not keep enough loads in flight for the HBM-bound sweep)
    for (uint32_t vb = blockIdx.x; vb < L.block_start[kMaxTensors]; vb += gridDim.x) {
    const int ti = find_tensor(L, vb);
    const FocAdamTensor& T = L.t[ti];
    const uint64_t base = (uint64_t)(vb - L.block_start[ti]) * kPerBlock;
    const bool skip = bad || T.grad == nullptr;
    // PULL: every peer load of this thread's kUnroll iterations is issued before the first one is used -- NVLink latency (~2 us) is paid
    // once per CTA, not once per iteration (a shard is at most a wave or two of CTAs, so nothing else would hide it)
    uint2 pre[PULL > 0 ? kUnroll : 1][PULL > 0 ? PULL : 1];
    if constexpr (PULL > 0) {
        if (!skip && ti == PS.tensor) {
            #pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                const uint64_t i = base + ((uint64_t)u * kThreads + threadIdx.x) * kVec;
                if (i < T.n) {
                    #pragma unroll
                    for (int r = 0; r < PULL; ++r) pre[u][r] = ld_peer8(PG.g16[r] + i);
                }
            }
        }
    }
    #pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
        const uint64_t i = base + ((uint64_t)u * kThreads + threadIdx.x) * kVec;
        if (i >= T.n) break;
        const bool full = i + 4 <= T.n;
        if (!skip) {
            float g[4], p[4], m[4], v[4];
            if constexpr (PULL > 0) {
                if (ti == PS.tensor) mean_table4<PULL>(pre[u], g);    // the shard is a multiple of 8 elements: never a partial vector
                else pull_tail4<PULL>(PG, T, i, g);
            } else load_grad4(T, i, T.n, g);
            if (full) {
                // the fp32 master values and moments stream through once per step (294 MB against a 126 MB L2): evict-first loads and
                // stores, so that what the NEXT step gathers from / scatters into -- the fp16 shadow and the zeroed fp16 gradient written
                // below with the default policy, 49 MB together -- is still L2-resident when its forward starts
                const float4 P = __ldcs(reinterpret_cast<const float4*>(T.param + i)), M = __ldcs(reinterpret_cast<const float4*>(T.exp_avg + i)),
                             V = __ldcs(reinterpret_cast<const float4*>(T.exp_avg_sq + i));
                p[0] = P.x; p[1] = P.y; p[2] = P.z; p[3] = P.w; m[0] = M.x; m[1] = M.y; m[2] = M.z; m[3] = M.w; v[0] = V.x; v[1] = V.y; v[2] = V.z; v[3] = V.w;
            } else {
                #pragma unroll
                for (int k = 0; k < 4; ++k) { const bool in = i + k < T.n; p[k] = in ? T.param[i + k] : 0.f; m[k] = in ? T.exp_avg[i + k] : 0.f; v[k] = in ? T.exp_avg_sq[i + k] : 0.f; }
            }
            #pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float gk = g[k] * inv_scale;
                m[k] = fmaf(w1, gk - m[k], m[k]);
                v[k] = fmaf(beta2, v[k], w2 * gk * gk);
                const float denom = fmaf(sqrtf(v[k]), rs, eps);
                p[k] = p[k] - step_size * (m[k] / denom);
            }
            if (full) {
                __stcs(reinterpret_cast<float4*>(T.param + i), make_float4(p[0], p[1], p[2], p[3]));
                __stcs(reinterpret_cast<float4*>(T.exp_avg + i), make_float4(m[0], m[1], m[2], m[3]));
                __stcs(reinterpret_cast<float4*>(T.exp_avg_sq + i), make_float4(v[0], v[1], v[2], v[3]));
                const __half2 a = __floats2half2_rn(p[0], p[1]), b = __floats2half2_rn(p[2], p[3]);
                const uint2 packed = make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
                if (PEERS && ti == PS.tensor) {
                    // EXPERIMENT (FOC_DP_PUSH_SKIP=1, off by default): every rank's table holds the same fp16 values, so a vector whose four
                    // fp16 values did not change (lr * m / sqrt(v) below half an fp16 ulp; untouched entries) needs no store.  Measured at
                    // N = 8: the sweep 87.0 -> 81.8 us, but the step 0.5605 -> 0.5688 ms (the ranks finish less evenly and wait longer at the
                    // barrier that follows); N = 2: 0.5634 -> 0.5702 ms.  Bit-identical replicas either way.
                    bool push = true;
                    if (PS.self >= 0) { const uint2 old = *reinterpret_cast<const uint2*>(PS.base[PS.self] + i); push = old.x != packed.x || old.y != packed.y; }
                    if (push) {
                        #pragma unroll
                        for (int r = 0; r < FOC_MAX_PEERS; ++r) if (r < PS.world) *reinterpret_cast<uint2*>(PS.base[r] + i) = packed;
                    }
                } else if (T.shadow16) *reinterpret_cast<uint2*>(T.shadow16 + i) = packed;
            } else {
                for (int k = 0; k < 4 && i + k < T.n; ++k) {
                    T.param[i + k] = p[k]; T.exp_avg[i + k] = m[k]; T.exp_avg_sq[i + k] = v[k];
                    const __half h = __float2half_rn(p[k]);
                    if (PEERS && ti == PS.tensor) { for (int r = 0; r < PS.world; ++r) reinterpret_cast<__half*>(PS.base[r])[i + k] = h; }
                    else if (T.shadow16) reinterpret_cast<__half*>(T.shadow16)[i + k] = h;
                }
            }
        }
        if (T.zero_grad && T.grad != nullptr) {
            if (T.grad_dtype == 1) {
                __half* gp = static_cast<__half*>(T.grad) + i;
                if (full) *reinterpret_cast<uint2*>(gp) = make_uint2(0u, 0u);
                else for (int k = 0; k < 4 && i + k < T.n; ++k) gp[k] = __float2half_rn(0.f);
            } else {
                float* gp = static_cast<float*>(T.grad) + i;
                if (full) *reinterpret_cast<float4*>(gp) = make_float4(0.f, 0.f, 0.f, 0.f);
                else for (int k = 0; k < 4 && i + k < T.n; ++k) gp[k] = 0.f;
            }
        }
    }
    }
    // the pushed fp16 copies must be visible to the peers before the barrier that follows: ONE system-scope fence per CTA (thread 0, after
    // a CTA barrier -- fence cumulativity covers the other threads' stores, the pattern of a cooperative-groups grid sync) instead of one
    // per thread: thousands of concurrent MEMBAR.SYS, each waiting for NVLink acknowledgements, kept the CTA slots occupied
    if (PEERS) { __syncthreads(); if (threadIdx.x == 0) __threadfence_system(); }
}

// torch.amp.GradScaler.update() (torch/amp/grad_scaler.py, amp_update_scale_cuda_kernel) + the step counter
// lr != nullptr: also the reference's per-iteration LambdaLR (main_nerf.py:477, scheduler_update_every_step=True):
//     lr(next iteration) = lr0 * 0.1 ** min(iteration / decay_iters, 1), iteration counted whether or not the step was skipped
__global__ void k_scaler_update(float* __restrict__ scale, int32_t* __restrict__ growth_tracker, float* __restrict__ found_inf,
                                int32_t* __restrict__ step, float growth, float backoff, int32_t interval,
                                float* __restrict__ lr, int32_t* __restrict__ sched_iter, float lr0, float decay_iters) {
    pdl_sync();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (lr != nullptr) {
        const int32_t it = *sched_iter + 1;
        *sched_iter = it;
        *lr = (float)((double)lr0 * pow(0.1, fmin((double)it / (double)decay_iters, 1.0)));
    }
    const bool bad = found_inf != nullptr && *found_inf != 0.0f;
    if (scale != nullptr) {
        if (bad) { *scale = *scale * backoff; *growth_tracker = 0; }
        else {
            const int32_t ok = *growth_tracker + 1;
            if (ok == interval) {
                const float grown = *scale * growth;
                if (fabsf(grown) <= 3.402823466e+38f) *scale = grown;         // never grow into inf
                *growth_tracker = 0;
            } else *growth_tracker = ok;
        }
    }
    if (!bad) *step = *step + 1;
    if (found_inf != nullptr) *found_inf = 0.0f;
}

int make_list(const FocAdamTensor* tensors, int count, AdamList& L) {
    if (count <= 0 || count > kMaxTensors) return FOC_ERR_UNSUPPORTED;
    L.count = count;
    uint64_t blocks = 0;
    for (int i = 0; i < count; ++i) {
        const FocAdamTensor& t = tensors[i];
        // 16-byte vector accesses: base pointers must be 16-byte aligned (torch allocations are)
        if ((reinterpret_cast<uintptr_t>(t.param) | reinterpret_cast<uintptr_t>(t.exp_avg) | reinterpret_cast<uintptr_t>(t.exp_avg_sq) |
             reinterpret_cast<uintptr_t>(t.grad) | reinterpret_cast<uintptr_t>(t.shadow16)) & 15u) return FOC_ERR_UNSUPPORTED;
        if (t.grad_dtype != 0 && t.grad_dtype != 1) return FOC_ERR_UNSUPPORTED;
        L.t[i] = t;
        L.block_start[i] = (uint32_t)blocks;
        blocks += (t.n + kPerBlock - 1) / kPerBlock;
    }
    for (int i = count; i <= kMaxTensors; ++i) L.block_start[i] = (uint32_t)blocks;
    if (blocks == 0 || blocks > 0x7fffffffull) return FOC_ERR_UNSUPPORTED;
    return 0;
}

}  // namespace

FOC_API int foc_adam_nonfinite(const FocAdamTensor* tensors, int count, float* found_inf, void* stream) {
    AdamList L;
    if (int rc = make_list(tensors, count, L)) return rc;
    foc::launch(k_adam_nonfinite, L.block_start[kMaxTensors], kThreads, 0, reinterpret_cast<cudaStream_t>(stream), L, found_inf);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_adam_step(const FocAdamTensor* tensors, int count, const float* lr, float beta1, float beta2, float eps,
                          const float* scale, const float* found_inf, const int32_t* step, void* stream) {
    AdamList L;
    if (int rc = make_list(tensors, count, L)) return rc;
    const uint32_t adam_grid = L.block_start[kMaxTensors];
    PeerShadow ps{};
    foc::launch(k_adam_step<false, 0>, adam_grid, kThreads, 0, reinterpret_cast<cudaStream_t>(stream), L, lr, beta1, beta2, eps, scale, found_inf, step, ps, PeerGrad{});
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_adam_step_peers(const FocAdamTensor* tensors, int count, const float* lr, float beta1, float beta2, float eps,
                                const float* scale, const float* found_inf, const int32_t* step, int peer_tensor, int world,
                                uint16_t* const* peer_shadow16, void* stream) {
    AdamList L;
    if (int rc = make_list(tensors, count, L)) return rc;
    if (world < 1 || world > FOC_MAX_PEERS || peer_tensor < 0 || peer_tensor >= count || peer_shadow16 == nullptr) return FOC_ERR_UNSUPPORTED;
    PeerShadow ps{};
    ps.world = world; ps.tensor = peer_tensor; ps.self = -1;
    for (int r = 0; r < world; ++r) {
        if (peer_shadow16[r] == nullptr || (reinterpret_cast<uintptr_t>(peer_shadow16[r]) & 15u)) return FOC_ERR_UNSUPPORTED;
        ps.base[r] = peer_shadow16[r];
    }
    foc::launch(k_adam_step<true, 0>, L.block_start[kMaxTensors], kThreads, 0, reinterpret_cast<cudaStream_t>(stream), L, lr, beta1, beta2, eps, scale,
                found_inf, step, ps, PeerGrad{});
    FOC_LAUNCH_CHECK(); return 0;
}

template <int W>
static void launch_dp(const AdamList& L, const float* lr, float beta1, float beta2, float eps, const float* scale, const float* found_inf,
                      const int32_t* step, const PeerShadow& ps, const PeerGrad& pg, void* stream) {
    foc::launch(k_adam_step<true, W>, L.block_start[kMaxTensors], kThreads, 0, reinterpret_cast<cudaStream_t>(stream), L, lr, beta1, beta2, eps, scale,
                found_inf, step, ps, pg);
}

FOC_API int foc_adam_step_dp(const FocAdamTensor* tensors, int count, const float* lr, float beta1, float beta2, float eps,
                             const float* scale, const float* found_inf, const int32_t* step, int peer_tensor, const FocDpGrads* grads,
                             uint16_t* const* peer_shadow16, void* stream) {
    AdamList L;
    if (int rc = make_list(tensors, count, L)) return rc;
    if (grads == nullptr || grads->world < 1 || grads->world > FOC_MAX_PEERS || grads->rank < 0 || grads->rank >= grads->world ||
        peer_tensor < 0 || peer_tensor >= count || peer_shadow16 == nullptr) return FOC_ERR_UNSUPPORTED;
    if (tensors[peer_tensor].n % 8u) return FOC_ERR_UNSUPPORTED;
    PeerShadow ps{};
    PeerGrad pg{};
    ps.world = grads->world; ps.tensor = peer_tensor; pg.rank = grads->rank;
    ps.self = getenv("FOC_DP_PUSH_SKIP") ? grads->rank : -1;                     // FOC_DP_PUSH_SKIP=1: do not push unchanged vectors (measured: not a win)
    for (int r = 0; r < grads->world; ++r) {
        if (peer_shadow16[r] == nullptr || grads->grad16[r] == nullptr || grads->tail[r] == nullptr) return FOC_ERR_UNSUPPORTED;
        if ((reinterpret_cast<uintptr_t>(peer_shadow16[r]) | reinterpret_cast<uintptr_t>(grads->grad16[r]) | reinterpret_cast<uintptr_t>(grads->tail[r])) & 15u) return FOC_ERR_UNSUPPORTED;
        ps.base[r] = peer_shadow16[r]; pg.g16[r] = grads->grad16[r]; pg.tail[r] = grads->tail[r];
    }
    for (int i = 0; i < count; ++i) {                                   // every tensor but the table takes its gradient from the fp32 tails
        if (i == peer_tensor) continue;
        const float* g = static_cast<const float*>(tensors[i].grad);
        if (g == nullptr || tensors[i].grad_dtype != 0 || g < pg.tail[pg.rank] || ((g - pg.tail[pg.rank]) & 3)) return FOC_ERR_UNSUPPORTED;
    }
    switch (grads->world) {
        case 1: launch_dp<1>(L, lr, beta1, beta2, eps, scale, found_inf, step, ps, pg, stream); break;
        case 2: launch_dp<2>(L, lr, beta1, beta2, eps, scale, found_inf, step, ps, pg, stream); break;
        case 3: launch_dp<3>(L, lr, beta1, beta2, eps, scale, found_inf, step, ps, pg, stream); break;
        case 4: launch_dp<4>(L, lr, beta1, beta2, eps, scale, found_inf, step, ps, pg, stream); break;
        case 5: launch_dp<5>(L, lr, beta1, beta2, eps, scale, found_inf, step, ps, pg, stream); break;
        case 6: launch_dp<6>(L, lr, beta1, beta2, eps, scale, found_inf, step, ps, pg, stream); break;
        case 7: launch_dp<7>(L, lr, beta1, beta2, eps, scale, found_inf, step, ps, pg, stream); break;
        default: launch_dp<8>(L, lr, beta1, beta2, eps, scale, found_inf, step, ps, pg, stream); break;
    }
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_scaler_update(float* scale, int32_t* growth_tracker, float* found_inf, int32_t* step, float growth_factor,
                              float backoff_factor, int32_t growth_interval, void* stream) {
    foc::launch(k_scaler_update, 1, 32, 0, reinterpret_cast<cudaStream_t>(stream), scale, growth_tracker, found_inf, step, growth_factor, backoff_factor,
                growth_interval, (float*)nullptr, (int32_t*)nullptr, 0.f, 1.f);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_scaler_update_sched(float* scale, int32_t* growth_tracker, float* found_inf, int32_t* step, float growth_factor,
                                    float backoff_factor, int32_t growth_interval, float* lr, int32_t* sched_iter, float lr0,
                                    float decay_iters, void* stream) {
    if (lr == nullptr || sched_iter == nullptr || !(decay_iters > 0.f)) return FOC_ERR_UNSUPPORTED;
    foc::launch(k_scaler_update, 1, 32, 0, reinterpret_cast<cudaStream_t>(stream), scale, growth_tracker, found_inf, step, growth_factor, backoff_factor,
                growth_interval, lr, sched_iter, lr0, decay_iters);
    FOC_LAUNCH_CHECK(); return 0;
}
