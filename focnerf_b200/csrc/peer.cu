// peer.cu -- data-parallel gradient exchange for the train step over NVLink peer memory.
//
// The reference trains on one GPU; its multi-GPU story would be torch DDP: NCCL all-reduce of every gradient, then every
// rank runs the full Adam sweep over the 12.2 M-entry hash table.  On B200 + NVSwitch every GPU can load and store any
// peer's HBM directly, so the exchange is fused into the optimiser instead of bolted on in front of it:
//
//   barrier            "my backward is done" (all gradient tables complete, nobody reads the fp16 parameter copies)
//   k_dp_reduce        rank r PULLS shard r of the fp16 table gradient from every peer (16-byte P2P loads), sums them in
//                      fp32 in rank order (deterministic, the same bits whoever owns the shard), averages, checks for
//                      inf/nan and keeps the fp32 result locally.  The small MLP gradients (the "tail", ~10 k floats) are
//                      reduced in full by every rank -- replicated, bit-identical.
//   barrier(+inf)      exchanges the ranks' non-finite flags (GradScaler semantics: ANY bad rank skips the step everywhere)
//   k_adam_step        the ordinary sweep (optim.cu) over shard r only -- 1/N of the HBM traffic of the replicated sweep --
//                      and PUSHES the fp16 copy of the updated parameters into every peer's table (P2P stores)
//   barrier            "all shards delivered": the next forward may gather from the fp16 table
//
// Per rank and step that is (N-1)/N of the table in (pulls) and out (pushes) -- the traffic of a ring all-reduce -- but
// no staging copies, no NCCL channels holding SMs, and an optimiser sweep that shrinks with N.  The fp32 master
// parameters and Adam moments are sharded (rank r's shard is the only current one); `foc_peer_*` exposes the memory.
//
// Barriers are one-warp kernels that write an epoch into every peer's signal pad (st.release.sys) and spin on their own
// pad (ld.acquire.sys).  Epochs only grow, a peer can be at most one barrier ahead, so one word per source rank is
// enough.  A spin that sees no progress for ~2 s gives up and raises a sticky status word (never hangs the GPU).
#include "common.cuh"
#include "../../include/foc_b200.h"

using namespace foc;

namespace {

constexpr int kThreads = 256;
constexpr int kPerBlock = 4096;                    // elements per CTA: matches optim.cu, shard boundaries are multiples of it

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) { asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
    uint32_t v; asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
// peer data: never served from a stale L1 line
__device__ __forceinline__ uint4 ld_peer16(const void* p) {
    uint4 v; asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory"); return v;
}

// signal pad layout (uint32 words): [0..7] arrival epoch by source rank; [8..15], [16..23] inf bits by source rank, two slots
// local state: [0] epoch, [1] local non-finite flag (set by k_dp_reduce), [2] sticky status (1 = a barrier timed out)
__global__ void __launch_bounds__(32) k_peer_barrier(const __grid_constant__ FocPeerGroup g, float* __restrict__ found_inf) {
    const int lane = threadIdx.x;
    const uint32_t epoch = g.state[0] + 1;
    const bool exch = found_inf != nullptr;
    const uint32_t bit = g.state[1] | ((exch && *found_inf != 0.0f) ? 1u : 0u);     // k_dp_reduce's flag, or a local foc_adam_nonfinite result
    __syncwarp();
    __threadfence_system();
    uint32_t peer_bit = 0;
    if (lane < g.world) {
        uint32_t* pad = g.flags[lane];
        if (exch) st_relaxed_sys(pad + 8 + 8 * (epoch & 1u) + g.rank, bit);
        st_release_sys(pad + g.rank, epoch);
        const uint32_t* mine = g.flags[g.rank];
        const long long t0 = clock64();
        bool ok = true;
        while ((int32_t)(ld_acquire_sys(mine + lane) - epoch) < 0) {
            if (clock64() - t0 > 4000000000ll) { ok = false; break; }
            __nanosleep(40);
        }
        if (!ok) g.state[2] = 1;
        if (exch) peer_bit = ld_relaxed_sys(mine + 8 + 8 * (epoch & 1u) + lane);
    }
    const bool any = __any_sync(0xffffffffu, peer_bit != 0);
    __threadfence_system();
    if (lane == 0) {
        g.state[0] = epoch;
        if (exch) { if (any) *found_inf = 1.0f; g.state[1] = 0; }
    }
}

__device__ __forceinline__ void acc8(const uint4& v, float (&a)[8]) {
    const __half2* h = reinterpret_cast<const __half2*>(&v);
    #pragma unroll
    for (int k = 0; k < 4; ++k) { const float2 f = __half22float2(h[k]); a[2 * k] += f.x; a[2 * k + 1] += f.y; }
}

// blocks [0, shard_blocks): this rank's shard of the fp16 table gradient; blocks after that: the fp32 tail (replicated)
template <int W>
__global__ void __launch_bounds__(kThreads) k_dp_reduce(const __grid_constant__ FocDpReduce a) {
    const float inv = 1.0f / (float)W;
    bool ok = true;
    if (blockIdx.x < a.shard_blocks) {
        #pragma unroll
        for (int u = 0; u < kPerBlock / (kThreads * 8); ++u) {
            const uint64_t i = (uint64_t)blockIdx.x * kPerBlock + ((uint64_t)u * kThreads + threadIdx.x) * 8;   // within the shard
            if (i >= a.shard_n) break;                                  // shard_n is a multiple of 8 (checked on the host)
            uint4 v[W];
            #pragma unroll
            for (int r = 0; r < W; ++r) v[r] = ld_peer16(a.grad16[r] + a.shard_start + i);
            float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            #pragma unroll
            for (int r = 0; r < W; ++r) acc8(v[r], s);
            #pragma unroll
            for (int k = 0; k < 8; ++k) { s[k] *= inv; ok = ok && (fabsf(s[k]) <= 3.402823466e+38f); }
            float4* out = reinterpret_cast<float4*>(a.reduced + i);
            out[0] = make_float4(s[0], s[1], s[2], s[3]);
            out[1] = make_float4(s[4], s[5], s[6], s[7]);
        }
    } else {
        const uint64_t base = (uint64_t)(blockIdx.x - a.shard_blocks) * kPerBlock;
        #pragma unroll
        for (int u = 0; u < kPerBlock / (kThreads * 4); ++u) {
            const uint64_t i = base + ((uint64_t)u * kThreads + threadIdx.x) * 4;
            if (i >= a.tail_n) break;                                   // tail_n is a multiple of 4
            uint4 v[W];
            #pragma unroll
            for (int r = 0; r < W; ++r) v[r] = ld_peer16(a.tail[r] + i);
            float s[4] = {0.f, 0.f, 0.f, 0.f};
            #pragma unroll
            for (int r = 0; r < W; ++r) { s[0] += __uint_as_float(v[r].x); s[1] += __uint_as_float(v[r].y); s[2] += __uint_as_float(v[r].z); s[3] += __uint_as_float(v[r].w); }
            #pragma unroll
            for (int k = 0; k < 4; ++k) { s[k] *= inv; ok = ok && (fabsf(s[k]) <= 3.402823466e+38f); }
            *reinterpret_cast<float4*>(a.reduced_tail + i) = make_float4(s[0], s[1], s[2], s[3]);
        }
    }
    if (!ok) *a.inf_flag = 1u;
}

// Epilogue of an inference render (legacy/nerf/renderer.py:369-372: background mix, depth normalisation), fused with the
// packing into [r, g, b, depth, weights_sum] rows and -- when the frame's rays were dealt to the ranks in kDeal-ray blocks
// (dist.cyclic_tile: block b -> rank b % world) -- with the exchange: one CTA per local block stages its 128 rows
// (2560 contiguous bytes of the frame) in shared memory and stores them, 16 bytes per thread, into EVERY rank's frame
// buffer over peer memory.  Arithmetic = the reference's separate torch ops (no fma contraction).
constexpr uint32_t kDeal = 128;

__global__ void __launch_bounds__(kDeal) k_render_finish(const float* __restrict__ image, const float* __restrict__ depth, const float* __restrict__ weights_sum,
                                                         const float* __restrict__ nears, const float* __restrict__ fars, uint32_t N, float bg,
                                                         int depth_from_first_sample, const __grid_constant__ FocFrameOut o) {
    __shared__ __align__(16) float rows[kDeal * 5];
    const uint32_t i = blockIdx.x * kDeal + threadIdx.x;
    if (i < N) {
        const float ws = weights_sum[i], near = nears[i], far = fars[i];
        const float mix = __fmul_rn(__fsub_rn(1.0f, ws), bg);
        float d = depth[i];
        if (depth_from_first_sample) d = __fadd_rn(d, __fmul_rn(ws > 0.0f ? near : 0.0f, ws));
        d = __fdiv_rn(fmaxf(__fsub_rn(d, near), 0.0f), __fsub_rn(far, near));
        float* r = rows + 5 * threadIdx.x;
        r[0] = __fadd_rn(image[3 * (size_t)i], mix); r[1] = __fadd_rn(image[3 * (size_t)i + 1], mix); r[2] = __fadd_rn(image[3 * (size_t)i + 2], mix);
        r[3] = d; r[4] = ws;
    }
    __syncthreads();
    const uint32_t g0 = ((uint32_t)blockIdx.x * (uint32_t)o.world + (uint32_t)o.rank) * kDeal;       // first frame row of this block
    if (g0 >= o.n_total) return;                                                                       // a padding block of the deal
    const uint32_t valid = min(min(kDeal, o.n_total - g0), N - blockIdx.x * kDeal);                    // rows of this block inside the frame
    const float4* src = reinterpret_cast<const float4*>(rows);
    for (uint32_t q = threadIdx.x; q < (valid * 5 + 3) / 4; q += kDeal) {
        const bool whole = (q + 1) * 4 <= valid * 5;
        #pragma unroll
        for (int r = 0; r < FOC_MAX_PEERS; ++r) {
            if (r >= o.world) break;
            float* dst = o.frame[r] + (size_t)g0 * 5;
            if (whole) reinterpret_cast<float4*>(dst)[q] = src[q];
            else for (uint32_t k = q * 4; k < valid * 5; ++k) dst[k] = rows[k];
        }
    }
    if (o.world > 1) { __syncthreads(); if (threadIdx.x == 0) __threadfence_system(); }      // one fence per CTA (see k_adam_step)
}

inline bool valid_group(const FocPeerGroup* g) {
    if (g == nullptr || g->world < 1 || g->world > FOC_MAX_PEERS || g->rank < 0 || g->rank >= g->world || g->state == nullptr) return false;
    for (int r = 0; r < g->world; ++r) if (g->flags[r] == nullptr) return false;
    return true;
}

}  // namespace

// ---- peer-visible memory: plain cudaMalloc (IPC-exportable), handles travel between the ranks' processes out of band
FOC_API int foc_peer_alloc(uint64_t bytes, void** ptr) {
    if (ptr == nullptr || bytes == 0) return FOC_ERR_UNSUPPORTED;
    cudaError_t e = cudaMalloc(ptr, bytes);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaMemset(*ptr, 0, bytes);
}
FOC_API int foc_peer_free(void* ptr) { return (int)cudaFree(ptr); }
FOC_API int foc_peer_export(void* ptr, uint8_t* handle64) {
    static_assert(sizeof(cudaIpcMemHandle_t) == FOC_PEER_HANDLE_BYTES, "handle size");
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, ptr);
    if (e != cudaSuccess) return (int)e;
    memcpy(handle64, &h, sizeof h);
    return 0;
}
FOC_API int foc_peer_open(const uint8_t* handle64, void** ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof h);
    return (int)cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
}
FOC_API int foc_peer_close(void* ptr) { return (int)cudaIpcCloseMemHandle(ptr); }

FOC_API int foc_peer_barrier(const FocPeerGroup* group, float* found_inf, void* stream) {
    if (!valid_group(group)) return FOC_ERR_UNSUPPORTED;
    k_peer_barrier<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(*group, found_inf);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_dp_reduce(const FocDpReduce* args, void* stream) {
    if (args == nullptr || args->world < 1 || args->world > FOC_MAX_PEERS) return FOC_ERR_UNSUPPORTED;
    FocDpReduce a = *args;
    if ((a.shard_n % 8) || (a.shard_start % 8) || (a.tail_n % 4) || a.inf_flag == nullptr) return FOC_ERR_UNSUPPORTED;
    if (a.shard_n && a.reduced == nullptr) return FOC_ERR_UNSUPPORTED;
    if (a.tail_n && a.reduced_tail == nullptr) return FOC_ERR_UNSUPPORTED;
    for (int r = 0; r < a.world; ++r) {
        if ((a.shard_n && (a.grad16[r] == nullptr || (reinterpret_cast<uintptr_t>(a.grad16[r]) & 15u))) ||
            (a.tail_n && (a.tail[r] == nullptr || (reinterpret_cast<uintptr_t>(a.tail[r]) & 15u)))) return FOC_ERR_UNSUPPORTED;
    }
    a.shard_blocks = (uint32_t)((a.shard_n + kPerBlock - 1) / kPerBlock);
    const uint32_t grid = a.shard_blocks + (uint32_t)((a.tail_n + kPerBlock - 1) / kPerBlock);
    if (grid == 0) return 0;
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    switch (a.world) {
        case 1: k_dp_reduce<1><<<grid, kThreads, 0, s>>>(a); break;
        case 2: k_dp_reduce<2><<<grid, kThreads, 0, s>>>(a); break;
        case 3: k_dp_reduce<3><<<grid, kThreads, 0, s>>>(a); break;
        case 4: k_dp_reduce<4><<<grid, kThreads, 0, s>>>(a); break;
        case 5: k_dp_reduce<5><<<grid, kThreads, 0, s>>>(a); break;
        case 6: k_dp_reduce<6><<<grid, kThreads, 0, s>>>(a); break;
        case 7: k_dp_reduce<7><<<grid, kThreads, 0, s>>>(a); break;
        default: k_dp_reduce<8><<<grid, kThreads, 0, s>>>(a); break;
    }
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_render_finish(const float* image, const float* depth, const float* weights_sum, const float* nears, const float* fars,
                              uint32_t N, float bg, int depth_from_first_sample, const FocFrameOut* out, void* stream) {
    if (N == 0) return 0;
    if (out == nullptr || out->world < 1 || out->world > FOC_MAX_PEERS || out->rank < 0 || out->rank >= out->world) return FOC_ERR_UNSUPPORTED;
    for (int r = 0; r < out->world; ++r)
        if (out->frame[r] == nullptr || (reinterpret_cast<uintptr_t>(out->frame[r]) & 15u)) return FOC_ERR_UNSUPPORTED;
    k_render_finish<<<ceil_div(N, kDeal), kDeal, 0, reinterpret_cast<cudaStream_t>(stream)>>>(image, depth, weights_sum, nears, fars, N, bg,
                                                                                             depth_from_first_sample, *out);
    FOC_LAUNCH_CHECK(); return 0;
}
