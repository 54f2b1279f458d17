// mlp_fwd.cu -- fully fused MLP forward / inference on the 5th-generation tensor cores (tcgen05 + TMEM + TMA).
//
// Replaces K15 / K18-forward (ffmlp/src/ffmlp.cu:331-407, 525-709): y = W_out . act(W_h ... act(W_in . x)), no bias,
// fp16 storage.  One persistent CTA (4 warps) walks 128-row batch tiles:
//   * all weight matrices are fetched ONCE per CTA by TMA (cp.async.bulk.tensor, hardware 64B/128B swizzle) and stay
//     resident in shared memory in the canonical K-major UMMA layout;
//   * the [128, in] input tile of the NEXT batch tile is prefetched by TMA into a second buffer while this one computes;
//   * each layer is K/16 `tcgen05.mma` (M=128 rows, N=64 hidden or 16 outputs, fp32 accumulators in TMEM) issued by a
//     single thread and tracked by `tcgen05.commit` -> mbarrier;
//   * the epilogue drains TMEM with `tcgen05.ld` (thread t of warp w owns row 32w+t), applies the activation, rounds
//     to fp16 exactly where the reference stores fp16 (ffmlp.cu:118), and writes the row back to shared memory in the
//     swizzled layout as the next layer's A operand -- activations never leave the SM between layers.  Training
//     additionally streams each layer's activations to forward_buffer[layer] (ffmlp.cu:395-404).
// Accumulation is fp32 (the reference's wmma path accumulates in fp16, ffmlp.cu:564): results agree within the
// fp16-accumulation error of the reference, well inside BASELINE.json's 1e-3.
#include "tc_common.cuh"

using namespace foc;
using namespace foc::tc;

namespace {

constexpr int kTile = 128;          // batch rows per tile = UMMA M
constexpr int kHidden = 64;         // hidden width (the FOC networks are 64 wide; ffmlp.py:103 allows 16..256)
constexpr int kOutPad = 16;         // padded output rows of W_out (ffmlp.cu:632)
constexpr int kMaxLayers = 8;

// ffmlp/src/utils.h:424-470.  Kept OUT of line: inlining this switch at each of the 64 elements of the unrolled epilogue
// made the kernel > 64 KB of SASS and a third of its stall samples were instruction-fetch misses (profiles/r1).
__device__ __noinline__ float act_fwd_generic(uint32_t a, float v) {
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

// ACT is a compile-time activation: 0 = ReLU, 6 = none, anything else = run-time value through act_fwd_generic
template <int ACT>
__device__ __forceinline__ float act_fwd(uint32_t a, float v) {
    if constexpr (ACT == 0) return fmaxf(v, 0.0f);
    else if constexpr (ACT == 6) return v;
    else return act_fwd_generic(a, v);
}

struct FwdParams {
    const __half* inputs;
    __half* forward_buffer;      // [num_layers, B, 64] or nullptr (inference)
    __half* outputs;             // [B, 16]
    uint32_t B, num_layers, activation, output_activation, num_tiles;
};

template <int IN>
struct FwdSmem {
    static constexpr uint32_t kInTileBytes = kTile * IN * 2;          // 8 KB (IN=32) / 16 KB (IN=64)
    static constexpr uint32_t kActBytes = kTile * kHidden * 2;        // 16 KB
    static constexpr uint32_t kWinBytes = kHidden * IN * 2;
    static constexpr uint32_t kWhBytes = kHidden * kHidden * 2;       // 8 KB per hidden matrix
    static constexpr uint32_t kWoutBytes = kOutPad * kHidden * 2;     // 2 KB
    __host__ __device__ static constexpr uint32_t off_in(int b) { return b * kInTileBytes; }
    __host__ __device__ static constexpr uint32_t off_act() { return 2 * kInTileBytes; }
    __host__ __device__ static constexpr uint32_t off_win() { return off_act() + kActBytes; }
    __host__ __device__ static constexpr uint32_t off_wh(int l) { return off_win() + ((kWinBytes + 1023u) & ~1023u) + l * kWhBytes; }
    __host__ __device__ static uint32_t off_wout(int n_hidden_mm) { return off_wh(n_hidden_mm); }
    __host__ __device__ static uint32_t off_bars(int n_hidden_mm) { return off_wout(n_hidden_mm) + ((kWoutBytes + 1023u) & ~1023u); }
    __host__ __device__ static uint32_t total(int n_hidden_mm) { return off_bars(n_hidden_mm) + 64 + 1024 /* base alignment slack */; }
};

template <int IN, int ACT>
__global__ void __launch_bounds__(128)
k_mlp_fwd(const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ CUtensorMap tm_win,
          const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_fb, FwdParams p) {
    using SM = FwdSmem<IN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int n_hmm = (int)p.num_layers - 1;                       // hidden x hidden matmuls
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SM::off_bars(n_hmm));
    uint64_t* bar_w = bars;            // weights landed
    uint64_t* bar_in = bars + 1;       // [2] input tile landed
    uint64_t* bar_mma = bars + 3;      // MMAs of the current layer complete
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tm_in); tma_prefetch_desc(&tm_win); tma_prefetch_desc(&tm_w);
        if (p.forward_buffer) tma_prefetch_desc(&tm_fb);
        mbar_init(bar_w, 1); mbar_init(bar_in, 1); mbar_init(bar_in + 1, 1); mbar_init(bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_slot, 64);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    pdl_sync();                                                      // everything above overlaps the previous kernel's tail

    const uint32_t first = blockIdx.x, stride = gridDim.x;
    if (first >= p.num_tiles) { __syncthreads(); if (warp == 0) tmem_dealloc(tmem, 64); return; }

    if (threadIdx.x == 0) {
        // weights: W_in as one box; hidden + output matrices as 16-row boxes of the [(L-1)*64+16, 64] tensor
        const uint32_t wbytes = SM::kWinBytes + (uint32_t)n_hmm * SM::kWhBytes + SM::kWoutBytes;
        mbar_expect_tx(bar_w, wbytes);
        tma_load_2d(smem + SM::off_win(), &tm_win, bar_w, 0, 0);
        for (int r16 = 0; r16 < n_hmm * 4 + 1; ++r16)
            tma_load_2d(smem + SM::off_wh(0) + r16 * 2048, &tm_w, bar_w, 0, r16 * 16);
        mbar_expect_tx(bar_in, SM::kInTileBytes);
        tma_load_2d(smem + SM::off_in(0), &tm_in, bar_in, 0, (int32_t)(first * kTile));
    }

    constexpr uint32_t kInSwz = IN == 32 ? SWZ_64B : SWZ_128B;
    constexpr uint32_t kInSbo = IN == 32 ? 512u : 1024u;
    constexpr uint32_t idesc_h = umma_idesc_f16(kTile, kHidden, 0, 0);
    constexpr uint32_t idesc_o = umma_idesc_f16(kTile, kOutPad, 0, 0);
    const uint32_t s_act = smem_u32(smem + SM::off_act());
    const uint32_t row = (uint32_t)(warp * 32 + lane);               // row of the tile this thread drains
    const uint32_t t_lane = tmem + ((uint32_t)(warp * 32) << 16);
    uint32_t mma_phase = 0;

    uint32_t it = 0;
    for (uint32_t tile = first; tile < p.num_tiles; tile += stride, ++it) {
        const uint32_t buf = it & 1u;
        const size_t grow = (size_t)tile * kTile + row;              // global batch row of this thread
        // ---- layer 0: x[128, IN] . W_in^T -> D[128, 64]
        if (threadIdx.x == 0) {
            const uint32_t next = tile + stride;
            if (next < p.num_tiles) {                                // prefetch the next tile's inputs (its buffer is free:
                mbar_expect_tx(bar_in + (buf ^ 1u), SM::kInTileBytes);   // the MMAs that read it completed last iteration)
                tma_load_2d(smem + SM::off_in(buf ^ 1u), &tm_in, bar_in + (buf ^ 1u), 0, (int32_t)(next * kTile));
            }
            if (it == 0) mbar_wait(bar_w, 0);
            mbar_wait(bar_in + buf, (it >> 1) & 1u);
            tc_fence_after();
            const uint32_t sa = smem_u32(smem + SM::off_in(buf)), sb = smem_u32(smem + SM::off_win());
            #pragma unroll
            for (int k = 0; k < IN / 16; ++k)
                umma_f16(tmem, umma_desc(sa + k * 32, 16, kInSbo, kInSwz), umma_desc(sb + k * 32, 16, kInSbo, kInSwz), idesc_h, k > 0);
            umma_commit(bar_mma);
        }
        for (uint32_t l = 0; l < p.num_layers; ++l) {
            // ---- epilogue of hidden layer l: TMEM -> act -> fp16 -> smem (next A operand) [+ forward_buffer]
            mbar_wait(bar_mma, mma_phase); mma_phase ^= 1u;
            tc_fence_after();
            uint32_t acc[4][16];
            #pragma unroll
            for (int q = 0; q < 4; ++q) tmem_ld16(t_lane + q * 16, acc[q]);
            tmem_ld_wait();
            uint4 packed[8];
            #pragma unroll
            for (int c = 0; c < 8; ++c) {
                uint32_t w[4];
                #pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int col = c * 8 + j * 2;
                    const float a = act_fwd<ACT>(p.activation, __uint_as_float(acc[col >> 4][col & 15]));
                    const float b = act_fwd<ACT>(p.activation, __uint_as_float(acc[(col + 1) >> 4][(col + 1) & 15]));
                    const __half2 h = __floats2half2_rn(a, b);
                    w[j] = *reinterpret_cast<const uint32_t*>(&h);
                }
                packed[c] = make_uint4(w[0], w[1], w[2], w[3]);
            }
            #pragma unroll
            for (int c = 0; c < 8; ++c)
                *reinterpret_cast<uint4*>(smem + SM::off_act() + sw128_off(row, c)) = packed[c];
            fence_proxy_async_smem();
            tc_fence_before();
            __syncthreads();
            // ---- next matmul: hidden l+1, or the output layer
            if (threadIdx.x == 0) {
                // training: the activation tile goes to forward_buffer as ONE bulk tensor store of the shared-memory tile (the layout
                // the backward's TMA loads expect), not as 8 x 16-byte stores per thread (the LSU queue was the kernel's lg_throttle)
                if (p.forward_buffer) { tma_store_2d(&tm_fb, smem + SM::off_act(), 0, (int32_t)((size_t)l * p.B + (size_t)tile * kTile)); bulk_commit(); }
                tc_fence_after();
                const bool last = (l + 1 == p.num_layers);
                const uint32_t sb = smem_u32(smem + (last ? SM::off_wout(n_hmm) : SM::off_wh((int)l)));
                #pragma unroll
                for (int k = 0; k < kHidden / 16; ++k)
                    umma_f16(tmem, umma_desc(s_act + k * 32, 16, 1024, SWZ_128B), umma_desc(sb + k * 32, 16, 1024, SWZ_128B),
                             last ? idesc_o : idesc_h, k > 0);
                // the tile is overwritten by the next epilogue, which starts when bar_mma flips: flip it only after the store has read
                if (p.forward_buffer) bulk_wait_read_all();
                umma_commit(bar_mma);
            }
        }
        // ---- output epilogue: D[128, 16] -> outputs[row, 0:16]
        mbar_wait(bar_mma, mma_phase); mma_phase ^= 1u;
        tc_fence_after();
        uint32_t o[16];
        tmem_ld16(t_lane, o);
        tmem_ld_wait();
        uint32_t w[8];
        #pragma unroll
        for (int j = 0; j < 8; ++j) {
            float v0 = __uint_as_float(o[2 * j]), v1 = __uint_as_float(o[2 * j + 1]);
            if (p.output_activation != 6) { v0 = act_fwd_generic(p.output_activation, v0); v1 = act_fwd_generic(p.output_activation, v1); }
            const __half2 h = __floats2half2_rn(v0, v1);
            w[j] = *reinterpret_cast<const uint32_t*>(&h);
        }
        uint4* dst = reinterpret_cast<uint4*>(p.outputs + grow * kOutPad);
        dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
        dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
        tc_fence_before();
        __syncthreads();                                             // every row drained before the next tile's MMA overwrites D
    }
    if (threadIdx.x == 0 && p.forward_buffer) bulk_wait_all();
    if (warp == 0) tmem_dealloc(tmem, 64);
}

template <int IN, int ACT>
int launch_fwd(const uint16_t* inputs, const uint16_t* weights, uint32_t B, uint32_t num_layers, uint32_t activation,
               uint32_t output_activation, uint16_t* forward_buffer, uint16_t* outputs, cudaStream_t st) {
    using SM = FwdSmem<IN>;
    const int n_hmm = (int)num_layers - 1;
    CUtensorMap tm_in, tm_win, tm_w, tm_fb;
    int rc = make_tmap_f16_2d(&tm_in, inputs, B, IN, kTile, IN);
    if (rc) return rc;
    if (forward_buffer) {
        rc = make_tmap_f16_2d(&tm_fb, forward_buffer, (uint64_t)num_layers * B, kHidden, kTile, kHidden);
        if (rc) return rc;
    } else tm_fb = tm_in;                                              // unused by the kernel in inference
    rc = make_tmap_f16_2d(&tm_win, weights, kHidden, IN, kHidden, IN);
    if (rc) return rc;
    rc = make_tmap_f16_2d(&tm_w, weights + (size_t)kHidden * IN, (uint64_t)n_hmm * kHidden + kOutPad, kHidden, 16, kHidden);
    if (rc) return rc;
    const uint32_t smem = SM::total(n_hmm);
    static DynSmemOptIn optin;                                         // per (instantiation, device)
    if (int rc2 = optin.ensure(k_mlp_fwd<IN, ACT>, 96 * 1024)) return rc2;
    if (smem > 96 * 1024) return FOC_ERR_UNSUPPORTED;
    FwdParams p;
    p.inputs = reinterpret_cast<const __half*>(inputs);
    p.forward_buffer = reinterpret_cast<__half*>(forward_buffer);
    p.outputs = reinterpret_cast<__half*>(outputs);
    p.B = B; p.num_layers = num_layers; p.activation = activation; p.output_activation = output_activation;
    p.num_tiles = B / kTile;
    const uint32_t ctas_per_sm = 4;                                    // 4 x 64 TMEM columns, ~4 x 45 KB shared memory
    const uint32_t cap = (uint32_t)foc::num_sms() * ctas_per_sm, grid = p.num_tiles < cap ? p.num_tiles : cap;
    foc::launch(k_mlp_fwd<IN, ACT>, grid, 128, smem, st, tm_in, tm_win, tm_w, tm_fb, p);
    FOC_LAUNCH_CHECK(); return 0;
}

}  // namespace

FOC_API int foc_ffmlp_forward(const uint16_t* inputs, const uint16_t* weights, uint32_t B, uint32_t input_dim,
                              uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers, uint32_t activation,
                              uint32_t output_activation, uint16_t* forward_buffer, uint16_t* outputs, void* stream) {
    if (B == 0) return 0;
    // the shapes of the FOC field networks (network_ff.py:31-49) take the tcgen05 kernel; whatever else ffmlp.py:99-123 allows (hidden
    // 16 / 32 / 128 / 256, input_dim = 16 m) goes to the CUDA-core kernels of mlp_generic.cu
    if (!foc::mlp_tensor_shape(input_dim, hidden_dim, num_layers))
        return foc::mlp_generic_forward(inputs, weights, B, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, forward_buffer,
                                        outputs, reinterpret_cast<cudaStream_t>(stream));
    if (hidden_dim != (uint32_t)kHidden || output_dim > (uint32_t)kOutPad || num_layers < 1 || num_layers > (uint32_t)kMaxLayers ||
        B % kTile != 0) return FOC_ERR_UNSUPPORTED;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#define FOC_FWD(IN_) \
    return activation == 0 ? launch_fwd<IN_, 0>(inputs, weights, B, num_layers, activation, output_activation, forward_buffer, outputs, st) \
         : activation == 6 ? launch_fwd<IN_, 6>(inputs, weights, B, num_layers, activation, output_activation, forward_buffer, outputs, st) \
                           : launch_fwd<IN_, -1>(inputs, weights, B, num_layers, activation, output_activation, forward_buffer, outputs, st)
    if (input_dim == 32) { FOC_FWD(32); }
    if (input_dim == 64) { FOC_FWD(64); }
#undef FOC_FWD
    return FOC_ERR_UNSUPPORTED;
}
