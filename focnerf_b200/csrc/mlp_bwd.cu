// mlp_bwd.cu -- fused MLP backward on tcgen05/TMEM: activation-gradient chain AND all weight gradients in ONE kernel.
//
// Replaces K16 + K17 + K18-backward (ffmlp/src/ffmlp.cu:410-518, 580-628, 749-895): the reference runs one fused
// dgrad kernel that materialises backward_buffer[num_layers, B, 64], then num_layers+1 CUTLASS split-K GEMMs on side
// streams (fp16 accumulation) that re-read forward_buffer and backward_buffer.  Here each persistent CTA, per 128-row
// batch tile:
//   * TMA-loads the tile's stored activations A_0..A_{L-1} and inputs X (hardware swizzle), copies the [128,16] output
//     gradient G into shared memory;
//   * dgrad:  dA_{l-1} = dZ_l . W_l   as tcgen05.mma M=128 (K-major A = dZ tile, K-major B = W_l^T staged once per CTA);
//             the epilogue drains TMEM, applies the activation derivative from the stored A_{l-1}, rounds to fp16
//             (the reference's backward_buffer precision) and writes dZ_{l-1} back to shared memory;
//   * wgrad:  dW_l += dZ_l^T . A_{l-1}  as tcgen05.mma M=64, K=128 batch rows, reading THE SAME shared-memory tiles as
//             MN-major operands (a [rows=batch][64 features] swizzled tile is the canonical MN-major layout with
//             K = batch), accumulating in fp32 TMEM columns that persist across all tiles of the CTA.
// dZ never goes to HBM unless the caller asks for backward_buffer.  At the end every CTA stores its fp32 partial dW
// and a small second kernel reduces the partials in a fixed order (deterministic) and narrows to fp16.
#include "tc_common.cuh"

using namespace foc;
using namespace foc::tc;

namespace {

constexpr int kTile = 128, kHidden = 64, kOutPad = 16, kMaxLayers = 8;
constexpr uint32_t kTileBytes = kTile * kHidden * 2;     // 16 KB: one [128, 64] fp16 tile

// ffmlp/src/utils.h:537-582: derivative through the stored POST-activation value.  Out of line (code size; see mlp_fwd.cu).
__device__ __noinline__ float act_bwd_generic(uint32_t a, float g, float fwd) {
    switch (a) {
        case 0: return fwd > 0.f ? g : 0.f;
        case 1: return g * fwd;
        case 3: return g * (fwd * (1.0f - fwd));
        case 4: { const float y = fwd * 10.0f; return g * (y * y / (y * y + 1.0f)); }
        case 5: return g * (1.0f - __expf(-fwd * 10.0f));
        default: return g;
    }
}

template <int ACT>
__device__ __forceinline__ float act_bwd(uint32_t a, float g, float fwd) {
    if constexpr (ACT == 0) return fwd > 0.f ? g : 0.f;
    else if constexpr (ACT == 6) return g;
    else return act_bwd_generic(a, g, fwd);
}

struct BwdParams {
    // head: how this thread's row of the [B,16] output gradient is obtained (fetch_grad_row / finish_grad_row):
    //   0  read from `grad` [B,16] f16                                                    (the reference operator, ffmlp.h:9)
    //   1  colour net: formed from grad_rgb [B,3] f32 and rgb [B,3] f32                   (field_glue.cu: k_rgb_bwd)
    //   2  sigma net : formed from grad_sigma [B] f32 (or NULL), sigma [B] f32 and the colour net's grad_inputs [B,32] f16
    //                                                                                     (field_glue.cu: k_field_mid_bwd)
    int head;
    const float* head_grad;      // grad_rgb / grad_sigma
    const float* head_fwd;       // rgb / sigma
    const __half* head_gcin;     // [B,32] (head 2)
    const int32_t* n_rows;       // optional: valid leading rows (a multiple of 128), read from device memory; tiles beyond are skipped
    float fifteen;               // 15.0f at run time: expf(+-15) must be the DEVICE's expf, as in k_field_mid_bwd, not a folded constant
    const __half* grad;          // [B, 16]
    const __half* weights;       // flat
    __half* backward_buffer;     // [num_layers, B, 64] or nullptr
    __half* grad_inputs;         // [B, IN] or nullptr
    float* partials;             // [gridDim.x, num_params] fp32
    uint32_t B, num_layers, activation, num_tiles, num_params, tmem_cols;
};

__device__ __forceinline__ uint32_t bw_pack2(float a, float b) { const __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<const uint32_t*>(&h); }
__device__ __forceinline__ uint32_t bw_pack2h(__half a, __half b) { const __half2 h = __halves2half2(a, b); return *reinterpret_cast<const uint32_t*>(&h); }

// This thread's row of the output gradient G [B,16] f16, in two steps so that the global loads of the NEXT tile's row can be in flight
// during the current tile: fetch_grad_row only issues loads into registers, finish_grad_row does the arithmetic one tile later.
struct RawGrad { uint4 a, b; float f[6]; };

__device__ __forceinline__ void fetch_grad_row(const BwdParams& p, size_t grow, RawGrad& r) {
    if (p.head == 0) {
        const uint4* g = reinterpret_cast<const uint4*>(p.grad + grow * kOutPad);
        r.a = __ldg(g); r.b = __ldg(g + 1);
    } else if (p.head == 1) {
        #pragma unroll
        for (int c = 0; c < 3; ++c) { r.f[c] = __ldg(p.head_fwd + 3 * grow + c); r.f[3 + c] = __ldg(p.head_grad + 3 * grow + c); }
    } else {
        r.a = __ldg(reinterpret_cast<const uint4*>(p.head_gcin + grow * 32) + 2);
        r.b = __ldg(reinterpret_cast<const uint4*>(p.head_gcin + grow * 32) + 3);
        r.f[0] = p.head_grad ? __ldg(p.head_grad + grow) : 0.0f;
        r.f[1] = __ldg(p.head_fwd + grow);
    }
}
__device__ __forceinline__ void finish_grad_row(const BwdParams& p, const RawGrad& r, uint4& g0, uint4& g1) {
    if (p.head == 0) { g0 = r.a; g1 = r.b; }
    else if (p.head == 1) {                       // k_rgb_bwd: the gradient of a half tensor is half; d sigmoid = y (1 - y); columns 3..15 zero
        float g[3];
        #pragma unroll
        for (int c = 0; c < 3; ++c) g[c] = __half2float(__float2half_rn(r.f[3 + c])) * (1.0f - r.f[c]) * r.f[c];
        g0 = make_uint4(bw_pack2(g[0], g[1]), bw_pack2(g[2], 0.f), 0u, 0u);
        g1 = make_uint4(0u, 0u, 0u, 0u);
    } else {                                      // k_field_mid_bwd, with exp(clamp(h0, -15, 15)) == clamp(exp(h0), exp(-15), exp(15)): expf is monotonic
        const __half* a = reinterpret_cast<const __half*>(&r.a);                               // and sigma = expf(h0) is the forward's own output
        const __half* c = reinterpret_cast<const __half*>(&r.b);
        const float gs = p.head_grad ? r.f[0] * fminf(expf(p.fifteen), fmaxf(expf(-p.fifteen), r.f[1])) : 0.0f;
        g0 = make_uint4(bw_pack2h(__float2half_rn(gs), a[0]), bw_pack2h(a[1], a[2]), bw_pack2h(a[3], a[4]), bw_pack2h(a[5], a[6]));
        g1 = make_uint4(bw_pack2h(a[7], c[0]), bw_pack2h(c[1], c[2]), bw_pack2h(c[3], c[4]), bw_pack2h(c[5], c[6]));
    }
}

template <int IN>
struct BwdSmem {
    static constexpr uint32_t kXBytes = kTile * IN * 2;
    __host__ __device__ static uint32_t off_act(int l) { return l * kTileBytes; }
    __host__ __device__ static uint32_t off_x(int L) { return L * kTileBytes; }
    __host__ __device__ static uint32_t off_gz(int L) { return off_x(L) + ((kXBytes + 1023u) & ~1023u); }
    // weights exactly as the forward kernel holds them (TMA, hardware swizzle): hidden matrices, then W_out, then W_in
    __host__ __device__ static uint32_t off_w_h(int L, int i) { return off_gz(L) + kTileBytes + i * 8192u; }                  // W_{i+1} [64 o, 64 j]
    __host__ __device__ static uint32_t off_w_out(int L) { return off_w_h(L, L - 1); }                                       // [16 o, 64 j], 2 KB
    __host__ __device__ static uint32_t off_w_in(int L) { return off_w_out(L) + 2048u; }                                     // [64 o, IN i]
    __host__ __device__ static uint32_t off_bars(int L) { return off_w_in(L) + ((kHidden * IN * 2u + 1023u) & ~1023u); }
    __host__ __device__ static uint32_t total(int L) { return off_bars(L) + 128 + 1024; }
};

template <int IN, int ACT>
__global__ void __launch_bounds__(128)
k_mlp_bwd(const __grid_constant__ CUtensorMap tm_fb, const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ CUtensorMap tm_win,
          const __grid_constant__ CUtensorMap tm_w, BwdParams p) {
    using SM = BwdSmem<IN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int L = (int)p.num_layers, n_hmm = L - 1;
    if (p.n_rows != nullptr) { const uint32_t t = (uint32_t)max(__ldg(p.n_rows), 0) / kTile; p.num_tiles = t < p.num_tiles ? t : p.num_tiles; }
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SM::off_bars(L));
    uint64_t* bar_act = bars;                     // [L] one per stored-activation tile + [L] = the input tile X: each buffer is
    uint64_t* bar_x = bars + L;                   // refilled for the CTA's NEXT tile as soon as its last reader of this tile is done
    uint64_t* bar_mma = bars + kMaxLayers + 1;
    uint64_t* bar_w = bars + kMaxLayers + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + kMaxLayers + 3);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tm_fb); tma_prefetch_desc(&tm_in); tma_prefetch_desc(&tm_win); tma_prefetch_desc(&tm_w);
        for (int l = 0; l <= L; ++l) mbar_init(bar_act + l, 1);
        mbar_init(bar_mma, 1); mbar_init(bar_w, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_slot, p.tmem_cols);

    // ---- weights: fetched once per CTA by TMA in their stored [fan-out o][fan-in j] layout.  The dgrad products
    // dA_{l-1}[rows, j] = dZ_l[rows, o] . W_l[o, j] read them as MN-major B operands (N = j contiguous, K = o: 16 rows per
    // K step), the same way the weight-gradient MMAs read the activation tiles -- no transposed copies are staged.
    const size_t in_sz = (size_t)kHidden * IN, hh = (size_t)kHidden * kHidden;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    pdl_sync();                                                      // everything above overlaps the previous kernel's tail
    if (threadIdx.x == 0 && blockIdx.x < p.num_tiles) {
        mbar_expect_tx(bar_w, (uint32_t)(kHidden * IN * 2) + (uint32_t)n_hmm * 8192u + 2048u);
        tma_load_2d(smem + SM::off_w_in(L), &tm_win, bar_w, 0, 0);
        for (int r16 = 0; r16 < n_hmm * 4 + 1; ++r16)
            tma_load_2d(smem + SM::off_w_h(L, 0) + r16 * 2048, &tm_w, bar_w, 0, r16 * 16);
        for (int l = 0; l < L; ++l) {                                // the CTA's first tile
            mbar_expect_tx(bar_act + l, kTileBytes);
            tma_load_2d(smem + SM::off_act(l), &tm_fb, bar_act + l, 0, (int32_t)((size_t)l * p.B + (size_t)blockIdx.x * kTile));
        }
        mbar_expect_tx(bar_x, SM::kXBytes);
        tma_load_2d(smem + SM::off_x(L), &tm_in, bar_x, 0, (int32_t)(blockIdx.x * kTile));
    }

    // TMEM columns: [0,64) dgrad D | dW_0 (IN) | dW_1..dW_{L-1} (64 each) | dW_L^T (16)
    const uint32_t col_dw0 = 64, col_dwh0 = 64 + IN, col_dwout = 64 + IN + 64 * (uint32_t)n_hmm;
    constexpr uint32_t id_dgrad_h = umma_idesc_f16(128, 64, 0, 1), id_dgrad_in = umma_idesc_f16(128, IN, 0, 1);      // B MN-major
    constexpr uint32_t id_wg_h = umma_idesc_f16(64, 64, 1, 1), id_wg_in = umma_idesc_f16(64, IN, 1, 1), id_wg_out = umma_idesc_f16(64, 16, 1, 1);
    constexpr uint32_t kXSwz = IN == 32 ? SWZ_64B : SWZ_128B, kXSbo = IN == 32 ? 512u : 1024u, kXk16 = IN == 32 ? 1024u : 2048u;
    const uint32_t s_gz = smem_u32(smem + SM::off_gz(L));
    const uint32_t row = (uint32_t)(warp * 32 + lane);
    const uint32_t t_lane = tmem + ((uint32_t)(warp * 32) << 16);
    uint32_t ld_phase = 0, mma_phase = 0, acc_w = 0;      // acc_w: 0 on the CTA's first tile (overwrite), 1 afterwards
    RawGrad raw;                                          // this thread's row of the NEXT tile's output gradient: loads in flight, prefetched
    raw.a = raw.b = make_uint4(0, 0, 0, 0);
    #pragma unroll
    for (int c = 0; c < 6; ++c) raw.f[c] = 0.f;
    if (blockIdx.x < p.num_tiles) fetch_grad_row(p, (size_t)blockIdx.x * kTile + row, raw);

    for (uint32_t tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        const size_t grow = (size_t)tile * kTile + row;
        const uint32_t next_tile = tile + gridDim.x;
        const bool has_next = next_tile < p.num_tiles;
        {   // G tile: this thread's row, 16 halfs = 2 chunks, into the 128-byte-pitch swizzled layout
            uint4 gn0, gn1;
            finish_grad_row(p, raw, gn0, gn1);
            *reinterpret_cast<uint4*>(smem + SM::off_gz(L) + sw128_off(row, 0)) = gn0;
            *reinterpret_cast<uint4*>(smem + SM::off_gz(L) + sw128_off(row, 1)) = gn1;
            if (has_next) fetch_grad_row(p, (size_t)next_tile * kTile + row, raw);
        }
        fence_proxy_async_smem();
        tc_fence_before();
        __syncthreads();
        if (threadIdx.x == 0) {
            if (acc_w == 0) mbar_wait(bar_w, 0);
            mbar_wait(bar_act + (L - 1), ld_phase);
            tc_fence_after();
            // dA_{L-1} = G . W_L  (K = 16 outputs = the 16 rows of the W_out tile)
            umma_f16(tmem, umma_desc(s_gz, 16, 1024, SWZ_128B), umma_desc(smem_u32(smem + SM::off_w_out(L)), 16, 1024, SWZ_128B), id_dgrad_h, 0);
            // dW_L^T [64 j, 16 o] += A_{L-1}^T . G  (K = 128 batch rows)
            const uint32_t s_a = smem_u32(smem + SM::off_act(L - 1));
            #pragma unroll
            for (int k = 0; k < kTile / 16; ++k)
                umma_f16(tmem + col_dwout, umma_desc(s_a + k * 2048, 16, 1024, SWZ_128B), umma_desc(s_gz + k * 2048, 16, 1024, SWZ_128B),
                         id_wg_out, acc_w | (k > 0));
            umma_commit(bar_mma);
        }
        for (int l = L - 1; l >= 0; --l) {
            // ---- epilogue: dZ_l = dA_l (.) act'(A_l)  -> fp16 -> GZ tile
            mbar_wait(bar_mma, mma_phase); mma_phase ^= 1u;
            // every thread reads the TMA-written activation tile with generic loads below, so every thread must itself
            // observe its load barrier's phase (acquire), not just the MMA-issuing thread
            mbar_wait(bar_act + l, ld_phase);
            tc_fence_after();
            uint32_t acc[4][16];
            #pragma unroll
            for (int q = 0; q < 4; ++q) tmem_ld16(t_lane + q * 16, acc[q]);
            uint4 fw[8];
            #pragma unroll
            for (int c = 0; c < 8; ++c) fw[c] = *reinterpret_cast<const uint4*>(smem + SM::off_act(l) + sw128_off(row, c));
            tmem_ld_wait();
            uint4 packed[8];
            #pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint32_t f[4] = {fw[c].x, fw[c].y, fw[c].z, fw[c].w};
                uint32_t w[4];
                #pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int col = c * 8 + j * 2;
                    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&f[j]));
                    // the reference keeps dA in fp16 fragments before the activation transfer (ffmlp.cu:459-470)
                    const float g0 = __half2float(__float2half_rn(__uint_as_float(acc[col >> 4][col & 15])));
                    const float g1 = __half2float(__float2half_rn(__uint_as_float(acc[(col + 1) >> 4][(col + 1) & 15])));
                    const __half2 h = __floats2half2_rn(act_bwd<ACT>(p.activation, g0, a.x), act_bwd<ACT>(p.activation, g1, a.y));
                    w[j] = *reinterpret_cast<const uint32_t*>(&h);
                }
                packed[c] = make_uint4(w[0], w[1], w[2], w[3]);
            }
            #pragma unroll
            for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(smem + SM::off_gz(L) + sw128_off(row, c)) = packed[c];
            if (p.backward_buffer) {
                uint4* dst = reinterpret_cast<uint4*>(p.backward_buffer + ((size_t)(L - 1 - l) * p.B + grow) * kHidden);
                #pragma unroll
                for (int c = 0; c < 8; ++c) dst[c] = packed[c];
            }
            fence_proxy_async_smem();
            tc_fence_before();
            __syncthreads();
            if (threadIdx.x == 0) {
                // A_l has had its last reader of this tile (dW_{l+1} completed before this epilogue began, the epilogue's own
                // reads are behind the barrier above): refill the buffer with the next tile's A_l while this tile goes on
                if (has_next) {
                    mbar_expect_tx(bar_act + l, kTileBytes);
                    tma_load_2d(smem + SM::off_act(l), &tm_fb, bar_act + l, 0, (int32_t)((size_t)l * p.B + (size_t)next_tile * kTile));
                }
                if (l > 0) mbar_wait(bar_act + (l - 1), ld_phase); else mbar_wait(bar_x, ld_phase);
                tc_fence_after();
                if (l > 0) {
                    const uint32_t s_prev = smem_u32(smem + SM::off_act(l - 1));
                    #pragma unroll
                    for (int k = 0; k < kTile / 16; ++k)       // dW_l [64 o, 64 j] += dZ_l^T . A_{l-1}
                        umma_f16(tmem + col_dwh0 + 64u * (uint32_t)(l - 1), umma_desc(s_gz + k * 2048, 16, 1024, SWZ_128B),
                                 umma_desc(s_prev + k * 2048, 16, 1024, SWZ_128B), id_wg_h, acc_w | (k > 0));
                    const uint32_t s_w = smem_u32(smem + SM::off_w_h(L, l - 1));
                    #pragma unroll
                    for (int k = 0; k < kHidden / 16; ++k)     // dA_{l-1} = dZ_l . W_l   (K step = 16 rows o of the W_l tile)
                        umma_f16(tmem, umma_desc(s_gz + k * 32, 16, 1024, SWZ_128B), umma_desc(s_w + k * 2048, 16, 1024, SWZ_128B), id_dgrad_h, k > 0);
                } else {
                    const uint32_t s_x = smem_u32(smem + SM::off_x(L));
                    #pragma unroll
                    for (int k = 0; k < kTile / 16; ++k)       // dW_0 [64 o, IN] += dZ_0^T . X
                        umma_f16(tmem + col_dw0, umma_desc(s_gz + k * 2048, 16, 1024, SWZ_128B), umma_desc(s_x + k * kXk16, 16, kXSbo, kXSwz),
                                 id_wg_in, acc_w | (k > 0));
                    if (p.grad_inputs) {
                        const uint32_t s_w = smem_u32(smem + SM::off_w_in(L));
                        #pragma unroll
                        for (int k = 0; k < kHidden / 16; ++k) // dX = dZ_0 . W_0   (W_0 tile rows are IN*2 bytes: same form as the X tile)
                            umma_f16(tmem, umma_desc(s_gz + k * 32, 16, 1024, SWZ_128B), umma_desc(s_w + k * kXk16, 16, kXSbo, kXSwz), id_dgrad_in, k > 0);
                    }
                }
                umma_commit(bar_mma);
            }
        }
        // ---- all MMAs of this tile done: optional dX epilogue; shared-memory tiles become free for the next TMA
        mbar_wait(bar_mma, mma_phase); mma_phase ^= 1u;
        tc_fence_after();
        if (p.grad_inputs) {
            uint32_t gi[IN / 16][16];
            #pragma unroll
            for (int q = 0; q < IN / 16; ++q) tmem_ld16(t_lane + q * 16, gi[q]);
            tmem_ld_wait();
            uint4* dst = reinterpret_cast<uint4*>(p.grad_inputs + grow * IN);
            #pragma unroll
            for (int c = 0; c < IN / 8; ++c) {
                uint32_t w[4];
                #pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int col = c * 8 + j * 2;
                    const __half2 h = __floats2half2_rn(__uint_as_float(gi[col >> 4][col & 15]), __uint_as_float(gi[(col + 1) >> 4][(col + 1) & 15]));
                    w[j] = *reinterpret_cast<const uint32_t*>(&h);
                }
                dst[c] = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
        if (threadIdx.x == 0 && has_next) {            // X's last reader (dW_0 / dX) is done
            mbar_expect_tx(bar_x, SM::kXBytes);
            tma_load_2d(smem + SM::off_x(L), &tm_in, bar_x, 0, (int32_t)(next_tile * kTile));
        }
        ld_phase ^= 1u;
        tc_fence_before();
        __syncthreads();
        acc_w = 1;
    }

    // ---- flush this CTA's fp32 weight-gradient partials.  M = 64 accumulators live in lanes {0-15, 32-47, 64-79, 96-111}:
    // row (16 * warp + lane) for lane < 16.
    float* part = p.partials + (size_t)blockIdx.x * p.num_params;
    const bool has_row = lane < 16;
    const uint32_t orow = (uint32_t)(warp * 16 + lane);
    const bool worked = blockIdx.x < p.num_tiles;
    {   // dW_0 [64, IN]
        #pragma unroll
        for (int q = 0; q < IN / 16; ++q) {
            uint32_t v[16];
            tmem_ld16(t_lane + col_dw0 + q * 16, v);
            tmem_ld_wait();
            if (has_row) {
                float4* d = reinterpret_cast<float4*>(part + (size_t)orow * IN + q * 16);
                #pragma unroll
                for (int j = 0; j < 4; ++j)
                    d[j] = worked ? make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]))
                                  : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    }
    for (int l = 1; l <= n_hmm; ++l) {      // dW_l [64, 64]
        #pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint32_t v[16];
            tmem_ld16(t_lane + col_dwh0 + 64u * (uint32_t)(l - 1) + q * 16, v);
            tmem_ld_wait();
            if (has_row) {
                float4* d = reinterpret_cast<float4*>(part + in_sz + (size_t)(l - 1) * hh + (size_t)orow * kHidden + q * 16);
                #pragma unroll
                for (int j = 0; j < 4; ++j)
                    d[j] = worked ? make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]))
                                  : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    }
    {   // dW_L^T [64 j, 16 o] -> dW_L [16 o, 64 j]
        uint32_t v[16];
        tmem_ld16(t_lane + col_dwout, v);
        tmem_ld_wait();
        if (has_row) {
            float* d = part + in_sz + (size_t)n_hmm * hh;
            #pragma unroll
            for (int o = 0; o < 16; ++o) d[(size_t)o * kHidden + orow] = worked ? __uint_as_float(v[o]) : 0.f;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, p.tmem_cols);
}

// fixed-order reduction of the per-CTA partials, narrowed to fp16 (the reference's grad_weights dtype).
// Block = 32 parameters x 8 slices of the CTA range: each thread sums its slice (coalesced 128-byte rows), the 8 slice
// sums are combined in a fixed order through shared memory -> deterministic, and 8x shorter dependent-load chains.
__global__ void __launch_bounds__(256)
k_mlp_wgrad_finish(const float* __restrict__ partials, uint32_t n_cta, uint32_t num_params, __half* __restrict__ grad_weights, int accumulate) {
    pdl_sync();
    __shared__ float s[8][32];
    const uint32_t lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
    const uint32_t i = blockIdx.x * 32 + lane;
    float acc = 0.f;
    if (i < num_params) {
        const uint32_t per = (n_cta + 7) / 8, c0 = slice * per, c1 = min(n_cta, c0 + per);
        for (uint32_t c = c0; c < c1; ++c) acc += __ldg(partials + (size_t)c * num_params + i);
    }
    s[slice][lane] = acc;
    __syncthreads();
    if (slice == 0 && i < num_params) {
        float t = 0.f;
        #pragma unroll
        for (int k = 0; k < 8; ++k) t += s[k][lane];
        if (accumulate) t += __half2float(grad_weights[i]);       // `.grad` semantics for a persistent gradient buffer (adding to +0 is exact)
        grad_weights[i] = __float2half_rn(t);
    }
}

uint32_t bwd_grid(uint32_t B) {
    const uint32_t tiles = B / kTile, cap = (uint32_t)foc::num_sms() * 2u;
    return tiles < cap ? tiles : cap;
}
uint32_t num_params_of(uint32_t input_dim, uint32_t num_layers) { return kHidden * (input_dim + kHidden * (num_layers - 1) + kOutPad); }

struct BwdHead { int head; const float* grad; const float* fwd; const uint16_t* gcin; int accumulate; const int32_t* n_rows; };

template <int IN, int ACT>
int launch_bwd(const uint16_t* grad, const uint16_t* inputs, const uint16_t* weights, const uint16_t* forward_buffer, uint32_t B,
               uint32_t num_layers, uint32_t activation, int calc_grad_inputs, uint16_t* backward_buffer, uint16_t* grad_inputs,
               uint16_t* grad_weights, void* scratch, cudaStream_t st, BwdHead hd = BwdHead{0, nullptr, nullptr, nullptr, 0, nullptr}) {
    using SM = BwdSmem<IN>;
    CUtensorMap tm_fb, tm_in, tm_win, tm_w;
    int rc = make_tmap_f16_2d(&tm_fb, forward_buffer, (uint64_t)num_layers * B, kHidden, kTile, kHidden);
    if (rc) return rc;
    rc = make_tmap_f16_2d(&tm_in, inputs, B, IN, kTile, IN);
    if (rc) return rc;
    rc = make_tmap_f16_2d(&tm_win, weights, kHidden, IN, kHidden, IN);
    if (rc) return rc;
    rc = make_tmap_f16_2d(&tm_w, weights + (size_t)kHidden * IN, (uint64_t)(num_layers - 1) * kHidden + kOutPad, kHidden, 16, kHidden);
    if (rc) return rc;
    const uint32_t smem = SM::total((int)num_layers);
    if (smem > 227 * 1024) return FOC_ERR_UNSUPPORTED;
    static DynSmemOptIn optin;                                         // per (instantiation, device)
    if (int rc2 = optin.ensure(k_mlp_bwd<IN, ACT>, 227 * 1024)) return rc2;
    BwdParams p;
    p.head = hd.head; p.head_grad = hd.grad; p.head_fwd = hd.fwd; p.head_gcin = reinterpret_cast<const __half*>(hd.gcin); p.fifteen = 15.0f; p.n_rows = hd.n_rows;
    p.grad = reinterpret_cast<const __half*>(grad);
    p.weights = reinterpret_cast<const __half*>(weights);
    p.backward_buffer = reinterpret_cast<__half*>(backward_buffer);
    p.grad_inputs = calc_grad_inputs ? reinterpret_cast<__half*>(grad_inputs) : nullptr;
    p.partials = static_cast<float*>(scratch);
    p.B = B; p.num_layers = num_layers; p.activation = activation; p.num_tiles = B / kTile;
    p.num_params = num_params_of(IN, num_layers);
    const uint32_t need = 64 + IN + 64 * (num_layers - 1) + 16;
    p.tmem_cols = need <= 128 ? 128 : need <= 256 ? 256 : 512;
    if (need > 512) return FOC_ERR_UNSUPPORTED;
    const uint32_t grid = bwd_grid(B);
    foc::launch(k_mlp_bwd<IN, ACT>, grid, 128, smem, st, tm_fb, tm_in, tm_win, tm_w, p);
    FOC_LAUNCH_CHECK();
    if (grad_weights == nullptr) return 0;                         // the caller reduces the partials itself (foc_ffmlp_wgrad_finish)
    foc::launch(k_mlp_wgrad_finish, ceil_div(p.num_params, 32), 256, 0, st, p.partials, grid, p.num_params, reinterpret_cast<__half*>(grad_weights),
                hd.accumulate);
    FOC_LAUNCH_CHECK(); return 0;
}

}  // namespace

FOC_API size_t foc_ffmlp_backward_scratch_bytes(uint32_t B, uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers) {
    if (!foc::mlp_tensor_shape(input_dim, hidden_dim, num_layers))
        return foc::mlp_generic_shape_ok(B, input_dim, 16, hidden_dim, num_layers) ? foc::mlp_generic_scratch_bytes(B, input_dim, hidden_dim, num_layers) : 0;
    if (hidden_dim != (uint32_t)kHidden || B % kTile != 0 || num_layers < 1) return 0;
    return (size_t)bwd_grid(B) * num_params_of(input_dim, num_layers) * sizeof(float);
}

// The second half of foc_ffmlp_backward on its own: reduce the per-CTA fp32 partials a backward call with grad_weights == NULL
// left in `scratch` (same B, input_dim, num_layers).  Lets a caller run the reduction on another stream, beside the rest of its
// backward pass: nothing but the optimiser reads the weight gradient.
FOC_API int foc_ffmlp_wgrad_finish(const void* scratch, uint32_t B, uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers,
                                   uint16_t* grad_weights, void* stream) {
    if (B == 0) return 0;
    if (!foc::mlp_tensor_shape(input_dim, hidden_dim, num_layers) && foc::mlp_generic_shape_ok(B, input_dim, 16, hidden_dim, num_layers) && scratch && grad_weights)
        return foc::mlp_generic_finish(scratch, input_dim, hidden_dim, num_layers, grad_weights, 0, reinterpret_cast<cudaStream_t>(stream));
    if (hidden_dim != (uint32_t)kHidden || B % kTile != 0 || num_layers < 1 || !scratch || !grad_weights) return FOC_ERR_UNSUPPORTED;
    const uint32_t n = num_params_of(input_dim, num_layers);
    foc::launch(k_mlp_wgrad_finish, ceil_div(n, 32), 256, 0, reinterpret_cast<cudaStream_t>(stream), static_cast<const float*>(scratch), bwd_grid(B), n,
                reinterpret_cast<__half*>(grad_weights), 0);
    FOC_LAUNCH_CHECK(); return 0;
}

// foc_ffmlp_wgrad_finish that ADDS the reduced gradient to what `grad_weights` holds: the semantics of `.grad` for a persistent
// fp16 gradient buffer that an optimiser sweep zeroes after use (two backward passes before one step accumulate).
FOC_API int foc_ffmlp_wgrad_accumulate(const void* scratch, uint32_t B, uint32_t input_dim, uint32_t hidden_dim, uint32_t num_layers,
                                       uint16_t* grad_weights, void* stream) {
    if (B == 0) return 0;
    if (!foc::mlp_tensor_shape(input_dim, hidden_dim, num_layers) && foc::mlp_generic_shape_ok(B, input_dim, 16, hidden_dim, num_layers) && scratch && grad_weights)
        return foc::mlp_generic_finish(scratch, input_dim, hidden_dim, num_layers, grad_weights, 1, reinterpret_cast<cudaStream_t>(stream));
    if (hidden_dim != (uint32_t)kHidden || B % kTile != 0 || num_layers < 1 || !scratch || !grad_weights) return FOC_ERR_UNSUPPORTED;
    const uint32_t n = num_params_of(input_dim, num_layers);
    foc::launch(k_mlp_wgrad_finish, ceil_div(n, 32), 256, 0, reinterpret_cast<cudaStream_t>(stream), static_cast<const float*>(scratch), bwd_grid(B), n,
                reinterpret_cast<__half*>(grad_weights), 1);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_ffmlp_backward(const uint16_t* grad, const uint16_t* inputs, const uint16_t* weights, const uint16_t* forward_buffer,
                               uint32_t B, uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers,
                               uint32_t activation, uint32_t output_activation, int calc_grad_inputs, uint16_t* backward_buffer,
                               uint16_t* grad_inputs, uint16_t* grad_weights, void* scratch, void* stream) {
    if (B == 0) return 0;
    // output activation is always `none` in the reference module (ffmlp.py:108); activation 2 (sine) has no stored-value derivative
    if (!foc::mlp_tensor_shape(input_dim, hidden_dim, num_layers)) {
        if (output_dim != (uint32_t)kOutPad || output_activation != 6) return FOC_ERR_UNSUPPORTED;
        return foc::mlp_generic_backward(grad, inputs, weights, forward_buffer, B, input_dim, output_dim, hidden_dim, num_layers, activation, calc_grad_inputs,
                                         backward_buffer, grad_inputs, grad_weights, 0, scratch, reinterpret_cast<cudaStream_t>(stream));
    }
    if (hidden_dim != (uint32_t)kHidden || output_dim != (uint32_t)kOutPad || num_layers < 1 || num_layers > (uint32_t)kMaxLayers ||
        B % kTile != 0 || output_activation != 6 || activation == 2 || !scratch) return FOC_ERR_UNSUPPORTED;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#define FOC_BWD(IN_, ACT_) launch_bwd<IN_, ACT_>(grad, inputs, weights, forward_buffer, B, num_layers, activation, calc_grad_inputs, \
                                               backward_buffer, grad_inputs, grad_weights, scratch, st)
    if (input_dim == 32) return activation == 0 ? FOC_BWD(32, 0) : activation == 6 ? FOC_BWD(32, 6) : FOC_BWD(32, -1);
    if (input_dim == 64) return activation == 0 ? FOC_BWD(64, 0) : activation == 6 ? FOC_BWD(64, 6) : FOC_BWD(64, -1);
#undef FOC_BWD
    return FOC_ERR_UNSUPPORTED;
}

// foc_ffmlp_backward for the two nets of the FOC field, taking what the compositor / the colour net's backward produce instead of a
// materialised [B,16] fp16 output gradient (no reference counterpart: network_ff.py's sigmoid / trunc_exp / slice backward kernels):
//   head 1 (colour net): head_grad = grad_rgb [B,3] f32, head_fwd = rgb [B,3] f32                -> G = half(grad_rgb) * y (1 - y), cols 3..15 zero
//   head 2 (sigma net) : head_grad = grad_sigma [B] f32 or NULL, head_fwd = sigma [B] f32, grad_cin [B,32] f16 (the colour net's grad_inputs)
//                                                                                               -> G = [grad_sigma * clamp(sigma, e^-15, e^15), grad_cin[16:31]]
// Same values as foc_rgb_sigmoid_backward / foc_field_mid_backward followed by foc_ffmlp_backward.  accumulate != 0: the reduced weight
// gradient is ADDED to grad_weights (a persistent buffer zeroed by the optimiser sweep), else it overwrites it.  ReLU nets, hidden 64.
// n_rows (optional, device memory): only the first *n_rows rows (a multiple of 128) take part; grad_inputs rows beyond are not written.
FOC_API int foc_ffmlp_backward_head(int head, const float* head_grad, const float* head_fwd, const uint16_t* grad_cin, const uint16_t* inputs,
                                    const uint16_t* weights, const uint16_t* forward_buffer, uint32_t B, uint32_t input_dim, uint32_t hidden_dim,
                                    uint32_t num_layers, int calc_grad_inputs, uint16_t* grad_inputs, uint16_t* grad_weights, int accumulate,
                                    const int32_t* n_rows, void* scratch, void* stream) {
    if (B == 0) return 0;
    if (hidden_dim != (uint32_t)kHidden || input_dim != 32 || num_layers < 1 || num_layers > (uint32_t)kMaxLayers || B % kTile != 0 || !scratch ||
        (head != 1 && head != 2) || head_fwd == nullptr || (head == 1 && head_grad == nullptr) || (head == 2 && grad_cin == nullptr)) return FOC_ERR_UNSUPPORTED;
    return launch_bwd<32, 0>(nullptr, inputs, weights, forward_buffer, B, num_layers, 0u, calc_grad_inputs, nullptr, grad_inputs, grad_weights, scratch,
                             reinterpret_cast<cudaStream_t>(stream), BwdHead{head, head_grad, head_fwd, grad_cin, accumulate, n_rows});
}
