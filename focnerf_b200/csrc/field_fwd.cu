// field_fwd.cu -- the whole field evaluation behind the hash grid in ONE tcgen05 kernel:
//     enc [B,32] f16 --sigma net--> h [16] --(sigma = exp(h0); cin = [SH4(dir) | h[1:16] | 0])--> colour net --> rgb = sigmoid(hc[0:3])
// Replaces, for the FOC field (nerf/network_ff.py:51-74), two K15 launches (ffmlp.cu:331-407) and the ~10 elementwise PyTorch kernels
// between and after them -- in this repo: 2 x k_mlp_fwd + k_field_mid_fwd + k_rgb_fwd.  h, cin and hc never travel through HBM
// (inference) and the four launches become one.  Same arithmetic and the same fp16 rounding points as the unfused kernels
// (mlp_fwd.cu, field_glue.cu): results are bit-identical to that path (tests/test_gpu_field.py).
//
// Structure: the stages of a tile are a dependent chain (MMA -> TMEM drain -> activation -> shared memory -> next MMA), each ~1 us
// of latency for ~0.1 us of work, so throughput comes from TILES IN FLIGHT per SM.  The unfused kernels get 4 per SM from 4 CTAs
// that each hold their own copy of the weights.  Here one CTA per SM holds ONE copy of both nets' weights (36 KB) and G = 6
// independent warpgroups, each walking its own 128-row tiles with its own input / activation tiles (24 KB), mbarriers, named
// barrier and 64 TMEM columns.  The input tile is single-buffered: it holds enc for layer 0, is overwritten with cin for the
// colour net's layer 0, and is refilled by TMA with the warpgroup's next tile as soon as those MMAs have completed.
#include "tc_common.cuh"
#include "../../include/foc_b200.h"

using namespace foc;
using namespace foc::tc;

namespace {

constexpr int kTile = 128, kHidden = 64, kOutPad = 16, kIn = 32;
constexpr int kGroups = 6;                       // warpgroups (tiles in flight) per CTA
constexpr int kMaxHmm = 3;                       // hidden x hidden matmuls per net (num_layers - 1)
constexpr uint32_t kInBytes = kTile * kIn * 2;   // 8 KB, 64-byte rows, SWIZZLE_64B
constexpr uint32_t kActBytes = kTile * kHidden * 2;
constexpr uint32_t kWinBytes = kHidden * kIn * 2, kWhBytes = kHidden * kHidden * 2, kWoutBytes = kOutPad * kHidden * 2;

struct FieldFwdParams {
    const float* dirs;           // [B,3]
    float* sigma;                // [B]
    float* rgb;                  // [B,3]
    uint32_t B, num_tiles, n_hmm_s, n_hmm_c, training;
    const int32_t* n_rows;       // optional: valid leading rows (a multiple of 128), read from device memory; tiles beyond are skipped
};

// shared-memory map (offsets from the 1024-aligned base)
struct FieldSmem {
    uint32_t win_s, wh_s, wout_s, win_c, wh_c, wout_c, tiles, bars, total;
    __host__ __device__ FieldSmem(uint32_t hs, uint32_t hc) {
        win_s = 0; wh_s = win_s + kWinBytes; wout_s = wh_s + hs * kWhBytes;
        win_c = wout_s + 2048u; wh_c = win_c + kWinBytes; wout_c = wh_c + hc * kWhBytes;
        tiles = wout_c + 2048u;                                  // every matrix starts on a 1024-byte boundary (4096 / 8192 / 2048-byte sizes)
        bars = tiles + kGroups * (kInBytes + kActBytes);
        total = bars + 256 + 1024;
    }
    __host__ __device__ uint32_t in(int g) const { return tiles + g * (kInBytes + kActBytes); }
    __host__ __device__ uint32_t act(int g) const { return in(g) + kInBytes; }
};

__device__ __forceinline__ void group_sync(int g) { asm volatile("bar.sync %0, 128;" ::"r"(g + 1) : "memory"); }

// field_glue.cu: sh4 / pack helpers (same operations, same order)
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

// TMEM accumulator [128, 64] f32 -> ReLU -> fp16 -> this thread's row of the swizzled activation tile (mlp_fwd.cu epilogue)
__device__ __forceinline__ void drain_hidden(uint32_t t_lane, uint8_t* act_tile, uint32_t row) {
    #pragma unroll
    for (int half_ = 0; half_ < 2; ++half_) {                    // two 32-column halves: 32 live accumulator registers instead of 64
        uint32_t acc[2][16];
        tmem_ld16(t_lane + half_ * 32, acc[0]);
        tmem_ld16(t_lane + half_ * 32 + 16, acc[1]);
        tmem_ld_wait();
        #pragma unroll
        for (int c = 0; c < 4; ++c) {
            uint32_t w[4];
            #pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int col = c * 8 + j * 2;
                w[j] = pack2(fmaxf(__uint_as_float(acc[col >> 4][col & 15]), 0.0f), fmaxf(__uint_as_float(acc[(col + 1) >> 4][(col + 1) & 15]), 0.0f));
            }
            *reinterpret_cast<uint4*>(act_tile + sw128_off(row, half_ * 4 + c)) = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
}

template <bool TRAIN>
__global__ void __launch_bounds__(128 * kGroups, 1)
k_field_fwd(const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ CUtensorMap tm_win_s, const __grid_constant__ CUtensorMap tm_w_s,
            const __grid_constant__ CUtensorMap tm_win_c, const __grid_constant__ CUtensorMap tm_w_c, const __grid_constant__ CUtensorMap tm_fb_s,
            const __grid_constant__ CUtensorMap tm_fb_c, const __grid_constant__ CUtensorMap tm_cin, FieldFwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const FieldSmem L(p.n_hmm_s, p.n_hmm_c);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bars);
    uint64_t* bar_w = bars;                       // weights landed
    uint64_t* bar_in = bars + 1;                  // [kGroups] input tile landed
    uint64_t* bar_mma = bars + 1 + kGroups;       // [kGroups] MMAs of the group's current stage complete
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 1 + 2 * kGroups);
    constexpr uint32_t kTmemCols = 512;           // kGroups x 64 accumulator columns, rounded up to a power of two; one CTA per SM

    const int g = threadIdx.x >> 7, wt = threadIdx.x & 127, warp = threadIdx.x >> 5;
    const bool issuer = wt == 0;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tm_in); tma_prefetch_desc(&tm_win_s); tma_prefetch_desc(&tm_w_s); tma_prefetch_desc(&tm_win_c); tma_prefetch_desc(&tm_w_c);
        if (TRAIN) { tma_prefetch_desc(&tm_fb_s); tma_prefetch_desc(&tm_fb_c); tma_prefetch_desc(&tm_cin); }
        mbar_init(bar_w, 1);
        for (int i = 0; i < kGroups; ++i) { mbar_init(bar_in + i, 1); mbar_init(bar_mma + i, 1); }
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_slot, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    pdl_sync();

    if (p.n_rows != nullptr) { const uint32_t t = (uint32_t)max(__ldg(p.n_rows), 0) / kTile; p.num_tiles = t < p.num_tiles ? t : p.num_tiles; }
    const uint32_t hs = p.n_hmm_s, hc = p.n_hmm_c;
    if (threadIdx.x == 0) {                       // both nets' weights, once per CTA
        mbar_expect_tx(bar_w, 2 * kWinBytes + (hs + hc) * kWhBytes + 2 * kWoutBytes);
        tma_load_2d(smem + L.win_s, &tm_win_s, bar_w, 0, 0);
        for (uint32_t r16 = 0; r16 < hs * 4 + 1; ++r16) tma_load_2d(smem + L.wh_s + r16 * 2048, &tm_w_s, bar_w, 0, (int32_t)(r16 * 16));
        tma_load_2d(smem + L.win_c, &tm_win_c, bar_w, 0, 0);
        for (uint32_t r16 = 0; r16 < hc * 4 + 1; ++r16) tma_load_2d(smem + L.wh_c + r16 * 2048, &tm_w_c, bar_w, 0, (int32_t)(r16 * 16));
    }
    const uint32_t stride = gridDim.x * kGroups;
    uint32_t tile = blockIdx.x * kGroups + (uint32_t)g;
    uint8_t* in_tile = smem + L.in(g);
    uint8_t* act_tile = smem + L.act(g);
    if (issuer && tile < p.num_tiles) {
        mbar_expect_tx(bar_in + g, kInBytes);
        tma_load_2d(in_tile, &tm_in, bar_in + g, 0, (int32_t)(tile * kTile));
    }
    constexpr uint32_t idesc_h = umma_idesc_f16(kTile, kHidden, 0, 0), idesc_o = umma_idesc_f16(kTile, kOutPad, 0, 0);
    const uint32_t s_in = smem_u32(in_tile), s_act = smem_u32(act_tile);
    const uint32_t tmem_g = tmem + (uint32_t)g * 64u;
    const uint32_t row = (uint32_t)wt;
    const uint32_t t_lane = tmem_g + ((uint32_t)((wt >> 5) * 32) << 16);
    uint32_t mma_phase = 0, in_phase = 0;

    // one matmul stage: K/16 MMAs into this group's accumulator, tracked by the group's mbarrier
    auto issue_in = [&](uint32_t s_w) {           // [128, 32] (64-byte rows) x W_in^T
        #pragma unroll
        for (int k = 0; k < kIn / 16; ++k)
            umma_f16(tmem_g, umma_desc(s_in + k * 32, 16, 512, SWZ_64B), umma_desc(s_w + k * 32, 16, 512, SWZ_64B), idesc_h, k > 0);
    };
    auto issue_hidden = [&](uint32_t s_w, bool last) {
        #pragma unroll
        for (int k = 0; k < kHidden / 16; ++k)
            umma_f16(tmem_g, umma_desc(s_act + k * 32, 16, 1024, SWZ_128B), umma_desc(s_w + k * 32, 16, 1024, SWZ_128B), last ? idesc_o : idesc_h, k > 0);
    };

    for (uint32_t it = 0; tile < p.num_tiles; tile += stride, ++it) {
        const size_t grow = (size_t)tile * kTile + row;
        const float dx = __ldg(p.dirs + 3 * grow), dy = __ldg(p.dirs + 3 * grow + 1), dz = __ldg(p.dirs + 3 * grow + 2);
        const bool has_next = tile + stride < p.num_tiles;
        // ---------------------------------------------------------------- sigma net
        if (issuer) {
            if (it == 0) mbar_wait(bar_w, 0);
            mbar_wait(bar_in + g, in_phase);
            tc_fence_after();
            issue_in(smem_u32(smem + L.win_s));
            umma_commit(bar_mma + g);
        }
        for (uint32_t l = 0; l <= hs; ++l) {       // hs + 1 hidden activations
            mbar_wait(bar_mma + g, mma_phase); mma_phase ^= 1u;
            tc_fence_after();
            drain_hidden(t_lane, act_tile, row);
            fence_proxy_async_smem();
            tc_fence_before();
            group_sync(g);
            if (issuer) {
                if (TRAIN) { tma_store_2d(&tm_fb_s, act_tile, 0, (int32_t)((size_t)l * p.B + (size_t)tile * kTile)); bulk_commit(); }
                tc_fence_after();
                const bool last = l == hs;
                issue_hidden(smem_u32(smem + (last ? L.wout_s : L.wh_s + l * kWhBytes)), last);
                if (TRAIN) bulk_wait_read_all();  // the tile is overwritten by the next epilogue, which starts when the barrier flips
                umma_commit(bar_mma + g);
            }
        }
        // ---------------------------------------------------------------- h -> sigma, cin (field_glue.cu: k_field_mid_fwd)
        mbar_wait(bar_mma + g, mma_phase); mma_phase ^= 1u;
        tc_fence_after();
        {
            uint32_t o[16];
            tmem_ld16(t_lane, o);
            tmem_ld_wait();
            __half hh[16];
            #pragma unroll
            for (int j = 0; j < 16; ++j) hh[j] = __float2half_rn(__uint_as_float(o[j]));       // h is an fp16 tensor (ffmlp.cu:118)
            p.sigma[grow] = expf(__half2float(hh[0]));                                           // activation.py:10
            float s[16];
            sh4(dx, dy, dz, s);
            *reinterpret_cast<uint4*>(in_tile + sw64_off(row, 0)) = make_uint4(pack2(s[0], s[1]), pack2(s[2], s[3]), pack2(s[4], s[5]), pack2(s[6], s[7]));
            *reinterpret_cast<uint4*>(in_tile + sw64_off(row, 1)) = make_uint4(pack2(s[8], s[9]), pack2(s[10], s[11]), pack2(s[12], s[13]), pack2(s[14], s[15]));
            *reinterpret_cast<uint4*>(in_tile + sw64_off(row, 2)) = make_uint4(pack2h(hh[1], hh[2]), pack2h(hh[3], hh[4]), pack2h(hh[5], hh[6]), pack2h(hh[7], hh[8]));
            *reinterpret_cast<uint4*>(in_tile + sw64_off(row, 3)) = make_uint4(pack2h(hh[9], hh[10]), pack2h(hh[11], hh[12]), pack2h(hh[13], hh[14]),
                                                                               pack2h(hh[15], __float2half_rn(0.f)));
        }
        fence_proxy_async_smem();
        tc_fence_before();
        group_sync(g);
        // ---------------------------------------------------------------- colour net
        if (issuer) {
            if (TRAIN) { tma_store_2d(&tm_cin, in_tile, 0, (int32_t)(tile * kTile)); bulk_commit(); }      // the backward's X of the colour net
            tc_fence_after();
            issue_in(smem_u32(smem + L.win_c));
            if (TRAIN) bulk_wait_read_all();
            umma_commit(bar_mma + g);
        }
        for (uint32_t l = 0; l <= hc; ++l) {
            mbar_wait(bar_mma + g, mma_phase); mma_phase ^= 1u;
            tc_fence_after();
            if (l == 0 && issuer && has_next) {   // the input tile's last reader (the colour net's first matmul) is done: refill it
                mbar_expect_tx(bar_in + g, kInBytes);
                tma_load_2d(in_tile, &tm_in, bar_in + g, 0, (int32_t)((tile + stride) * kTile));
            }
            drain_hidden(t_lane, act_tile, row);
            fence_proxy_async_smem();
            tc_fence_before();
            group_sync(g);
            if (issuer) {
                if (TRAIN) { tma_store_2d(&tm_fb_c, act_tile, 0, (int32_t)((size_t)l * p.B + (size_t)tile * kTile)); bulk_commit(); }
                tc_fence_after();
                const bool last = l == hc;
                issue_hidden(smem_u32(smem + (last ? L.wout_c : L.wh_c + l * kWhBytes)), last);
                if (TRAIN) bulk_wait_read_all();
                umma_commit(bar_mma + g);
            }
        }
        // ---------------------------------------------------------------- hc -> rgb (field_glue.cu: k_rgb_fwd)
        mbar_wait(bar_mma + g, mma_phase); mma_phase ^= 1u;
        tc_fence_after();
        {
            uint32_t o[16];
            tmem_ld16(t_lane, o);
            tmem_ld_wait();
            #pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float hcv = __half2float(__float2half_rn(__uint_as_float(o[c])));
                p.rgb[3 * grow + c] = __half2float(__float2half_rn(1.0f / (1.0f + expf(-hcv))));
            }
        }
        tc_fence_before();
        group_sync(g);                             // every row drained before the next tile's first MMA overwrites the accumulator
        in_phase ^= 1u;
    }
    if (TRAIN && issuer) bulk_wait_all();
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, kTmemCols);
}

template <bool TRAIN>
int launch_field_fwd(const uint16_t* enc, const float* dirs, const uint16_t* w_s, const uint16_t* w_c, uint32_t B, uint32_t L_s, uint32_t L_c,
                     uint16_t* fb_s, uint16_t* fb_c, uint16_t* cin, float* sigma, float* rgb, const int32_t* n_rows, cudaStream_t st) {
    const uint32_t hs = L_s - 1, hc = L_c - 1;
    CUtensorMap tm_in, tm_win_s, tm_w_s, tm_win_c, tm_w_c, tm_fb_s, tm_fb_c, tm_cin;
    int rc;
    if ((rc = make_tmap_f16_2d(&tm_in, enc, B, kIn, kTile, kIn))) return rc;
    if ((rc = make_tmap_f16_2d(&tm_win_s, w_s, kHidden, kIn, kHidden, kIn))) return rc;
    if ((rc = make_tmap_f16_2d(&tm_w_s, w_s + (size_t)kHidden * kIn, (uint64_t)hs * kHidden + kOutPad, kHidden, 16, kHidden))) return rc;
    if ((rc = make_tmap_f16_2d(&tm_win_c, w_c, kHidden, kIn, kHidden, kIn))) return rc;
    if ((rc = make_tmap_f16_2d(&tm_w_c, w_c + (size_t)kHidden * kIn, (uint64_t)hc * kHidden + kOutPad, kHidden, 16, kHidden))) return rc;
    if (TRAIN) {
        if ((rc = make_tmap_f16_2d(&tm_fb_s, fb_s, (uint64_t)L_s * B, kHidden, kTile, kHidden))) return rc;
        if ((rc = make_tmap_f16_2d(&tm_fb_c, fb_c, (uint64_t)L_c * B, kHidden, kTile, kHidden))) return rc;
        if ((rc = make_tmap_f16_2d(&tm_cin, cin, B, kIn, kTile, kIn))) return rc;
    } else { tm_fb_s = tm_in; tm_fb_c = tm_in; tm_cin = tm_in; }       // unused in inference
    const FieldSmem L(hs, hc);
    if (L.total > 227 * 1024) return FOC_ERR_UNSUPPORTED;
    static DynSmemOptIn optin;
    if ((rc = optin.ensure(k_field_fwd<TRAIN>, 227 * 1024))) return rc;
    FieldFwdParams p;
    p.dirs = dirs; p.sigma = sigma; p.rgb = rgb; p.B = B; p.num_tiles = B / kTile; p.n_hmm_s = hs; p.n_hmm_c = hc; p.training = TRAIN; p.n_rows = n_rows;
    const uint32_t want = ceil_div(p.num_tiles, (uint32_t)kGroups);
    const uint32_t sms = (uint32_t)foc::num_sms(), grid = want < sms ? want : sms;
    foc::launch(k_field_fwd<TRAIN>, grid, 128 * kGroups, L.total, st, tm_in, tm_win_s, tm_w_s, tm_win_c, tm_w_c, tm_fb_s, tm_fb_c, tm_cin, p);
    FOC_LAUNCH_CHECK(); return 0;
}

}  // namespace

// enc [B,32] f16 (hash-grid encodings), dirs [B,3] f32, the two FFMLP weight blobs (fp16, ffmlp.cu:632 layout) -> sigma [B] f32, rgb [B,3] f32.
// Training (fb_s != NULL): also forward_buffer_s [L_s,B,64], forward_buffer_c [L_c,B,64] and cin [B,32] f16 for the backward kernels.
// n_rows (optional, device memory): only the first *n_rows rows (a multiple of 128) are evaluated.
FOC_API int foc_field_forward(const uint16_t* enc, const float* dirs, const uint16_t* weights_sigma, const uint16_t* weights_color, uint32_t B,
                              uint32_t num_layers_sigma, uint32_t num_layers_color, uint16_t* forward_buffer_sigma, uint16_t* forward_buffer_color,
                              uint16_t* cin, float* sigma, float* rgb, const int32_t* n_rows, void* stream) {
    if (B == 0) return 0;
    if (B % kTile != 0 || num_layers_sigma < 1 || num_layers_color < 1 || num_layers_sigma - 1 > (uint32_t)kMaxHmm ||
        num_layers_color - 1 > (uint32_t)kMaxHmm) return FOC_ERR_UNSUPPORTED;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const bool train = forward_buffer_sigma != nullptr;
    if (train && (forward_buffer_color == nullptr || cin == nullptr)) return FOC_ERR_UNSUPPORTED;
    return train ? launch_field_fwd<true>(enc, dirs, weights_sigma, weights_color, B, num_layers_sigma, num_layers_color, forward_buffer_sigma,
                                          forward_buffer_color, cin, sigma, rgb, n_rows, st)
                 : launch_field_fwd<false>(enc, dirs, weights_sigma, weights_color, B, num_layers_sigma, num_layers_color, nullptr, nullptr, nullptr,
                                           sigma, rgb, n_rows, st);
}
