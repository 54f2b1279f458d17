// hashgrid_fwd.cu -- multiresolution hash-grid encoding, forward (K11: gridencoder/src/gridencoder.cu:87-245).
//
// One thread per (point, level).  All 2^D corner rows of a thread are fetched as ONE vector load each (half2 for the
// C=2 fp16 table of the FOC configuration: 4 B; float2 / 8-16 B vectors otherwise) and all 2^D loads are issued
// before the first use, so a warp keeps 256 independent gathers in flight.  On B200 the whole 23 MB fp16 table
// is L2-resident (126 MB L2), so the reference's level-major launch order is no longer needed for locality; two
// thread mappings are offered:
//   layout 0  [L,B,C]  the reference's output order (grid.y = level), what `_backend.grid_encode_forward` returns;
//   layout 1  [B,L,C]  point-major: a warp covers 32/L points x L levels and writes a contiguous 128-byte run of the
//                      [B, L*C] tensor the network consumes -- this removes the permute+reshape copy the reference's
//                      Python wrapper does after the kernel (grid.py:57).
// fp16 numerics follow the reference's c10::Half rounding sequence (see hashgrid.cuh) so encodings agree bit for bit.
#include <cstdlib>
#include "hashgrid.cuh"
#include "tc_common.cuh"

using namespace foc;

namespace {

template <typename T, int C> struct RowVec;     // one table row (C features) as a single aligned vector load
template <> struct RowVec<__half, 1> { static __device__ __forceinline__ void ld(const __half* p, float (&v)[1]) { v[0] = __half2float(__ldg(p)); } };
template <> struct RowVec<__half, 2> { static __device__ __forceinline__ void ld(const __half* p, float (&v)[2]) {
    const __half2 h = __ldg(reinterpret_cast<const __half2*>(p)); v[0] = __low2float(h); v[1] = __high2float(h); } };
template <> struct RowVec<__half, 4> { static __device__ __forceinline__ void ld(const __half* p, float (&v)[4]) {
    const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
    const __half2 a = *reinterpret_cast<const __half2*>(&u.x), b = *reinterpret_cast<const __half2*>(&u.y);
    v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b); } };
template <> struct RowVec<__half, 8> { static __device__ __forceinline__ void ld(const __half* p, float (&v)[8]) {
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    #pragma unroll
    for (int i = 0; i < 4; ++i) { const __half2 h = *reinterpret_cast<const __half2*>(&w[i]); v[2 * i] = __low2float(h); v[2 * i + 1] = __high2float(h); } } };
template <> struct RowVec<float, 1> { static __device__ __forceinline__ void ld(const float* p, float (&v)[1]) { v[0] = __ldg(p); } };
template <> struct RowVec<float, 2> { static __device__ __forceinline__ void ld(const float* p, float (&v)[2]) {
    const float2 f = __ldg(reinterpret_cast<const float2*>(p)); v[0] = f.x; v[1] = f.y; } };
template <> struct RowVec<float, 4> { static __device__ __forceinline__ void ld(const float* p, float (&v)[4]) {
    const float4 f = __ldg(reinterpret_cast<const float4*>(p)); v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w; } };
template <> struct RowVec<float, 8> { static __device__ __forceinline__ void ld(const float* p, float (&v)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w; } };

template <typename T, int C>
__device__ __forceinline__ void store_row(T* p, const float (&v)[C]) {
    if constexpr (sizeof(T) == 2 && C == 2) {
        *reinterpret_cast<__half2*>(p) = __floats2half2_rn(v[0], v[1]);
    } else if constexpr (sizeof(T) == 4 && C == 2) {
        *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
    } else {
        #pragma unroll
        for (int c = 0; c < C; ++c) Elem<T>::st(p + c, v[c]);
    }
}

// one (point, level): locate the cell, gather its 2^D corner rows, blend them in the reference's order -> res[C]
template <typename T, int D, int C>
__device__ __forceinline__ void encode_level(const float (&x)[D], const T* __restrict__ tab, const GridLevel& lv, const GridIndexer<D>& gidx,
                                             bool align_corners, uint32_t interp, float (&pos)[D], float (&pder)[D], uint32_t (&pg)[D],
                                             float (&res)[C]) {
    grid_locate<D>(x, lv.scale, align_corners, interp, pos, pder, pg);

    // gather: all corner rows first (independent loads), then the ordered accumulation of the reference
    float w[1 << D], rows[1 << D][C];
    #pragma unroll
    for (int idx = 0; idx < (1 << D); idx += 2) {
        // corners (idx, idx | 1) differ in x only: neighbouring rows on dense levels, and on hashed levels whenever x is even.
        // An aligned neighbour pair of the C = 2 fp16 table is fetched with one 8-byte load.
        uint32_t row[2];
        #pragma unroll
        for (int s = 0; s < 2; ++s) {
            float wt = 1.0f;
            uint32_t pl[D];
            #pragma unroll
            for (int d = 0; d < D; ++d) {
                if (((idx + s) >> d) & 1) { wt = __fmul_rn(wt, pos[d]); pl[d] = pg[d] + 1; }
                else { wt = __fmul_rn(wt, __fsub_rn(1.0f, pos[d])); pl[d] = pg[d]; }
            }
            w[idx + s] = wt; row[s] = gidx(pl);
        }
        if constexpr (sizeof(T) == 2 && C == 2) {
            if (row[1] == row[0] + 1 && (row[0] & 1u) == 0) {
                const uint2 u = __ldg(reinterpret_cast<const uint2*>(tab + (size_t)row[0] * 2));
                const __half2 a = *reinterpret_cast<const __half2*>(&u.x), b = *reinterpret_cast<const __half2*>(&u.y);
                rows[idx][0] = __low2float(a); rows[idx][1] = __high2float(a); rows[idx + 1][0] = __low2float(b); rows[idx + 1][1] = __high2float(b);
                continue;
            }
        }
        RowVec<T, C>::ld(tab + (size_t)row[0] * C, rows[idx]);
        RowVec<T, C>::ld(tab + (size_t)row[1] * C, rows[idx + 1]);
    }
    if constexpr (sizeof(T) == 2 && C % 2 == 0) {
        // c10::Half `results += w * g`: the product is rounded to half, the sum is formed in float and rounded to half
        // (hashgrid.cuh Elem<__half>::acc).  A native fp16 add rounds the exact sum once; rounding through fp32 first is
        // innocuous because fp32 carries 24 >= 2 * 11 + 2 significand bits, so HADD2 gives the same bits with a third of
        // the conversion instructions (the kernel was throttled on the conversion pipe).
        __half2 r2[C / 2];
        #pragma unroll
        for (int c = 0; c < C / 2; ++c) r2[c] = __floats2half2_rn(0.f, 0.f);
        #pragma unroll
        for (int idx = 0; idx < (1 << D); ++idx) {
            #pragma unroll
            for (int c = 0; c < C / 2; ++c)
                r2[c] = __hadd2(r2[c], __floats2half2_rn(__fmul_rn(w[idx], rows[idx][2 * c]), __fmul_rn(w[idx], rows[idx][2 * c + 1])));
        }
        #pragma unroll
        for (int c = 0; c < C / 2; ++c) { res[2 * c] = __low2float(r2[c]); res[2 * c + 1] = __high2float(r2[c]); }
    } else {
        #pragma unroll
        for (int c = 0; c < C; ++c) res[c] = 0.f;
        #pragma unroll
        for (int idx = 0; idx < (1 << D); ++idx) {
            #pragma unroll
            for (int c = 0; c < C; ++c) res[c] = Elem<T>::acc(res[c], w[idx], rows[idx][c]);
        }
    }
}

// Per-level constants of the tile kernel, resolved once per CTA (grid_level + GridIndexer cost ~45 instructions per use).
template <int D>
struct LevelParams {
    float scale;
    uint32_t offset, and_mask, hashed, need_mod, hashmap_size, resolution;
    uint32_t st[D];
};

// encode_level for a level whose lanes all take the same path (the tile kernel: the level is warp-uniform), index terms
// formed once per dimension: (p + 1) * k = p * k + k in the wrapping 32-bit arithmetic of gridencoder.cu:50-84, so the 2^D corner
// indices are D multiplies + D adds + (D - 1) combines each, instead of 2 D multiplies each.  Same indices, same float
// operations in the same order, same bits out.
// SMEM: `tab` is a shared-memory copy of the level (k_grid_fwd_staged): plain loads instead of ld.global.nc
template <typename T, int D, int C, bool HASHED, bool SMEM = false>
__device__ __forceinline__ void encode_level_uniform(const float (&x)[D], const T* __restrict__ tab, const LevelParams<D>& P,
                                                     bool align_corners, uint32_t interp, float (&res)[C]) {
    constexpr uint32_t primes[7] = {1u, 2654435761u, 805459861u, 3674653429u, 2097192037u, 1434869437u, 2165219737u};
    float pos[D], pder[D];
    uint32_t pg[D];
    grid_locate<D>(x, P.scale, align_corners, interp, pos, pder, pg);
    uint32_t term[D][2];
    float wd[D][2];
    #pragma unroll
    for (int d = 0; d < D; ++d) {
        const uint32_t k = HASHED ? primes[d] : P.st[d];
        term[d][0] = pg[d] * k; term[d][1] = term[d][0] + k;
        wd[d][0] = __fsub_rn(1.0f, pos[d]); wd[d][1] = pos[d];
    }
    float w[1 << D], rows[1 << D][C];
    #pragma unroll
    for (int idx = 0; idx < (1 << D); idx += 2) {
        uint32_t row[2];
        #pragma unroll
        for (int s = 0; s < 2; ++s) {
            float wt = wd[0][(idx + s) & 1];                             // 1 * w_0 is exact: the reference's running product starts at 1
            uint32_t r = term[0][(idx + s) & 1];
            #pragma unroll
            for (int d = 1; d < D; ++d) {
                const int bit = ((idx + s) >> d) & 1;
                wt = __fmul_rn(wt, wd[d][bit]);
                r = HASHED ? (r ^ term[d][bit]) : (r + term[d][bit]);
            }
            w[idx + s] = wt; row[s] = r & P.and_mask;
        }
        if constexpr (sizeof(T) == 2 && C == 2) {
            if (row[1] == row[0] + 1 && (row[0] & 1u) == 0) {
                const uint2 u = SMEM ? *reinterpret_cast<const uint2*>(tab + (size_t)row[0] * 2) : __ldg(reinterpret_cast<const uint2*>(tab + (size_t)row[0] * 2));
                const __half2 a = *reinterpret_cast<const __half2*>(&u.x), b = *reinterpret_cast<const __half2*>(&u.y);
                rows[idx][0] = __low2float(a); rows[idx][1] = __high2float(a); rows[idx + 1][0] = __low2float(b); rows[idx + 1][1] = __high2float(b);
                continue;
            }
            if constexpr (SMEM) {
                const __half2 h0 = *reinterpret_cast<const __half2*>(tab + (size_t)row[0] * 2), h1 = *reinterpret_cast<const __half2*>(tab + (size_t)row[1] * 2);
                rows[idx][0] = __low2float(h0); rows[idx][1] = __high2float(h0); rows[idx + 1][0] = __low2float(h1); rows[idx + 1][1] = __high2float(h1);
                continue;
            }
        }
        RowVec<T, C>::ld(tab + (size_t)row[0] * C, rows[idx]);
        RowVec<T, C>::ld(tab + (size_t)row[1] * C, rows[idx + 1]);
    }
    if constexpr (sizeof(T) == 2 && C % 2 == 0) {
        __half2 r2[C / 2];
        #pragma unroll
        for (int c = 0; c < C / 2; ++c) r2[c] = __floats2half2_rn(0.f, 0.f);
        #pragma unroll
        for (int idx = 0; idx < (1 << D); ++idx) {
            #pragma unroll
            for (int c = 0; c < C / 2; ++c)
                r2[c] = __hadd2(r2[c], __floats2half2_rn(__fmul_rn(w[idx], rows[idx][2 * c]), __fmul_rn(w[idx], rows[idx][2 * c + 1])));
        }
        #pragma unroll
        for (int c = 0; c < C / 2; ++c) { res[2 * c] = __low2float(r2[c]); res[2 * c + 1] = __high2float(r2[c]); }
    } else {
        #pragma unroll
        for (int c = 0; c < C; ++c) res[c] = 0.f;
        #pragma unroll
        for (int idx = 0; idx < (1 << D); ++idx) {
            #pragma unroll
            for (int c = 0; c < C; ++c) res[c] = Elem<T>::acc(res[c], w[idx], rows[idx][c]);
        }
    }
}

// Layout 1 without dy_dx, the hot configuration: a CTA takes 64 consecutive points x all L levels.  Lanes of a warp hold
// 32 CONSECUTIVE POINTS OF ONE LEVEL (neighbouring samples of a ray share cells on the coarse and middle levels, so a
// gather request touches few distinct sectors; with levels across the lanes every request touched >= 16), each thread
// walks levels (warp >> 1) + 4k for its point, and the [64, L*C] tile leaves through shared memory as full rows.
// ncu (800x800 frame, 23 M points): the point-major mapping was bound by l1tex__data_pipe_lsu_wavefronts (73 %).
constexpr int kTilePts = 64, kTileMaxLevels = 32;

template <typename T, int D, int C>
__global__ void __launch_bounds__(256)
k_grid_fwd_tile(const float* __restrict__ inputs, const T* __restrict__ table, const int32_t* __restrict__ offsets,
                T* __restrict__ outputs, uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners, uint32_t interp,
                const int32_t* __restrict__ n_rows) {
    pdl_sync();
    // n_rows (optional): the number of VALID leading rows, read from device memory -- a caller that replays this launch inside a CUDA
    // graph sizes the buffers once (B) and lets the march publish how many rows this step really has; CTAs beyond it leave at once
    if (n_rows != nullptr) { const uint32_t v = (uint32_t)max(__ldg(n_rows), 0); B = v < B ? v : B; }
    if (blockIdx.x * kTilePts >= B) return;
    extern __shared__ __align__(16) uint8_t tile_raw[];
    // [64][L*C (+ one 4-byte pad word when a row element is exactly one word: the fp16 C = 2 table -> conflict-free banks)]
    constexpr bool kPad = sizeof(T) * C == 4;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t pt = (warp & 1u) * 32u + lane;
    const uint32_t b0 = blockIdx.x * kTilePts, b = b0 + pt;
    const uint32_t row_elems = L * C;
    const uint32_t row_stride = row_elems + (kPad ? 4 / sizeof(T) : 0);  // in elements
    T* tile = reinterpret_cast<T*>(tile_raw);
    __shared__ LevelParams<D> s_lv[kTileMaxLevels];
    if (threadIdx.x < L) {
        const GridLevel lv = grid_level(offsets, threadIdx.x, S, H);
        const GridIndexer<D> gi(gridtype, align_corners, lv.hashmap_size, lv.resolution);
        LevelParams<D>& P = s_lv[threadIdx.x];
        P.scale = lv.scale; P.offset = lv.offset; P.and_mask = gi.and_mask; P.hashed = gi.hashed; P.need_mod = gi.need_mod;
        P.hashmap_size = lv.hashmap_size; P.resolution = lv.resolution;
        #pragma unroll
        for (int d = 0; d < D; ++d) P.st[d] = gi.st[d];
    }
    __syncthreads();
    if (b < B) {
        float x[D];
        #pragma unroll
        for (int d = 0; d < D; ++d) x[d] = __ldg(inputs + (size_t)b * D + d);
        const bool oob = grid_oob<D>(x);                               // gridencoder.cu:110-135
        for (uint32_t level = warp >> 1; level < L; level += 4) {
            float res[C];
            if (oob) {
                #pragma unroll
                for (int c = 0; c < C; ++c) res[c] = 0.f;
            } else {
                const LevelParams<D>& P = s_lv[level];
                const T* tab = table + (size_t)P.offset * C;
                if (P.need_mod) {                                        // tiled grids / odd table sizes: the general indexer
                    GridLevel lv; lv.scale = P.scale; lv.resolution = P.resolution; lv.hashmap_size = P.hashmap_size; lv.offset = P.offset;
                    const GridIndexer<D> gidx(gridtype, align_corners, lv.hashmap_size, lv.resolution);
                    float pos[D], pder[D];
                    uint32_t pg[D];
                    encode_level<T, D, C>(x, tab, lv, gidx, align_corners, interp, pos, pder, pg, res);
                } else if (P.hashed) encode_level_uniform<T, D, C, true>(x, tab, P, align_corners, interp, res);
                else encode_level_uniform<T, D, C, false>(x, tab, P, align_corners, interp, res);
            }
            store_row<T, C>(tile + (size_t)pt * row_stride + level * C, res);
        }
    }
    __syncthreads();
    // the tile is a contiguous run of the [B, L*C] output: consecutive threads store consecutive words
    const uint32_t npts = min((uint32_t)kTilePts, B - b0);
    if constexpr (kPad) {
        const uint32_t* tw = reinterpret_cast<const uint32_t*>(tile_raw);
        uint32_t* dst = reinterpret_cast<uint32_t*>(outputs + (size_t)b0 * row_elems);      // row_elems * sizeof(T) = 4 L bytes: word aligned
        for (uint32_t i = threadIdx.x; i < npts * L; i += 256) { const uint32_t r = i / L, c = i - r * L; dst[i] = tw[r * (L + 1) + c]; }
    } else {
        for (uint32_t i = threadIdx.x; i < npts * row_elems; i += 256) outputs[(size_t)b0 * row_elems + i] = tile[i];
    }
}

// EXPERIMENT (BASELINE.json north_star item 2; measured and NOT the default): the tile kernel as a persistent kernel whose CTAs
// first copy the coarsest `n_stage` levels of the fp16 table (level 0: 4 920 rows = 19.7 KB, level 1: 12 168 = 48.7 KB, level 2: 32 768 =
// 131 KB) into shared memory with one cp.async.bulk, and gather those levels from there.  THREADS / 4 points per tile iteration.
// Selected by FOC_GRID_FWD_STAGE=2 (two CTAs of 512 threads per SM) or =3 (one CTA of 1024 threads per SM).  Same arithmetic, same bits.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
k_grid_fwd_staged(const float* __restrict__ inputs, const __half* __restrict__ table, const int32_t* __restrict__ offsets,
                  __half* __restrict__ outputs, uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners, uint32_t interp,
                  const int32_t* __restrict__ n_rows, uint32_t n_stage, uint32_t stage_rows_cap) {
    using T = __half;
    constexpr int D = 3, C = 2, TP = THREADS / 4;
    pdl_sync();
    if (n_rows != nullptr) { const uint32_t v = (uint32_t)max(__ldg(n_rows), 0); B = v < B ? v : B; }
    extern __shared__ __align__(128) uint8_t tile_raw[];
    // rows of levels [0, n_stage) as the offsets tensor holds them; the host sized the buffer from its own restatement of the sizing rule
    uint32_t stage_rows = (uint32_t)__ldg(offsets + n_stage);
    if (stage_rows > stage_rows_cap) { n_stage = 0; stage_rows = 0; }                       // disagreement: gather everything from global memory
    __half* s_tab = reinterpret_cast<__half*>(tile_raw);                                    // [stage_rows][2]
    uint32_t* tw = reinterpret_cast<uint32_t*>(tile_raw + (size_t)stage_rows_cap * 4);      // [TP][L + 1] words
    __shared__ LevelParams<D> s_lv[kTileMaxLevels];
    __shared__ __align__(8) uint64_t s_bar;
    if (threadIdx.x < L) {
        const GridLevel lv = grid_level(offsets, threadIdx.x, S, H);
        const GridIndexer<D> gi(gridtype, align_corners, lv.hashmap_size, lv.resolution);
        LevelParams<D>& P = s_lv[threadIdx.x];
        P.scale = lv.scale; P.offset = lv.offset; P.and_mask = gi.and_mask; P.hashed = gi.hashed; P.need_mod = gi.need_mod;
        P.hashmap_size = lv.hashmap_size; P.resolution = lv.resolution;
        #pragma unroll
        for (int d = 0; d < D; ++d) P.st[d] = gi.st[d];
    }
    if (threadIdx.x == 0) { foc::tc::mbar_init(&s_bar, 1); foc::tc::fence_barrier_init(); }
    __syncthreads();
    if (threadIdx.x == 0 && stage_rows != 0) {
        const uint32_t bytes = stage_rows * 4u;
        foc::tc::mbar_expect_tx(&s_bar, bytes);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(foc::tc::smem_u32(s_tab)), "l"(table), "r"(bytes), "r"(foc::tc::smem_u32(&s_bar)) : "memory");
    }
    if (stage_rows != 0) foc::tc::mbar_wait(&s_bar, 0);
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    constexpr uint32_t kPtWarps = TP / 32;                              // warps side by side over the points of a tile
    const uint32_t pt = (warp % kPtWarps) * 32u + lane;
    for (uint32_t tile = blockIdx.x; (uint64_t)tile * TP < B; tile += gridDim.x) {
        const uint32_t b0 = tile * TP, b = b0 + pt;
        if (b < B) {
            float x[D];
            #pragma unroll
            for (int d = 0; d < D; ++d) x[d] = __ldg(inputs + (size_t)b * D + d);
            const bool oob = grid_oob<D>(x);
            for (uint32_t level = warp / kPtWarps; level < L; level += 4) {
                float res[C];
                if (oob) { res[0] = 0.f; res[1] = 0.f; }
                else {
                    const LevelParams<D>& P = s_lv[level];
                    if (level < n_stage && !P.need_mod && !P.hashed) encode_level_uniform<T, D, C, false, true>(x, s_tab + (size_t)P.offset * C, P, align_corners, interp, res);
                    else {
                        const T* tab = table + (size_t)P.offset * C;
                        if (P.need_mod) {
                            GridLevel lv; lv.scale = P.scale; lv.resolution = P.resolution; lv.hashmap_size = P.hashmap_size; lv.offset = P.offset;
                            const GridIndexer<D> gidx(gridtype, align_corners, lv.hashmap_size, lv.resolution);
                            float pos[D], pder[D];
                            uint32_t pg[D];
                            encode_level<T, D, C>(x, tab, lv, gidx, align_corners, interp, pos, pder, pg, res);
                        } else if (P.hashed) encode_level_uniform<T, D, C, true>(x, tab, P, align_corners, interp, res);
                        else encode_level_uniform<T, D, C, false>(x, tab, P, align_corners, interp, res);
                    }
                }
                store_row<T, C>(reinterpret_cast<T*>(tw + (size_t)pt * (L + 1) + level), res);
            }
        }
        __syncthreads();
        const uint32_t npts = min((uint32_t)TP, B - b0);
        uint32_t* dst = reinterpret_cast<uint32_t*>(outputs + (size_t)b0 * L * C);
        for (uint32_t i = threadIdx.x; i < npts * L; i += THREADS) { const uint32_t r = i / L, c = i - r * L; dst[i] = tw[r * (L + 1) + c]; }
        __syncthreads();
    }
}

template <typename T, int D, int C, int LAYOUT>
__global__ void __launch_bounds__(256)
k_grid_fwd(const float* __restrict__ inputs, const T* __restrict__ table, const int32_t* __restrict__ offsets,
           T* __restrict__ outputs, uint32_t B, uint32_t L, float S, uint32_t H, T* __restrict__ dy_dx,
           uint32_t gridtype, bool align_corners, uint32_t interp) {
    pdl_sync();
    uint32_t b, level;
    if constexpr (LAYOUT == 0) { b = blockIdx.x * blockDim.x + threadIdx.x; level = blockIdx.y; if (b >= B) return; }
    else {
        const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
        if (t >= (uint64_t)B * L) return;
        split_index(t, L, b, level);
    }
    T* out = LAYOUT == 0 ? outputs + ((size_t)level * B + b) * C : outputs + ((size_t)b * L + level) * C;
    T* dyo = dy_dx ? dy_dx + (((size_t)b * L + level) * D) * C : nullptr;       // [B, L, D, C]

    float x[D];
    #pragma unroll
    for (int d = 0; d < D; ++d) x[d] = __ldg(inputs + (size_t)b * D + d);
    if (grid_oob<D>(x)) {                                                        // gridencoder.cu:110-135
        float z[C];
        #pragma unroll
        for (int c = 0; c < C; ++c) z[c] = 0.f;
        store_row<T, C>(out, z);
        if (dyo) { for (int i = 0; i < D * C; ++i) Elem<T>::st(dyo + i, 0.f); }
        return;
    }
    const GridLevel lv = grid_level(offsets, level, S, H);
    const GridIndexer<D> gidx(gridtype, align_corners, lv.hashmap_size, lv.resolution);
    const T* tab = table + (size_t)lv.offset * C;
    float pos[D], pder[D], res[C];
    uint32_t pg[D];
    encode_level<T, D, C>(x, tab, lv, gidx, align_corners, interp, pos, pder, pg, res);
    store_row<T, C>(out, res);

    if (!dyo) return;
    // gridencoder.cu:201-243: d(out)/d(x_gd) = scale * sum over the other dims' corners of w' * (right - left) * deriv
    #pragma unroll
    for (int gd = 0; gd < D; ++gd) {
        float rg[C];
        #pragma unroll
        for (int c = 0; c < C; ++c) rg[c] = 0.f;
        #pragma unroll
        for (int idx = 0; idx < (1 << (D - 1)); ++idx) {
            float wt = lv.scale;
            uint32_t pl[D];
            #pragma unroll
            for (int nd = 0; nd < D - 1; ++nd) {
                const int d = (nd >= gd) ? nd + 1 : nd;
                if ((idx >> nd) & 1) { wt = __fmul_rn(wt, pos[d]); pl[d] = pg[d] + 1; }
                else { wt = __fmul_rn(wt, __fsub_rn(1.0f, pos[d])); pl[d] = pg[d]; }
            }
            float left[C], right[C];
            pl[gd] = pg[gd];
            RowVec<T, C>::ld(tab + (size_t)gidx(pl) * C, left);
            pl[gd] = pg[gd] + 1;
            RowVec<T, C>::ld(tab + (size_t)gidx(pl) * C, right);
            #pragma unroll
            for (int c = 0; c < C; ++c) {
                if constexpr (sizeof(T) == 2) {
                    const float diff = Elem<T>::rnd(__fsub_rn(right[c], left[c]));
                    rg[c] = Elem<T>::rnd(__fadd_rn(rg[c], Elem<T>::rnd(__fmul_rn(__fmul_rn(wt, diff), pder[gd]))));
                } else {
                    rg[c] += wt * (right[c] - left[c]) * pder[gd];
                }
            }
        }
        #pragma unroll
        for (int c = 0; c < C; ++c) Elem<T>::st(dyo + gd * C + c, rg[c]);
    }
}

// FOC_GRID_FWD_STAGE (read once): 0 = the tile kernel (default); 2 / 3 = k_grid_fwd_staged with that many coarse levels in shared memory
inline int grid_fwd_stage_levels() { static const int v = [] { const char* e = getenv("FOC_GRID_FWD_STAGE"); return e ? atoi(e) : 0; }(); return v; }

inline int launch_staged(const float* inputs, const __half* table, const int32_t* offsets, __half* outputs, uint32_t B, uint32_t L, float S, uint32_t H,
                         uint32_t gridtype, bool ac, uint32_t interp, const int32_t* n_rows, uint32_t ns, cudaStream_t st) {
    // rows of levels [0, ns): the gridencoder's sizing rule (grid.py:117-128), as the offsets tensor holds it (device memory: restated here)
    uint32_t rows = 0;
    for (uint32_t l = 0; l < ns; ++l) {
        const uint32_t res = (uint32_t)ceilf(exp2f((float)l * S) * (float)H);
        uint64_t n = (uint64_t)(res + 1) * (res + 1) * (res + 1);
        rows += (uint32_t)((n + 7) / 8 * 8);
    }
    const bool big = ns >= 3;
    const uint32_t threads = big ? 1024u : 512u, tp = threads / 4;
    const size_t smem = (size_t)rows * 4 + (size_t)tp * (L + 1) * 4;
    if (smem > 220 * 1024) return FOC_ERR_UNSUPPORTED;
    static DynSmemOptIn optin_a, optin_b;
    const int sms = foc::num_sms();
    if (big) {
        if (int rc = optin_b.ensure(k_grid_fwd_staged<1024>, 220 * 1024)) return rc;
        foc::launch(k_grid_fwd_staged<1024>, (uint32_t)sms, 1024, smem, st, inputs, table, offsets, outputs, B, L, S, H, gridtype, ac, interp, n_rows, ns, rows);
    } else {
        if (int rc = optin_a.ensure(k_grid_fwd_staged<512>, 110 * 1024)) return rc;
        foc::launch(k_grid_fwd_staged<512>, (uint32_t)sms * 2u, 512, smem, st, inputs, table, offsets, outputs, B, L, S, H, gridtype, ac, interp, n_rows, ns, rows);
    }
    FOC_LAUNCH_CHECK(); return 0;
}

template <typename T, int D, int C>
int launch_fwd(const float* inputs, const void* table, const int32_t* offsets, void* outputs, uint32_t B, uint32_t L,
               float S, uint32_t H, void* dy_dx, uint32_t gridtype, bool ac, uint32_t interp, int layout, cudaStream_t st, const int32_t* n_rows = nullptr) {
    const bool tile_ok = layout != 0 && !dy_dx && L <= (uint32_t)kTileMaxLevels && (size_t)kTilePts * (L * C * sizeof(T) + 4) <= 40 * 1024;
    if (n_rows != nullptr && !tile_ok) return FOC_ERR_UNSUPPORTED;     // the device-side row count is implemented by the tile kernel only
    if (layout == 0) {
        dim3 grid(ceil_div(B, 256), L);
        foc::launch(k_grid_fwd<T, D, C, 0>, grid, 256, 0, st, inputs, (const T*)table, offsets, (T*)outputs, B, L, S, H, (T*)dy_dx, gridtype, ac, interp);
    } else if (!dy_dx && L <= (uint32_t)kTileMaxLevels && (size_t)kTilePts * (L * C * sizeof(T) + 4) <= 40 * 1024) {
        if constexpr (sizeof(T) == 2 && D == 3 && C == 2) {
            if (const int ns = grid_fwd_stage_levels(); ns > 0 && gridtype == 0 && (uint32_t)ns < L)
                return launch_staged(inputs, (const __half*)table, offsets, (__half*)outputs, B, L, S, H, gridtype, ac, interp, n_rows, (uint32_t)ns, st);
        }
        foc::launch(k_grid_fwd_tile<T, D, C>, ceil_div(B, (uint32_t)kTilePts), 256, (size_t)kTilePts * (L * C * sizeof(T) + 4), st, inputs, (const T*)table, offsets,
                    (T*)outputs, B, L, S, H, gridtype, ac, interp, n_rows);
    } else {
        const uint64_t total = (uint64_t)B * L;
        foc::launch(k_grid_fwd<T, D, C, 1>, (uint32_t)((total + 255) / 256), 256, 0, st, inputs, (const T*)table, offsets, (T*)outputs, B, L, S, H, (T*)dy_dx, gridtype, ac, interp);
    }
    FOC_LAUNCH_CHECK(); return 0;
}

template <typename T, int D>
int dispatch_c(uint32_t C, const float* inputs, const void* table, const int32_t* offsets, void* outputs, uint32_t B,
               uint32_t L, float S, uint32_t H, void* dy_dx, uint32_t gridtype, bool ac, uint32_t interp, int layout, cudaStream_t st,
               const int32_t* n_rows = nullptr) {
    switch (C) {
        case 1: return launch_fwd<T, D, 1>(inputs, table, offsets, outputs, B, L, S, H, dy_dx, gridtype, ac, interp, layout, st, n_rows);
        case 2: return launch_fwd<T, D, 2>(inputs, table, offsets, outputs, B, L, S, H, dy_dx, gridtype, ac, interp, layout, st, n_rows);
        case 4: return launch_fwd<T, D, 4>(inputs, table, offsets, outputs, B, L, S, H, dy_dx, gridtype, ac, interp, layout, st, n_rows);
        case 8: return launch_fwd<T, D, 8>(inputs, table, offsets, outputs, B, L, S, H, dy_dx, gridtype, ac, interp, layout, st, n_rows);
        default: return FOC_ERR_UNSUPPORTED;     // the reference throws for other C (gridencoder.cu:381)
    }
}

}  // namespace

// The per-level scale exp2f(level * S) * H - 1 exactly as the encoder kernels evaluate it (device exp2f is an
// approximation for non-integer arguments).  Test hook: lets the CPU oracle reproduce fp16 encodings bit for bit.
__global__ void k_grid_level_scales(float S, uint32_t H, uint32_t L, float* __restrict__ scales) {
    const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l < L) scales[l] = fmaf(exp2f((float)l * S), (float)H, -1.0f);
}
FOC_API int foc_grid_level_scales(float S, uint32_t H, uint32_t L, float* scales, void* stream) {
    if (L == 0) return 0;
    k_grid_level_scales<<<ceil_div(L, 32), 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(S, H, L, scales);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_grid_encode_forward(const float* inputs, const void* embeddings, const int32_t* offsets, void* outputs,
                                    uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, void* dy_dx,
                                    uint32_t gridtype, int align_corners, uint32_t interp, int dtype, int layout, void* stream) {
    if (B == 0 || L == 0) return 0;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const bool ac = align_corners != 0;
    if (dtype == 1) {
        if (D == 2) return dispatch_c<__half, 2>(C, inputs, embeddings, offsets, outputs, B, L, S, H, dy_dx, gridtype, ac, interp, layout, st);
        if (D == 3) return dispatch_c<__half, 3>(C, inputs, embeddings, offsets, outputs, B, L, S, H, dy_dx, gridtype, ac, interp, layout, st);
    } else if (dtype == 0) {
        if (D == 2) return dispatch_c<float, 2>(C, inputs, embeddings, offsets, outputs, B, L, S, H, dy_dx, gridtype, ac, interp, layout, st);
        if (D == 3) return dispatch_c<float, 3>(C, inputs, embeddings, offsets, outputs, B, L, S, H, dy_dx, gridtype, ac, interp, layout, st);
    }
    return FOC_ERR_UNSUPPORTED;                  // the reference supports D in {2,3,4,5}; FOC uses 3 (gridencoder.cu:388-398)
}

// foc_grid_encode_forward over the first *n_rows rows only (n_rows in device memory, clamped to B): for launches replayed inside a CUDA
// graph whose buffers are sized once.  Point-major layout without dy_dx (the training / render path); rows >= *n_rows are not written.
FOC_API int foc_grid_encode_forward_rows(const float* inputs, const void* embeddings, const int32_t* offsets, void* outputs,
                                         uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, uint32_t gridtype,
                                         int align_corners, uint32_t interp, int dtype, const int32_t* n_rows, void* stream) {
    if (B == 0 || L == 0) return 0;
    if (n_rows == nullptr) return FOC_ERR_UNSUPPORTED;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const bool ac = align_corners != 0;
    if (dtype == 1 && D == 3) return dispatch_c<__half, 3>(C, inputs, embeddings, offsets, outputs, B, L, S, H, nullptr, gridtype, ac, interp, 1, st, n_rows);
    if (dtype == 0 && D == 3) return dispatch_c<float, 3>(C, inputs, embeddings, offsets, outputs, B, L, S, H, nullptr, gridtype, ac, interp, 1, st, n_rows);
    return FOC_ERR_UNSUPPORTED;
}
