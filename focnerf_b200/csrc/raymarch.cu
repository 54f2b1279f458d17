// raymarch.cu -- occupancy-bitfield ray marching for sm_100a.
//
// Replaces the reference kernels K1-K6, K9 (raymarching/src/raymarching.cu:92-490, 701-815) behind the C ABI of
// include/foc_b200.h.  Integer results (sample counts, cell indices, offsets) are bit-exact with the reference
// given the same bitfield: every float operation below states its rounding (foc::fmul/fadd/ffma), placed to match
// the FMA contractions nvcc makes for the reference source (verified against the SASS of oracle/_ref).
//
// Training march = three launches, no spin-waits, no atomics:
//   record : walk each ray once, recording the parameter t of every occupied step -> scratch counts[N], tbuf[N,max_steps]
//            (constant step: warp per ray, probes in parallel -- k_march_record_const; dt_gamma > 0: serial walk)
//   scan   : 1 CTA, warp-shuffle exclusive prefix scan over counts                 -> rays[N,3], counter[2]
//   expand : warp per ray turns the recorded t's into samples, coalesced (slots are ray-ordered and deterministic; the
//            reference hands them out in atomic arrival order, raymarching.cu:405-406)
#include "common.cuh"

namespace foc {

struct MarchParams {
    const uint8_t* grid;
    float bound, dt_gamma, dt_min, dt_max, rH, H3, Hf, Hm1, halfH;
    int C;
    uint32_t H;
};

__host__ inline MarchParams make_march_params(const uint8_t* grid, float bound, float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t H) {
    MarchParams p;
    p.grid = grid; p.bound = bound; p.dt_gamma = dt_gamma; p.C = (int)C; p.H = H;
    const float two_sqrt3 = 2.0f * 1.7320508075688772f;                 // raymarching.cu:345-346
    p.dt_min = two_sqrt3 / (float)max_steps;                            // IEEE division on the host == device __fdiv_rn
    p.dt_max = (two_sqrt3 * (float)(1 << (C - 1))) / (float)H;
    p.rH = 1.0f / (float)H; p.Hf = (float)H; p.Hm1 = (float)(H - 1); p.H3 = (float)(H * H * H); p.halfH = 0.5f * (float)H;
    return p;
}

struct Ray {
    float ox, oy, oz, dx, dy, dz, rdx, rdy, rdz, sx, sy, sz, far;
};

__device__ __forceinline__ Ray load_ray(const float* __restrict__ rays_o, const float* __restrict__ rays_d, uint32_t id, float far) {
    Ray r;
    r.ox = __ldg(rays_o + 3 * (size_t)id); r.oy = __ldg(rays_o + 3 * (size_t)id + 1); r.oz = __ldg(rays_o + 3 * (size_t)id + 2);
    r.dx = __ldg(rays_d + 3 * (size_t)id); r.dy = __ldg(rays_d + 3 * (size_t)id + 1); r.dz = __ldg(rays_d + 3 * (size_t)id + 2);
    r.rdx = __frcp_rn(r.dx); r.rdy = __frcp_rn(r.dy); r.rdz = __frcp_rn(r.dz);       // 1/d, correctly rounded (l.337)
    // 0.5 * sign(d) folded once: (n + 0.5) + 0.5*sign(d) is exact in fp32 for n < 2^22
    r.sx = copysignf(0.5f, r.dx); r.sy = copysignf(0.5f, r.dy); r.sz = copysignf(0.5f, r.dz);
    r.far = far;
    return r;
}

__device__ __forceinline__ float step_dt(const MarchParams& p, float t) { return clampf(fmul(t, p.dt_gamma), p.dt_min, p.dt_max); }

// frexp exponent of a non-negative finite float, clamped to [0, C-1] (raymarching.cu:42-54)
__device__ __forceinline__ int mip_level(float mx, int C) {
    int e = (int)((__float_as_uint(mx) >> 23) & 0xffu) - 126;   // normal numbers; zero/denormals give e <= 0
    return min(C - 1, max(0, e));
}

// The cell test at parameter t (raymarching.cu:360-392): fills x,y,z,dt; returns true when the cell is occupied, else
// sets tt = the parameter at which the ray leaves the (empty) cell.
__device__ __forceinline__ bool probe_cell(const MarchParams& p, const Ray& r, float tc, float& x, float& y, float& z, float& dt, float& tt) {
    x = clampf(ffma(tc, r.dx, r.ox), -p.bound, p.bound);
    y = clampf(ffma(tc, r.dy, r.oy), -p.bound, p.bound);
    z = clampf(ffma(tc, r.dz, r.oz), -p.bound, p.bound);
    dt = step_dt(p, tc);
    int level = 0;
    float mip_bound = fminf(1.0f, p.bound), mip_rbound;
    if (p.C > 1) {
        const float mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
        level = max(mip_level(mx, p.C), mip_level(fmul(fmul(dt, p.Hf), 0.5f), p.C));
        mip_bound = fminf(__uint_as_float((uint32_t)(127 + level) << 23), p.bound);   // scalbnf(1, level)
    }
    mip_rbound = __frcp_rn(mip_bound);
    // 0.5 * (x * rbound + 1) * H : the reference evaluates both products in double and narrows once (l.374).
    // (0.5*v)*H is exact in double (24-bit x <=24-bit significands), so the single narrowing is the correctly
    // rounded fp32 product v * (H/2) -- one FMUL here, no FP64 pipe.
    const float fx = fmul(ffma(x, mip_rbound, 1.0f), p.halfH);
    const float fy = fmul(ffma(y, mip_rbound, 1.0f), p.halfH);
    const float fz = fmul(ffma(z, mip_rbound, 1.0f), p.halfH);
    const int nx = (int)clampf(fx, 0.0f, p.Hm1), ny = (int)clampf(fy, 0.0f, p.Hm1), nz = (int)clampf(fz, 0.0f, p.Hm1);
    const uint32_t index = (uint32_t)ffma((float)level, p.H3, (float)morton3((uint32_t)nx, (uint32_t)ny, (uint32_t)nz));
    const bool occ = (__ldg(p.grid + (index >> 3)) >> (index & 7u)) & 1u;
    if (occ) return true;
    const float tx = fmul(ffma(mip_bound, ffma(fmul(fadd(fadd((float)nx, 0.5f), r.sx), p.rH), 2.0f, -1.0f), -x), r.rdx);
    const float ty = fmul(ffma(mip_bound, ffma(fmul(fadd(fadd((float)ny, 0.5f), r.sy), p.rH), 2.0f, -1.0f), -y), r.rdy);
    const float tz = fmul(ffma(mip_bound, ffma(fmul(fadd(fadd((float)nz, 0.5f), r.sz), p.rH), 2.0f, -1.0f), -z), r.rdz);
    tt = fadd(tc, fmaxf(0.0f, fminf(tx, fminf(ty, tz))));
    return false;
}

// One probe at parameter t: returns true when the cell is occupied (x,y,z,dt filled);
// otherwise advances t past the empty cell (raymarching.cu:393-399).
__device__ __forceinline__ bool probe(const MarchParams& p, const Ray& r, float& t, float& x, float& y, float& z, float& dt) {
    float tt;
    if (probe_cell(p, r, t, x, y, z, dt, tt)) return true;
    float tn = t;
    do { tn = fadd(tn, step_dt(p, tn)); } while (tn < tt);
    t = tn;
    return false;
}

}  // namespace foc
using namespace foc;

// ------------------------------------------------------------------------------------------------ utility kernels

// K1 (raymarching.cu:92-145): slab test; near clamped to min_near; miss -> FLT_MAX for both.
__global__ void __launch_bounds__(256) k_near_far(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                                  const float* __restrict__ aabb, uint32_t N, float min_near,
                                                  float* __restrict__ nears, float* __restrict__ fars) {
    pdl_sync();
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    float tn = 0.f, tf = 0.f;
    bool miss = false;
    #pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float o = __ldg(rays_o + 3 * (size_t)n + a), r = __frcp_rn(__ldg(rays_d + 3 * (size_t)n + a));
        float lo = fmul(fsub(__ldg(aabb + a), o), r), hi = fmul(fsub(__ldg(aabb + a + 3), o), r);
        if (lo > hi) { const float s = lo; lo = hi; hi = s; }
        if (a == 0) { tn = lo; tf = hi; }
        else if (!miss) {
            if (tn > hi || lo > tf) miss = true;
            else { if (lo > tn) tn = lo; if (hi < tf) tf = hi; }
        }
    }
    if (miss) { tn = tf = 3.402823466e+38f; }
    else if (tn < min_near) tn = min_near;
    nears[n] = tn; fars[n] = tf;
}

// K2 (raymarching.cu:163-198)
__global__ void __launch_bounds__(256) k_sph_from_ray(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                                      float radius, uint32_t N, float* __restrict__ coords) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float ox = rays_o[3 * (size_t)n], oy = rays_o[3 * (size_t)n + 1], oz = rays_o[3 * (size_t)n + 2];
    const float dx = rays_d[3 * (size_t)n], dy = rays_d[3 * (size_t)n + 1], dz = rays_d[3 * (size_t)n + 2];
    // roundings stated as nvcc contracts the reference's expressions (raymarching.cu:184-197, -fmad=true; read off the SASS of
    // oracle/_ref/raymarching): `a*b + c*d + e*f` -> fma(e, f, fma(a, b, c*d)), `s - r*r` -> fma(-r, r, s), `x*x + z*z` -> fma(x, x, z*z),
    // `2 * theta * RPI - 1` -> fma(theta + theta, RPI, -1); oracle/orc_raymarch.c spells out the same sequence
    const float A = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
    const float B = __fmaf_rn(oz, dz, __fmaf_rn(ox, dx, __fmul_rn(oy, dy)));
    const float Cc = __fmaf_rn(-radius, radius, __fmaf_rn(oz, oz, __fmaf_rn(ox, ox, __fmul_rn(oy, oy))));
    const float t = __fdiv_rn(__fadd_rn(-B, sqrtf(__fmaf_rn(B, B, -__fmul_rn(A, Cc)))), A);
    const float x = __fmaf_rn(t, dx, ox), y = __fmaf_rn(t, dy, oy), z = __fmaf_rn(t, dz, oz);
    const float theta = atan2f(sqrtf(__fmaf_rn(x, x, __fmul_rn(z, z))), y), phi = atan2f(z, x);
    const float RPI = 0.3183098861837907f;
    coords[2 * (size_t)n] = __fmaf_rn(__fmul_rn(2.0f, theta), RPI, -1.0f);
    coords[2 * (size_t)n + 1] = __fmul_rn(phi, RPI);
}

// K3 / K4 (raymarching.cu:214-254)
__global__ void __launch_bounds__(256) k_morton3d(const int32_t* __restrict__ coords, uint32_t N, int32_t* __restrict__ indices) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    indices[n] = (int32_t)morton3((uint32_t)coords[3 * (size_t)n], (uint32_t)coords[3 * (size_t)n + 1], (uint32_t)coords[3 * (size_t)n + 2]);
}
__global__ void __launch_bounds__(256) k_morton3d_invert(const int32_t* __restrict__ indices, uint32_t N, int32_t* __restrict__ coords) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const uint32_t v = (uint32_t)indices[n];
    coords[3 * (size_t)n] = (int32_t)compact3(v);
    coords[3 * (size_t)n + 1] = (int32_t)compact3(v >> 1);
    coords[3 * (size_t)n + 2] = (int32_t)compact3(v >> 2);
}

// K5 (raymarching.cu:268-289): one thread packs 32 cells = 4 bytes: eight float4 loads, one 32-bit store.
__global__ void __launch_bounds__(256) k_packbits(const float* __restrict__ grid, uint32_t N, float thresh, uint8_t* __restrict__ bitfield) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;       // word index (4 bytes of bitfield)
    const uint32_t n_words = N >> 2;
    if (w < n_words) {
        const float4* g = reinterpret_cast<const float4*>(grid) + (size_t)w * 8;
        uint32_t bits = 0;
        #pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float4 v = __ldg(g + i);
            bits |= (uint32_t)(v.x > thresh) << (4 * i) | (uint32_t)(v.y > thresh) << (4 * i + 1) |
                    (uint32_t)(v.z > thresh) << (4 * i + 2) | (uint32_t)(v.w > thresh) << (4 * i + 3);
        }
        reinterpret_cast<uint32_t*>(bitfield)[w] = bits;
    } else if (w == n_words) {                                       // tail bytes when N % 4 != 0
        for (uint32_t n = n_words * 4; n < N; ++n) {
            uint8_t b = 0;
            for (int i = 0; i < 8; ++i) b |= (uint8_t)((grid[(size_t)n * 8 + i] > thresh) << i);
            bitfield[n] = b;
        }
    }
}
__global__ void __launch_bounds__(256) k_packbits_bytes(const float* __restrict__ grid, uint32_t N, float thresh, uint8_t* __restrict__ bitfield) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    uint8_t b = 0;
    #pragma unroll
    for (int i = 0; i < 8; ++i) b |= (uint8_t)((__ldg(grid + (size_t)n * 8 + i) > thresh) << i);
    bitfield[n] = b;
}

// ------------------------------------------------------------------------------------------------ training march

// Thread -> ray mapping of the serial pass.  A ray's march is a dependent chain of ~100-250 probes; with a 4096-ray
// training batch, one ray per lane gives only 128 warps for 592 warp schedulers, and each warp iterates to the LONGEST of
// its 32 rays while executing both sides of the occupied/empty branch (profiles/r1: IPC 0.15, 1.5 % of warp slots).
// For small batches only RPW lanes of each warp carry a ray: 32/RPW x more warps in flight, max over RPW rays not 32.
template <int RPW>
__device__ __forceinline__ bool ray_of_thread(uint32_t N, uint32_t& n) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x, lane = g & 31u;
    n = (g >> 5) * RPW + lane;
    return lane < (uint32_t)RPW && n < N;
}

// pass 1 (raymarching.cu:353-400): walk the ray ONCE, count its occupied steps and record the parameter t of each.
// tbuf[n * max_steps + s] = t at which sample s of ray n is taken.
template <int RPW>
__global__ void __launch_bounds__(128) k_march_record(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                                      const float* __restrict__ nears, const float* __restrict__ fars,
                                                      const float* __restrict__ noises, uint32_t N, uint32_t max_steps,
                                                      MarchParams p, uint32_t* __restrict__ counts, float* __restrict__ tbuf) {
    uint32_t n;
    if (!ray_of_thread<RPW>(N, n)) return;
    const Ray r = load_ray(rays_o, rays_d, n, __ldg(fars + n));
    const float near = __ldg(nears + n);
    float t = ffma(step_dt(p, near), __ldg(noises + n), near);          // t0 = near + clamp(near*gamma)*noise (one FMA, l.351)
    float* tb = tbuf + (size_t)n * max_steps;
    uint32_t num = 0;
    float x, y, z, dt;
    while (t < r.far && num < max_steps) {
        const float tc = t;
        if (probe(p, r, t, x, y, z, dt)) { tb[num++] = tc; t = fadd(tc, dt); }
    }
    counts[n] = num;
}

// pass 1, constant step (dt_gamma == 0, the setting of bound-1 scenes): ONE WARP PER RAY, data-parallel.
//
// With a constant step d the serial walk only ever visits parameters of ONE fixed sequence t_0, t_{k+1} = fl(t_k + d):
// an occupied cell advances by one element, an empty cell skips to the first element >= tt (raymarching.cu:393-399 adds
// d repeatedly).  So the walk splits into
//   A  generate a chunk of 512 elements of the sequence: every lane runs a 16-add chain from a speculated seed
//      (bits(t_s) + 16*lane*c, c = the per-step bit increment inside one binade); a chain is accepted only when its
//      last value equals the next lane's seed BITWISE, otherwise the first unverified lane restarts from the exact
//      value -- the result is the sequential sum by induction, no assumption about rounding survives unverified;
//   B  probe all elements < far in parallel (the global bitfield loads overlap instead of serialising), and resolve
//      each element's successor: k+1 when occupied, else the first later element >= tt;
//   C  follow the successor links from element 0 (one shared-memory load per visited cell), recording occupied ones.
// Same probes, same floats, same order of recorded samples as the serial walk.
constexpr int kSeqChunk = 512, kSeqPerLane = kSeqChunk / 32, kSeqShift = 4;      // 2^kSeqShift == kSeqPerLane
constexpr int kSeqWarps = 1;                 // warps (= rays) per CTA: 5.2 KB of shared memory per ray, so a 4096-ray batch
                                             // (28 rays per SM) is resident in ONE wave
struct MarchSeq {
    float t[kSeqChunk + 32];             // element k at k + (k >> kSeqShift): per-lane chains and strided reads are conflict-free
    uint16_t next[kSeqChunk];            // bit 15: occupied; low bits: successor index (kSeqChunk = beyond this chunk)
    union {
        float tt[kSeqChunk];             // B: exit parameter of an empty cell, -1 for an occupied one
        uint16_t samp[kSeqChunk];        // C: indices of the occupied cells the walk visited, in order
    };
};
// ---- conservative whole-ray rejection for frame-sized marches (render): most rays of a frame cross the box without touching an
// occupied cell, and the exact walk costs ~5000 warp instructions per ray whatever it finds.  A 32^3 "any bit set" map per cascade
// level, dilated by one coarse cell, is built from the bitfield (Morton order: a coarse cell is (H/32)^3 CONSECUTIVE bits), and a
// warp samples its ray no farther apart than one coarse cell per coordinate: every point of the ray then lies in a sample's coarse
// cell or one of its 26 neighbours, so "no dilated cell hit on any level" proves that no probe of the exact walk can find an
// occupied cell -> count 0 without generating the sequence.  Rays that may hit take the exact path unchanged (bit-exact results).
constexpr uint32_t kCoarse = 32, kCoarseCells = kCoarse * kCoarse * kCoarse;

__global__ void __launch_bounds__(256) k_coarse_any(const uint8_t* __restrict__ grid, uint32_t C, uint32_t H, uint8_t* __restrict__ any) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= C * kCoarseCells) return;
    const uint32_t c = i / kCoarseCells, m = i - c * kCoarseCells;
    const uint32_t f = H / kCoarse, bytes = f * f * f / 8;                   // H a power of two >= 64: bytes = 1, 8, 64, ...
    const uint8_t* src = grid + (size_t)c * ((size_t)H * H * H / 8) + (size_t)m * bytes;
    uint32_t acc = 0;
    for (uint32_t k = 0; k < bytes; ++k) acc |= src[k];
    any[i] = acc != 0;
}

__global__ void __launch_bounds__(256) k_coarse_dilate(const uint8_t* __restrict__ any, uint32_t C, uint32_t* __restrict__ dil) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;               // (c, Z, Y, X), X fastest: a warp = one word of the map
    if (i >= C * kCoarseCells) return;
    const uint32_t c = i / kCoarseCells, q = i - c * kCoarseCells;
    const int X = (int)(q & 31u), Y = (int)((q >> 5) & 31u), Z = (int)(q >> 10);
    bool occ = false;
    for (int dz = -1; dz <= 1; ++dz) for (int dy = -1; dy <= 1; ++dy) for (int dx = -1; dx <= 1; ++dx) {
        const int x = X + dx, y = Y + dy, z = Z + dz;
        if ((unsigned)x < kCoarse && (unsigned)y < kCoarse && (unsigned)z < kCoarse)
            occ |= any[(size_t)c * kCoarseCells + morton3((uint32_t)x, (uint32_t)y, (uint32_t)z)] != 0;
    }
    const uint32_t word = __ballot_sync(0xffffffffu, occ);
    if ((threadIdx.x & 31u) == 0) dil[i >> 5] = word;
}

// warp-collective; false only when the ray provably meets no occupied cell on any cascade level
__device__ __forceinline__ bool warp_ray_may_hit(const MarchParams& p, const Ray& r, float near, const uint32_t* __restrict__ dil, uint32_t lane) {
    const float dmax = fmaxf(fabsf(r.dx), fmaxf(fabsf(r.dy), fabsf(r.dz)));
    const float span = r.far - near;
    const float step = (2.0f * fminf(1.0f, p.bound) / (float)kCoarse) / dmax;       // <= one finest-level coarse cell per coordinate
    const float fn = ceilf(span / step);
    if (!(span >= 0.0f) || !(step > 0.0f) || !(fn < 4000.0f)) return true;             // degenerate: let the exact walk decide
    const uint32_t n = (uint32_t)fn + 1u;
    for (uint32_t base = 0; base < n; base += 32) {
        const float t = fminf(near + (float)(base + lane) * step, r.far);
        const float x = clampf(fmaf(t, r.dx, r.ox), -p.bound, p.bound), y = clampf(fmaf(t, r.dy, r.oy), -p.bound, p.bound),
                    z = clampf(fmaf(t, r.dz, r.oz), -p.bound, p.bound);
        bool hit = false;
        for (int c = 0; c < p.C; ++c) {
            const float rb = 1.0f / fminf(__uint_as_float((uint32_t)(127 + c) << 23), p.bound);
            const int X = (int)clampf((x * rb + 1.0f) * (0.5f * kCoarse), 0.0f, (float)(kCoarse - 1));
            const int Y = (int)clampf((y * rb + 1.0f) * (0.5f * kCoarse), 0.0f, (float)(kCoarse - 1));
            const int Z = (int)clampf((z * rb + 1.0f) * (0.5f * kCoarse), 0.0f, (float)(kCoarse - 1));
            hit |= (__ldg(dil + (size_t)c * (kCoarseCells / 32) + (uint32_t)(Y + 32 * Z)) >> X) & 1u;
        }
        if (__any_sync(0xffffffffu, hit)) return true;
    }
    return false;
}

__device__ __forceinline__ int seq_pos(int k) { return k + (k >> kSeqShift); }

__global__ void __launch_bounds__(32 * kSeqWarps) k_march_record_const(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                                            const float* __restrict__ nears, const float* __restrict__ fars,
                                                            const float* __restrict__ noises, uint32_t N, uint32_t max_steps,
                                                            MarchParams p, uint32_t* __restrict__ counts, float* __restrict__ tbuf,
                                                            const uint32_t* __restrict__ dil, const int32_t* __restrict__ counter_in,
                                                            uint32_t* __restrict__ base_slot) {
    pdl_sync();
    __shared__ MarchSeq seq[kSeqWarps];
    const uint32_t lane = threadIdx.x & 31u, n = blockIdx.x * (uint32_t)kSeqWarps + (threadIdx.x >> 5);
    if (base_slot != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *base_slot = (uint32_t)counter_in[0];   // see k_march_expand<true>
    if (n >= N) return;                                              // warp-uniform
    MarchSeq& w = seq[threadIdx.x >> 5];
    const uint32_t full = 0xffffffffu;
    const Ray r = load_ray(rays_o, rays_d, n, __ldg(fars + n));
    const float near = __ldg(nears + n);
    if (dil != nullptr && !warp_ray_may_hit(p, r, near, dil, lane)) { if (lane == 0) counts[n] = 0; return; }
    const float d = step_dt(p, 1.0f);                                // clamp(t * 0) for every finite t
    float t_start = ffma(d, __ldg(noises + n), near);
    float* tb = tbuf + (size_t)n * max_steps;
    uint32_t num = 0;
    float pending = -1.0f;                                           // tt of a skip that ran past the previous chunk (t >= 0)
    bool have_pending = false;
    for (int chunk = 0; chunk < 4096 && t_start < r.far && num < max_steps; ++chunk) {
        // ---- A: t[0..1023] (+ t_next), exact
        uint32_t first = 0, below = 0;
        float exact = t_start, seed = t_start, v = t_start;
        while (true) {
            const uint32_t c = __float_as_uint(fadd(exact, d)) - __float_as_uint(exact);
            if (lane >= first) {
                seed = __uint_as_float(__float_as_uint(exact) + (lane - first) * (uint32_t)kSeqPerLane * c);
                v = seed; below = 0;
                float* row = w.t + lane * (uint32_t)(kSeqPerLane + 1);
                #pragma unroll 8
                for (int i = 0; i < kSeqPerLane; ++i) { row[i] = v; below += (v < r.far) ? 1u : 0u; v = fadd(v, d); }
            }
            const float nxt = __shfl_down_sync(full, seed, 1);
            const bool bad = lane >= first && lane < 31u && __float_as_uint(v) != __float_as_uint(nxt);
            const uint32_t m = __ballot_sync(full, bad);
            if (m == 0) break;
            const int l = __ffs((int)m) - 1;
            first = (uint32_t)l + 1u;
            exact = __shfl_sync(full, v, l);
        }
        const float t_next = __shfl_sync(full, v, 31);
        const int Kc = (int)__reduce_add_sync(full, below);         // elements of this chunk with t < far (a prefix: t is monotone)
        __syncwarp();
        // ---- B1: probe every element (straight-line body, unrolled: the bitfield loads of several elements are in flight
        // together);  B2: successor = k+1 when occupied, else the first later element >= tt
        #pragma unroll 4
        for (int k = (int)lane; k < Kc; k += 32) {
            float x, y, z, dt, tt;
            const bool occ = probe_cell(p, r, w.t[seq_pos(k)], x, y, z, dt, tt);
            w.tt[k] = occ ? -1.0f : tt;                              // tt >= t >= 0 for an empty cell
        }
        __syncwarp();
        const float rd = __frcp_rn(d);
        for (int k = (int)lane; k < Kc; k += 32) {
            const float tt = w.tt[k];
            uint32_t j = 0x8000u | (uint32_t)(k + 1);
            if (tt >= 0.0f) {
                // first q > k with t[q] >= tt: guess from the spacing, then correct against the exact sequence
                int q = k + max(1, (int)fminf(ceilf((tt - w.t[seq_pos(k)]) * rd), 2048.0f));
                q = min(q, kSeqChunk);
                while (q > k + 1 && w.t[seq_pos(q - 1)] >= tt) --q;
                while (q < kSeqChunk && w.t[seq_pos(q)] < tt) ++q;
                j = (uint32_t)q;
            }
            w.next[k] = (uint16_t)j;
        }
        __syncwarp();
        // B3: shorten runs of empty cells.  An empty cell records nothing, so its link may point past any number of further
        // empty cells (pointer doubling, 3 rounds: runs shrink 8x).  Links that leave the chunk are kept: the walk must end on
        // the cell whose own skip crosses the boundary, its tt is what carries over.  Lanes may read a link another lane is
        // rewriting -- old and new value are both valid targets.
        #pragma unroll 1
        for (int round = 0; round < 3; ++round) {
            for (int k = (int)lane; k < Kc; k += 32) {
                const uint32_t j = w.next[k];
                if (j < (uint32_t)Kc) {                               // k empty (bit 15 clear), j inside the probed range
                    const uint32_t j2 = w.next[j];
                    if (j2 < (uint32_t)kSeqChunk) w.next[k] = (uint16_t)j2;   // j empty too, and its skip stays in the chunk
                }
            }
            __syncwarp();
        }
        // ---- C: follow the links.  This is the only serial part (one dependent shared-memory load per visited cell), so
        // the loop body is kept branch-free: the index is stored unconditionally and the count advances on occupied cells.
        int k = 0;
        if (have_pending) {
            while (k < kSeqChunk && w.t[seq_pos(k)] < pending) ++k;
            have_pending = (k == kSeqChunk);
        }
        const uint32_t num0 = num;
        int last_k = -1;
        uint32_t last_j = 0x8000u;
        while (k < Kc && num < max_steps) {
            const uint32_t j = w.next[k];
            const uint32_t occ = j >> 15;
            w.samp[num - num0] = (uint16_t)k;                        // kept only when occ (num advances past it)
            num += occ;
            last_k = k; last_j = j;
            k = (int)(j & 0x7fffu);
        }
        if (last_j == (uint32_t)kSeqChunk) {                          // an empty cell whose skip leaves this chunk: carry its target
            float x, y, z, dt;
            probe_cell(p, r, w.t[seq_pos(last_k)], x, y, z, dt, pending);
            have_pending = true;
        }
        __syncwarp();
        for (uint32_t s = lane; s < num - num0; s += 32) tb[num0 + s] = w.t[seq_pos(w.samp[s])];
        if (Kc < kSeqChunk) break;                                   // the sequence passed `far` inside this chunk
        t_start = t_next;
        __syncwarp();
    }
    if (lane == 0) counts[n] = num;
}

// exclusive prefix scan of counts in ray order by ONE CTA of 1024 threads (N is a few thousand in training; at
// 640k rays this is ~600 block iterations of coalesced loads).  rays[n] = (n, base + excl, count);
// counter[0] += total, counter[1] += N  (raymarching.cu:405-413).
__global__ void __launch_bounds__(1024) k_march_scan(const uint32_t* __restrict__ counts, uint32_t N,
                                                     int32_t* __restrict__ rays, int32_t* __restrict__ counter) {
    pdl_sync();
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t s_base;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_base = (uint32_t)counter[0];
    __syncthreads();
    for (uint32_t start = 0; start < N; start += 1024) {
        const uint32_t n = start + threadIdx.x;
        const uint32_t c = n < N ? counts[n] : 0u;
        const uint32_t incl = warp_incl_sum_u32(c, lane);
        if (lane == 31) warp_tot[wid] = incl;
        __syncthreads();
        const uint32_t base = s_base;
        uint32_t wv = warp_tot[lane];                       // every warp scans the 32 warp totals
        const uint32_t wincl = warp_incl_sum_u32(wv, lane);
        const uint32_t woff = __shfl_sync(0xffffffffu, wincl - wv, wid);
        const uint32_t total = __shfl_sync(0xffffffffu, wincl, 31);
        if (n < N) {
            rays[3 * (size_t)n] = (int32_t)n;
            rays[3 * (size_t)n + 1] = (int32_t)(base + woff + incl - c);
            rays[3 * (size_t)n + 2] = (int32_t)c;
        }
        __syncthreads();
        if (threadIdx.x == 0) s_base = base + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) { counter[0] = (int32_t)s_base; counter[1] += (int32_t)N; }
}

// The same scan for a whole frame of rays (march_rays_all: 640 000 rays), three launches: per-4096-ray block sums, one CTA
// scanning the block sums (and updating the counter), every block scanning its own rays from its offset.
constexpr uint32_t kScanBlock = 4096;               // 1024 threads x 4 consecutive rays

__device__ __forceinline__ uint32_t block_excl_scan_1024(uint32_t v, uint32_t* warp_tot, uint32_t& total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t incl = warp_incl_sum_u32(v, lane);
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    const uint32_t wv = warp_tot[lane];
    const uint32_t wincl = warp_incl_sum_u32(wv, lane);
    const uint32_t woff = __shfl_sync(0xffffffffu, wincl - wv, wid);
    total = __shfl_sync(0xffffffffu, wincl, 31);
    __syncthreads();                                 // warp_tot may be reused
    return woff + incl - v;
}

__global__ void __launch_bounds__(1024) k_scan_partial(const uint32_t* __restrict__ counts, uint32_t N, uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t warp_tot[32];
    const uint32_t i = blockIdx.x * kScanBlock + threadIdx.x * 4;
    uint32_t v = 0;
    if (i + 4 <= N) { const uint4 c = *reinterpret_cast<const uint4*>(counts + i); v = c.x + c.y + c.z + c.w; }
    else for (uint32_t k = i; k < N; ++k) v += counts[k];
    uint32_t total;
    block_excl_scan_1024(v, warp_tot, total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) k_scan_tops(uint32_t* __restrict__ block_sums, uint32_t nb, uint32_t N, int32_t* __restrict__ counter) {
    __shared__ uint32_t warp_tot[32];
    uint32_t base = (uint32_t)counter[0];
    for (uint32_t start = 0; start < nb; start += 1024) {
        const uint32_t b = start + threadIdx.x;
        const uint32_t v = b < nb ? block_sums[b] : 0u;
        uint32_t total;
        const uint32_t excl = block_excl_scan_1024(v, warp_tot, total);
        if (b < nb) block_sums[b] = base + excl;
        base += total;
    }
    if (threadIdx.x == 0) { counter[0] = (int32_t)base; counter[1] += (int32_t)N; }
}

__global__ void __launch_bounds__(1024) k_scan_apply(const uint32_t* __restrict__ counts, uint32_t N, const uint32_t* __restrict__ block_sums,
                                                     int32_t* __restrict__ rays) {
    __shared__ uint32_t warp_tot[32];
    const uint32_t i = blockIdx.x * kScanBlock + threadIdx.x * 4;
    uint32_t c[4] = {0u, 0u, 0u, 0u};
    if (i + 4 <= N) { const uint4 q = *reinterpret_cast<const uint4*>(counts + i); c[0] = q.x; c[1] = q.y; c[2] = q.z; c[3] = q.w; }
    else for (uint32_t k = 0; k < 4 && i + k < N; ++k) c[k] = counts[i + k];
    uint32_t total;
    uint32_t off = block_sums[blockIdx.x] + block_excl_scan_1024(c[0] + c[1] + c[2] + c[3], warp_tot, total);
    #pragma unroll
    for (uint32_t k = 0; k < 4; ++k) {
        if (i + k < N) {
            rays[3 * (size_t)(i + k)] = (int32_t)(i + k);
            rays[3 * (size_t)(i + k) + 1] = (int32_t)off;
            rays[3 * (size_t)(i + k) + 2] = (int32_t)c[k];
        }
        off += c[k];
    }
}

// pass 2 (raymarching.cu:404-479), now data-parallel: ONE WARP PER RAY expands the recorded t's into samples with the
// same float operations the serial walk would have applied, and writes its slice with coalesced 12/12/8-byte rows.
//   x = clamp(o + t d);  dt = clamp(t gamma);  delta1 = (t + dt) - last_t,  last_t = previous sample's t + dt (t0 first).
// SELF_SCAN (training-sized batches, foc_march_rays_train): there is no scan launch -- every CTA sums the counts of the rays before
// its own four (<= 32 independent loads per thread at N = 4096, from L2), writes rays[n] = (n, base + prefix, count) itself, and the
// warp of the last ray publishes counter[0] = base + total, counter[1] += N.  `base` is the counter[0] the record kernel saw on
// entry (parked in scratch), as the reference's atomicAdd would have returned it.
template <bool SELF_SCAN>
__global__ void __launch_bounds__(128) k_march_expand(const float* __restrict__ rays_o, const float* __restrict__ rays_d,
                                                      const float* __restrict__ nears, const float* __restrict__ noises,
                                                      uint32_t N, uint32_t M, uint32_t max_steps, uint32_t start, MarchParams p,
                                                      int32_t* __restrict__ rays, const float* __restrict__ tbuf,
                                                      float* __restrict__ xyzs, float* __restrict__ dirs, float* __restrict__ deltas,
                                                      const uint32_t* __restrict__ counts, const uint32_t* __restrict__ base_slot,
                                                      int32_t* __restrict__ counter, int32_t* __restrict__ rows_out) {
    pdl_sync();
    const uint32_t n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    uint32_t off, num;
    uint64_t total;                                                   // end of the last ray's slice: every warp knows it
    if constexpr (SELF_SCAN) {
        __shared__ uint32_t s_part[4], s_all[4];
        const uint32_t first = blockIdx.x * 4u;                         // rays before this CTA's
        uint32_t acc = 0, all = 0;
        for (uint32_t i = threadIdx.x; i < N; i += 128) { const uint32_t c = __ldg(counts + i); all += c; if (i < first) acc += c; }
        acc = __reduce_add_sync(0xffffffffu, acc);
        all = __reduce_add_sync(0xffffffffu, all);
        if (lane == 0) { s_part[threadIdx.x >> 5] = acc; s_all[threadIdx.x >> 5] = all; }
        __syncthreads();
        if (n >= N) return;
        total = (uint64_t)__ldg(base_slot) + s_all[0] + s_all[1] + s_all[2] + s_all[3];
        off = __ldg(base_slot) + s_part[0] + s_part[1] + s_part[2] + s_part[3];
        for (uint32_t i = first; i < n; ++i) off += __ldg(counts + i);
        num = __ldg(counts + n);
        if (lane == 0) {
            rays[3 * (size_t)n] = (int32_t)n; rays[3 * (size_t)n + 1] = (int32_t)off; rays[3 * (size_t)n + 2] = (int32_t)num;
            if (n == N - 1) { counter[0] = (int32_t)(off + num); counter[1] += (int32_t)N; }
        }
    } else {
        if (n >= N) return;
        off = (uint32_t)__ldg(rays + 3 * (size_t)n + 1); num = (uint32_t)__ldg(rays + 3 * (size_t)n + 2);
        total = (uint64_t)(uint32_t)__ldg(rays + 3 * (size_t)(N - 1) + 1) + (uint32_t)__ldg(rays + 3 * (size_t)(N - 1) + 2);
    }
    // Rows no ray writes must read as zero (delta == 0 marks padding for the compositors, raymarching.cu:858; the reference
    // wrapper pre-zeroes all three buffers).  Slices are in ascending ray order, so the kept rays are a prefix: the first
    // dropped ray (off <= M < off + num) zeroes its own truncated slice [off, M) (< max_steps rows); when nothing is dropped the
    // tail [total, M) is dealt to ALL warps in 32-row pieces (M = N * max_steps while mean_count <= 0 or with force_all_rays:
    // 134 MB at 4096 rays, which one warp would take milliseconds to clear).
    {
        const bool kept = (uint64_t)off + num <= M;
        if (!kept && off < M) {
            for (uint64_t i = (uint64_t)off + lane; i < M; i += 32) {
                xyzs[3 * i] = 0.f; xyzs[3 * i + 1] = 0.f; xyzs[3 * i + 2] = 0.f;
                dirs[3 * i] = 0.f; dirs[3 * i + 1] = 0.f; dirs[3 * i + 2] = 0.f;
                reinterpret_cast<float2*>(deltas)[i] = make_float2(0.f, 0.f);
            }
        }
        // rows_out (optional): publish the number of valid rows -- the total rounded up to the 128-row tile of the MLP kernels, at
        // most M -- for consumers that read their row count from device memory (foc_*_rows); only those rows need clearing then
        uint64_t hi = M;
        if (rows_out != nullptr) {
            const uint64_t r128 = (total + 127u) & ~(uint64_t)127u;
            hi = r128 < (uint64_t)M ? r128 : (uint64_t)M;
            if (hi == 0) hi = M < 128u ? M : 128u;                     // an empty march still hands one (zero-filled) tile on
            if (n == N - 1 && lane == 0) rows_out[0] = (int32_t)hi;
        }
        for (uint64_t i = total + (uint64_t)n * 32u + lane; i < hi; i += (uint64_t)N * 32u) {
            xyzs[3 * i] = 0.f; xyzs[3 * i + 1] = 0.f; xyzs[3 * i + 2] = 0.f;
            dirs[3 * i] = 0.f; dirs[3 * i + 1] = 0.f; dirs[3 * i + 2] = 0.f;
            reinterpret_cast<float2*>(deltas)[i] = make_float2(0.f, 0.f);
        }
        if (num == 0 || !kept) return;                                // raymarching.cu:415-416
    }
    const float ox = __ldg(rays_o + 3 * (size_t)n), oy = __ldg(rays_o + 3 * (size_t)n + 1), oz = __ldg(rays_o + 3 * (size_t)n + 2);
    const float dx = __ldg(rays_d + 3 * (size_t)n), dy = __ldg(rays_d + 3 * (size_t)n + 1), dz = __ldg(rays_d + 3 * (size_t)n + 2);
    const float near = __ldg(nears + n);
    const float t0 = ffma(step_dt(p, near), __ldg(noises + n), near);
    const float* tb = tbuf + (size_t)n * max_steps + start;          // start > 0: a later depth window of the same records
    for (uint32_t s = lane; s < num; s += 32) {
        const float t = __ldg(tb + s);
        float last = t0;
        if (s + start > 0) { const float tp = __ldg(tb + (ptrdiff_t)s - 1); last = fadd(tp, step_dt(p, tp)); }
        const float dt = step_dt(p, t), tn = fadd(t, dt);
        const size_t i = (size_t)off + s;
        xyzs[3 * i] = clampf(ffma(t, dx, ox), -p.bound, p.bound);
        xyzs[3 * i + 1] = clampf(ffma(t, dy, oy), -p.bound, p.bound);
        xyzs[3 * i + 2] = clampf(ffma(t, dz, oz), -p.bound, p.bound);
        dirs[3 * i] = dx; dirs[3 * i + 1] = dy; dirs[3 * i + 2] = dz;
        reinterpret_cast<float2*>(deltas)[i] = make_float2(dt, fsub(tn, last));
    }
}

// ------------------------------------------------------------------------------------------------ inference march (K9)

// raymarching.cu:701-805: each alive ray takes <= n_step occupied samples from rays_t[id]; slot n*n_step.
__global__ void __launch_bounds__(128) k_march_rays(uint32_t n_alive, uint32_t n_step, const int32_t* __restrict__ rays_alive,
                                                    const float* __restrict__ rays_t, const float* __restrict__ rays_o,
                                                    const float* __restrict__ rays_d, const float* __restrict__ fars,
                                                    const float* __restrict__ noises, MarchParams p,
                                                    float* __restrict__ xyzs, float* __restrict__ dirs, float* __restrict__ deltas) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_alive) return;
    const uint32_t id = (uint32_t)__ldg(rays_alive + n);
    const Ray r = load_ray(rays_o, rays_d, id, __ldg(fars + id));
    float t = __ldg(rays_t + id);
    t = ffma(step_dt(p, t), __ldg(noises + n), t);
    float last_t = t, x, y, z, dt;
    const size_t base = (size_t)n * n_step;
    float* px = xyzs + 3 * base; float* pd = dirs + 3 * base; float* pl = deltas + 2 * base;
    uint32_t step = 0;
    while (t < r.far && step < n_step) {
        if (!probe(p, r, t, x, y, z, dt)) continue;
        px[0] = x; px[1] = y; px[2] = z;
        pd[0] = r.dx; pd[1] = r.dy; pd[2] = r.dz;
        t = fadd(t, dt);
        pl[0] = dt; pl[1] = fsub(t, last_t);
        last_t = t;
        px += 3; pd += 3; pl += 2; ++step;
    }
}

// ------------------------------------------------------------------------------------------------ C ABI
static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

FOC_API int foc_near_far_from_aabb(const float* rays_o, const float* rays_d, const float* aabb, uint32_t N, float min_near,
                                   float* nears, float* fars, void* stream) {
    if (N == 0) return 0;
    foc::launch(k_near_far, ceil_div(N, 256), 256, 0, S(stream), rays_o, rays_d, aabb, N, min_near, nears, fars);
    FOC_LAUNCH_CHECK(); return 0;
}
FOC_API int foc_sph_from_ray(const float* rays_o, const float* rays_d, float radius, uint32_t N, float* coords, void* stream) {
    if (N == 0) return 0;
    k_sph_from_ray<<<ceil_div(N, 256), 256, 0, S(stream)>>>(rays_o, rays_d, radius, N, coords);
    FOC_LAUNCH_CHECK(); return 0;
}
FOC_API int foc_morton3D(const int32_t* coords, uint32_t N, int32_t* indices, void* stream) {
    if (N == 0) return 0;
    k_morton3d<<<ceil_div(N, 256), 256, 0, S(stream)>>>(coords, N, indices);
    FOC_LAUNCH_CHECK(); return 0;
}
FOC_API int foc_morton3D_invert(const int32_t* indices, uint32_t N, int32_t* coords, void* stream) {
    if (N == 0) return 0;
    k_morton3d_invert<<<ceil_div(N, 256), 256, 0, S(stream)>>>(indices, N, coords);
    FOC_LAUNCH_CHECK(); return 0;
}
FOC_API int foc_packbits(const float* grid, uint32_t N, float density_thresh, uint8_t* bitfield, void* stream) {
    if (N == 0) return 0;
    const bool aligned = ((reinterpret_cast<uintptr_t>(grid) & 15u) == 0) && ((reinterpret_cast<uintptr_t>(bitfield) & 3u) == 0);
    if (aligned) k_packbits<<<ceil_div((N >> 2) + 1, 256), 256, 0, S(stream)>>>(grid, N, density_thresh, bitfield);
    else k_packbits_bytes<<<ceil_div(N, 256), 256, 0, S(stream)>>>(grid, N, density_thresh, bitfield);
    FOC_LAUNCH_CHECK(); return 0;
}

// scratch = counts [N] u32 (rounded up to 256 B) + block sums of the multi-block scan + recorded sample parameters [N, max_steps] f32
static inline size_t march_block_sums_offset(uint32_t N) { return (((size_t)N * sizeof(uint32_t)) + 255) & ~(size_t)255; }
static inline size_t march_coarse_offset(uint32_t N) {
    return march_block_sums_offset(N) + ((((size_t)ceil_div(N, kScanBlock) + 1) * sizeof(uint32_t) + 255) & ~(size_t)255);
}
constexpr uint32_t kRejectMinRays = 16384;          // below this the march is one latency-bound wave: the pre-test would only add launches
constexpr size_t kCoarseBytes = 8 * ((size_t)kCoarseCells + kCoarseCells / 8);      // C <= 8: any-bytes + dilated bits
static inline size_t march_counts_bytes(uint32_t N) { return march_coarse_offset(N) + (N >= kRejectMinRays ? kCoarseBytes : 0); }
FOC_API size_t foc_march_rays_train_scratch_bytes(uint32_t N, uint32_t max_steps) {
    return march_counts_bytes(N) + (size_t)N * max_steps * sizeof(float);
}

// small N (a training batch) is latency-bound: one warp per CTA spreads the rays over all SMs
static inline uint32_t march_block(uint32_t n) { return n <= (uint32_t)kNumSMs * 32u * 8u ? 32u : 128u; }

static int scan_counts(const uint32_t* counts, uint32_t N, int32_t* rays, int32_t* counter, uint32_t* block_sums, cudaStream_t st);

// record + scan: rays[n] = (n, offset, count), counter[0] += total, counter[1] += N; the recorded parameters stay in `scratch`
// self_scan: leave the scan to k_march_expand<true> (the record kernel parks counter[0] in the first block-sum slot, unused at that size)
static inline bool march_self_scan_ok(float dt_gamma, uint32_t N) { return dt_gamma == 0.0f && N <= 8192u; }   // every CTA re-reads all N counts: beyond that the scan kernels are cheaper

static int march_count_impl(const float* rays_o, const float* rays_d, const uint8_t* grid, float bound, float dt_gamma,
                            uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, const float* nears, const float* fars,
                            int32_t* rays, int32_t* counter, const float* noises, void* scratch, void* stream, bool self_scan) {
    if (N == 0) return 0;
    if (C == 0 || C > 8 || H == 0 || H > 1024 || (uint64_t)C * H * H * H > (1ull << 24)) return FOC_ERR_UNSUPPORTED;  // fp32 cell index exact (H1)
    const MarchParams p = make_march_params(grid, bound, dt_gamma, max_steps, C, H);
    uint32_t* counts = static_cast<uint32_t*>(scratch);
    float* tbuf = reinterpret_cast<float*>(static_cast<uint8_t*>(scratch) + march_counts_bytes(N));
    // constant step: warp-per-ray parallel walk; else serial walk, 4 rays per warp for small batches (see ray_of_thread)
    if (dt_gamma == 0.0f) {
        const uint32_t* dil = nullptr;
        if (N >= kRejectMinRays && H >= 64 && (H & (H - 1)) == 0) {      // frame-sized call: build the dilated coarse occupancy first
            uint8_t* any = static_cast<uint8_t*>(scratch) + march_coarse_offset(N);
            uint32_t* d32 = reinterpret_cast<uint32_t*>(any + 8 * (size_t)kCoarseCells);
            k_coarse_any<<<ceil_div(C * kCoarseCells, 256), 256, 0, S(stream)>>>(grid, C, H, any);
            k_coarse_dilate<<<ceil_div(C * kCoarseCells, 256), 256, 0, S(stream)>>>(any, C, d32);
            FOC_LAUNCH_CHECK();
            dil = d32;
        }
        uint32_t* base_slot = self_scan ? reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(scratch) + march_block_sums_offset(N)) : nullptr;
        foc::launch(k_march_record_const, ceil_div(N, (uint32_t)kSeqWarps), 32 * kSeqWarps, 0, S(stream), rays_o, rays_d, nears, fars, noises, N, max_steps, p, counts, tbuf, dil,
                    (const int32_t*)counter, base_slot);
    }
    else if (N <= (uint32_t)kNumSMs * 32u * 8u)
        k_march_record<4><<<ceil_div(ceil_div(N, 4) * 32, 128), 128, 0, S(stream)>>>(rays_o, rays_d, nears, fars, noises, N, max_steps, p, counts, tbuf);
    else
        k_march_record<32><<<ceil_div(N, 128), 128, 0, S(stream)>>>(rays_o, rays_d, nears, fars, noises, N, max_steps, p, counts, tbuf);
    FOC_LAUNCH_CHECK();
    if (self_scan) return 0;
    return scan_counts(counts, N, rays, counter, reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(scratch) + march_block_sums_offset(N)), S(stream));
}

FOC_API int foc_march_rays_count(const float* rays_o, const float* rays_d, const uint8_t* grid, float bound, float dt_gamma,
                                 uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, const float* nears, const float* fars,
                                 int32_t* rays, int32_t* counter, const float* noises, void* scratch, void* stream) {
    return march_count_impl(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, nears, fars, rays, counter, noises, scratch, stream, false);
}

// expand the records of foc_march_rays_count (same scratch, same rays) into samples; rays whose slice would pass M write nothing
FOC_API int foc_march_rays_expand(const float* rays_o, const float* rays_d, float bound, float dt_gamma, uint32_t max_steps, uint32_t N,
                                  uint32_t C, uint32_t H, uint32_t M, const float* nears, const int32_t* rays, const float* noises,
                                  const void* scratch, float* xyzs, float* dirs, float* deltas, void* stream) {
    if (N == 0) return 0;
    if (C == 0 || C > 8 || H == 0 || H > 1024) return FOC_ERR_UNSUPPORTED;
    const MarchParams p = make_march_params(nullptr, bound, dt_gamma, max_steps, C, H);
    const float* tbuf = reinterpret_cast<const float*>(static_cast<const uint8_t*>(scratch) + march_counts_bytes(N));
    foc::launch(k_march_expand<false>, ceil_div(N, 4), 128, 0, S(stream), rays_o, rays_d, nears, noises, N, M, max_steps, 0u, p, const_cast<int32_t*>(rays), tbuf, xyzs, dirs, deltas,
                (const uint32_t*)nullptr, (const uint32_t*)nullptr, (int32_t*)nullptr, (int32_t*)nullptr);
    FOC_LAUNCH_CHECK(); return 0;
}

// ---- depth windows over the records of foc_march_rays_count (render: early termination without a host-driven loop)
__global__ void __launch_bounds__(256) k_window_counts(const uint32_t* __restrict__ counts, const uint8_t* __restrict__ alive,
                                                       uint32_t N, uint32_t start, uint32_t width, uint32_t* __restrict__ wcounts) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const uint32_t c = counts[n];
    uint32_t w = c > start ? min(c - start, width) : 0u;
    if (alive && !alive[n]) w = 0u;
    wcounts[n] = w;
}

static int scan_counts(const uint32_t* counts, uint32_t N, int32_t* rays, int32_t* counter, uint32_t* block_sums, cudaStream_t st) {
    if (N <= 8 * kScanBlock) {
        foc::launch(k_march_scan, 1, 1024, 0, st, counts, N, rays, counter);
    } else {
        const uint32_t nb = ceil_div(N, kScanBlock);
        k_scan_partial<<<nb, 1024, 0, st>>>(counts, N, block_sums);
        k_scan_tops<<<1, 1024, 0, st>>>(block_sums, nb, N, counter);
        k_scan_apply<<<nb, 1024, 0, st>>>(counts, N, block_sums, rays);
    }
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_march_window_count(void* scratch, uint32_t N, uint32_t start, uint32_t width, const uint8_t* alive,
                                   uint32_t* wcounts, int32_t* rays_w, int32_t* counter, void* stream) {
    if (N == 0) return 0;
    const uint32_t* counts = static_cast<const uint32_t*>(scratch);
    uint32_t* block_sums = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(scratch) + march_block_sums_offset(N));
    k_window_counts<<<ceil_div(N, 256), 256, 0, S(stream)>>>(counts, alive, N, start, width, wcounts);
    FOC_LAUNCH_CHECK();
    return scan_counts(wcounts, N, rays_w, counter, block_sums, S(stream));
}

FOC_API int foc_march_window_expand(const float* rays_o, const float* rays_d, float bound, float dt_gamma, uint32_t max_steps, uint32_t N,
                                    uint32_t C, uint32_t H, uint32_t M, uint32_t start, const float* nears, const int32_t* rays_w,
                                    const float* noises, const void* scratch, float* xyzs, float* dirs, float* deltas, void* stream) {
    if (N == 0) return 0;
    if (C == 0 || C > 8 || H == 0 || H > 1024 || start >= max_steps) return FOC_ERR_UNSUPPORTED;
    const MarchParams p = make_march_params(nullptr, bound, dt_gamma, max_steps, C, H);
    const float* tbuf = reinterpret_cast<const float*>(static_cast<const uint8_t*>(scratch) + march_counts_bytes(N));
    foc::launch(k_march_expand<false>, ceil_div(N, 4), 128, 0, S(stream), rays_o, rays_d, nears, noises, N, M, max_steps, start, p, const_cast<int32_t*>(rays_w), tbuf, xyzs, dirs, deltas,
                (const uint32_t*)nullptr, (const uint32_t*)nullptr, (int32_t*)nullptr, (int32_t*)nullptr);
    FOC_LAUNCH_CHECK(); return 0;
}

static int march_train_impl(const float* rays_o, const float* rays_d, const uint8_t* grid, float bound, float dt_gamma,
                            uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, uint32_t M, const float* nears,
                            const float* fars, float* xyzs, float* dirs, float* deltas, int32_t* rays, int32_t* counter,
                            const float* noises, void* scratch, void* stream, int32_t* rows_out) {
    if (N != 0 && march_self_scan_ok(dt_gamma, N)) {                   // two launches: record, then an expansion that scans for itself
        if (int rc = march_count_impl(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, nears, fars, rays, counter, noises, scratch, stream, true)) return rc;
        const MarchParams p = make_march_params(nullptr, bound, dt_gamma, max_steps, C, H);
        const uint32_t* counts = static_cast<const uint32_t*>(scratch);
        const uint32_t* base_slot = reinterpret_cast<const uint32_t*>(static_cast<const uint8_t*>(scratch) + march_block_sums_offset(N));
        const float* tbuf = reinterpret_cast<const float*>(static_cast<const uint8_t*>(scratch) + march_counts_bytes(N));
        foc::launch(k_march_expand<true>, ceil_div(N, 4), 128, 0, S(stream), rays_o, rays_d, nears, noises, N, M, max_steps, 0u, p, rays, tbuf, xyzs, dirs, deltas,
                    counts, base_slot, counter, rows_out);
        FOC_LAUNCH_CHECK(); return 0;
    }
    if (N == 0) return 0;
    if (int rc = foc_march_rays_count(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, nears, fars, rays, counter, noises, scratch, stream)) return rc;
    if (C == 0 || C > 8 || H == 0 || H > 1024) return FOC_ERR_UNSUPPORTED;
    const MarchParams p = make_march_params(nullptr, bound, dt_gamma, max_steps, C, H);
    const float* tbuf = reinterpret_cast<const float*>(static_cast<const uint8_t*>(scratch) + march_counts_bytes(N));
    foc::launch(k_march_expand<false>, ceil_div(N, 4), 128, 0, S(stream), rays_o, rays_d, nears, noises, N, M, max_steps, 0u, p, rays, tbuf, xyzs, dirs, deltas,
                (const uint32_t*)nullptr, (const uint32_t*)nullptr, (int32_t*)nullptr, rows_out);
    FOC_LAUNCH_CHECK(); return 0;
}

FOC_API int foc_march_rays_train(const float* rays_o, const float* rays_d, const uint8_t* grid, float bound, float dt_gamma,
                                 uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, uint32_t M, const float* nears,
                                 const float* fars, float* xyzs, float* dirs, float* deltas, int32_t* rays, int32_t* counter,
                                 const float* noises, void* scratch, void* stream) {
    return march_train_impl(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, M, nears, fars, xyzs, dirs, deltas, rays, counter, noises,
                            scratch, stream, nullptr);
}

// foc_march_rays_train that also PUBLISHES the number of valid sample rows in device memory: rows_out[0] = min(M, total rounded up to 128).
// Buffers sized once (M rows, M % 128 == 0) can then be reused by a CUDA-graph replay for steps of any sample count <= M: the consumers
// (foc_grid_encode_forward_rows, foc_field_forward, foc_ffmlp_backward_head, foc_grid_encode_backward_rows, foc_composite_mse_backward_rows)
// read the count instead of processing padding rows, and only rows [total, rows_out[0]) are zero-filled.  No reference counterpart: the
// reference re-allocates per step and drops the rays that do not fit its running-mean-sized buffers (raymarching.py:198-203).
FOC_API int foc_march_rays_train_rows(const float* rays_o, const float* rays_d, const uint8_t* grid, float bound, float dt_gamma,
                                      uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, uint32_t M, const float* nears,
                                      const float* fars, float* xyzs, float* dirs, float* deltas, int32_t* rays, int32_t* counter,
                                      const float* noises, int32_t* rows_out, void* scratch, void* stream) {
    if (rows_out == nullptr || M % 128u != 0) return FOC_ERR_UNSUPPORTED;
    return march_train_impl(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, M, nears, fars, xyzs, dirs, deltas, rays, counter, noises,
                            scratch, stream, rows_out);
}

FOC_API int foc_march_rays(uint32_t n_alive, uint32_t n_step, const int32_t* rays_alive, const float* rays_t,
                           const float* rays_o, const float* rays_d, float bound, float dt_gamma, uint32_t max_steps,
                           uint32_t C, uint32_t H, const uint8_t* grid, const float* nears, const float* fars, float* xyzs,
                           float* dirs, float* deltas, const float* noises, void* stream) {
    (void)nears;
    if (n_alive == 0) return 0;
    if (C == 0 || C > 8 || H == 0 || H > 1024 || (uint64_t)C * H * H * H > (1ull << 24)) return FOC_ERR_UNSUPPORTED;
    const MarchParams p = make_march_params(grid, bound, dt_gamma, max_steps, C, H);
    const uint32_t bs = march_block(n_alive);
    k_march_rays<<<ceil_div(n_alive, bs), bs, 0, S(stream)>>>(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, fars, noises, p, xyzs, dirs, deltas);
    FOC_LAUNCH_CHECK(); return 0;
}
