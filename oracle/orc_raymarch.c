/*
 * oracle/orc_raymarch.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C CPU restatement of the ray-marching / compositing half of FOC-NeRF's per-ray hot path.
 * Each function cites the reference source it follows (paths relative to /root/reference).
 * The reference has no CPU implementation and no tests for this path;
 * this restatement is pinned against golden vectors produced by running the reference's own
 * CUDA extension (oracle/_ref, built by oracle/build_ref.py) on a B200: see tests/golden/.
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp (see oracle/Makefile).  -ffp-contract=off matters:
 * every fused multiply-add nvcc emits for the reference kernel (checked in the SASS of
 * oracle/_ref/raymarching) is written here as an explicit fmaf(), everything else must stay
 * an unfused IEEE binary32 operation so integer results (sample counts, offsets) are bit-exact.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <float.h>

#define ORC_API __attribute__((visibility("default")))

static inline float orc_clampf(float x, float lo, float hi) { return fminf(hi, fmaxf(lo, x)); }
static inline float orc_sign1(float x) { return copysignf(1.0f, x); }

/* raymarching.cu:42-47 : cascade level from the largest |coordinate| (frexp exponent, clamped) */
static inline int orc_mip_from_pos(float x, float y, float z, int cascades) {
    float mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
    int e; (void)frexpf(mx, &e);
    return (int)fminf((float)(cascades - 1), fmaxf(0.0f, (float)e));
}
/* raymarching.cu:49-54 : cascade level from the step size */
static inline int orc_mip_from_dt(float dt, float H, int cascades) {
    float mx = (dt * H) * 0.5f;   /* reference multiplies by the double literal 0.5: exact either way */
    int e; (void)frexpf(mx, &e);
    return (int)fminf((float)(cascades - 1), fmaxf(0.0f, (float)e));
}
/* raymarching.cu:56-71 : 10-bit x 3 Morton interleave */
static inline uint32_t orc_spread3(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
static inline uint32_t orc_morton(uint32_t x, uint32_t y, uint32_t z) {
    return orc_spread3(x) | (orc_spread3(y) << 1) | (orc_spread3(z) << 2);
}
/* raymarching.cu:73-81 */
static inline uint32_t orc_compact3(uint32_t x) {
    x &= 0x49249249u;
    x = (x | (x >> 2)) & 0xc30c30c3u;
    x = (x | (x >> 4)) & 0x0f00f00fu;
    x = (x | (x >> 8)) & 0xff0000ffu;
    x = (x | (x >> 16)) & 0x0000ffffu;
    return x;
}

/* raymarching.cu:92-145 : slab test, near clamped to min_near, miss -> FLT_MAX for both */
ORC_API void orc_near_far_from_aabb(const float *rays_o, const float *rays_d, const float *aabb,
                                    uint32_t N, float min_near, float *nears, float *fars) {
    #pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < (int64_t)N; ++n) {
        const float *o = rays_o + 3 * n, *d = rays_d + 3 * n;
        float tn = 0.f, tf = 0.f; int miss = 0;
        for (int a = 0; a < 3 && !miss; ++a) {
            float r = 1.0f / d[a];
            float lo = (aabb[a] - o[a]) * r, hi = (aabb[a + 3] - o[a]) * r;
            if (lo > hi) { float s = lo; lo = hi; hi = s; }
            if (a == 0) { tn = lo; tf = hi; continue; }
            if (tn > hi || lo > tf) { miss = 1; break; }
            if (lo > tn) tn = lo;
            if (hi < tf) tf = hi;
        }
        if (miss) { nears[n] = fars[n] = FLT_MAX; continue; }
        if (tn < min_near) tn = min_near;
        nears[n] = tn; fars[n] = tf;
    }
}

/* raymarching.cu:163-198 : far hit with the background sphere -> (theta, phi) in [-1,1]^2 */
ORC_API void orc_sph_from_ray(const float *rays_o, const float *rays_d, float radius, uint32_t N, float *coords) {
    const float RPI = 0.3183098861837907f;
    for (uint32_t n = 0; n < N; ++n) {
        const float *o = rays_o + 3 * n, *d = rays_d + 3 * n;
        /* FMA placement as nvcc contracts the reference source (read off the SASS of the reference build, see csrc/raymarch.cu) */
        float A = fmaf(d[2], d[2], fmaf(d[0], d[0], d[1] * d[1]));
        float B = fmaf(o[2], d[2], fmaf(o[0], d[0], o[1] * d[1]));
        float Cc = fmaf(-radius, radius, fmaf(o[2], o[2], fmaf(o[0], o[0], o[1] * o[1])));
        float t = (-B + sqrtf(fmaf(B, B, -(A * Cc)))) / A;
        float x = fmaf(t, d[0], o[0]), y = fmaf(t, d[1], o[1]), z = fmaf(t, d[2], o[2]);
        float theta = atan2f(sqrtf(fmaf(x, x, z * z)), y);
        float phi = atan2f(z, x);
        coords[2 * n] = fmaf(2.0f * theta, RPI, -1.0f);
        coords[2 * n + 1] = phi * RPI;
    }
}

/* raymarching.cu:214-226 */
ORC_API void orc_morton3D(const int32_t *coords, uint32_t N, int32_t *indices) {
    for (uint32_t n = 0; n < N; ++n)
        indices[n] = (int32_t)orc_morton((uint32_t)coords[3 * n], (uint32_t)coords[3 * n + 1], (uint32_t)coords[3 * n + 2]);
}
/* raymarching.cu:237-254 */
ORC_API void orc_morton3D_invert(const int32_t *indices, uint32_t N, int32_t *coords) {
    for (uint32_t n = 0; n < N; ++n) {
        uint32_t v = (uint32_t)indices[n];
        coords[3 * n] = (int32_t)orc_compact3(v);
        coords[3 * n + 1] = (int32_t)orc_compact3(v >> 1);
        coords[3 * n + 2] = (int32_t)orc_compact3(v >> 2);
    }
}
/* raymarching.cu:268-289 : bit i of byte n = grid[8n+i] > thresh */
ORC_API void orc_packbits(const float *grid, uint32_t N, float thresh, uint8_t *bitfield) {
    #pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < (int64_t)N; ++n) {
        uint8_t b = 0;
        for (int i = 0; i < 8; ++i) if (grid[8 * n + i] > thresh) b |= (uint8_t)(1u << i);
        bitfield[n] = b;
    }
}

/* ---- the marcher: one ray's walk through the cascaded Morton bitfield (raymarching.cu:335-399) ---- */
typedef struct {
    float ox, oy, oz, dx, dy, dz, rdx, rdy, rdz;
    float bound, dt_gamma, dt_min, dt_max, rH, H3, Hf, far;
    int C; uint32_t H; const uint8_t *grid;
} orc_ray_t;

static inline void orc_ray_setup(orc_ray_t *r, const float *o, const float *d, const uint8_t *grid, float bound,
                                 float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t H, float far) {
    r->ox = o[0]; r->oy = o[1]; r->oz = o[2]; r->dx = d[0]; r->dy = d[1]; r->dz = d[2];
    r->rdx = 1.0f / d[0]; r->rdy = 1.0f / d[1]; r->rdz = 1.0f / d[2];
    r->bound = bound; r->dt_gamma = dt_gamma;
    const float two_sqrt3 = 2.0f * 1.7320508075688772f;                       /* = 3.46410155296..f in the SASS */
    r->dt_min = two_sqrt3 / (float)max_steps;
    r->dt_max = (two_sqrt3 * (float)(1 << (C - 1))) / (float)H;
    r->rH = 1.0f / (float)H; r->Hf = (float)H; r->H3 = (float)(H * H * H);
    r->C = (int)C; r->H = H; r->grid = grid; r->far = far;
}
static inline float orc_dt(const orc_ray_t *r, float t) { return orc_clampf(t * r->dt_gamma, r->dt_min, r->dt_max); }

/* One step at parameter t. Returns 1 and fills (x,y,z,dt) if the cell is occupied; otherwise
 * returns 0 and advances *t past the empty cell (at least one dt). */
static inline int orc_probe(const orc_ray_t *r, float *t, float *px, float *py, float *pz, float *pdt) {
    const float tc = *t;
    const float x = orc_clampf(fmaf(tc, r->dx, r->ox), -r->bound, r->bound);
    const float y = orc_clampf(fmaf(tc, r->dy, r->oy), -r->bound, r->bound);
    const float z = orc_clampf(fmaf(tc, r->dz, r->oz), -r->bound, r->bound);
    const float dt = orc_dt(r, tc);
    int lp = orc_mip_from_pos(x, y, z, r->C), ld = orc_mip_from_dt(dt, r->Hf, r->C);
    const int level = lp > ld ? lp : ld;
    const float mip_bound = fminf(scalbnf(1.0f, level), r->bound);
    const float mip_rbound = 1.0f / mip_bound;
    /* 0.5 * (x*rb + 1) * H is evaluated in double by the reference (literal 0.5) then narrowed */
    const float Hm1 = (float)(r->H - 1);
    const int nx = (int)orc_clampf((float)(0.5 * (double)fmaf(x, mip_rbound, 1.0f) * (double)r->H), 0.0f, Hm1);
    const int ny = (int)orc_clampf((float)(0.5 * (double)fmaf(y, mip_rbound, 1.0f) * (double)r->H), 0.0f, Hm1);
    const int nz = (int)orc_clampf((float)(0.5 * (double)fmaf(z, mip_rbound, 1.0f) * (double)r->H), 0.0f, Hm1);
    const uint32_t index = (uint32_t)fmaf((float)level, r->H3, (float)orc_morton((uint32_t)nx, (uint32_t)ny, (uint32_t)nz));
    const int occ = (r->grid[index >> 3] >> (index & 7u)) & 1;
    if (occ) { *px = x; *py = y; *pz = z; *pdt = dt; return 1; }
    /* distance to the next cell boundary along each axis */
    const float tx = fmaf(mip_bound, fmaf((fmaf(0.5f, orc_sign1(r->dx), (float)nx + 0.5f)) * r->rH, 2.0f, -1.0f), -x) * r->rdx;
    const float ty = fmaf(mip_bound, fmaf((fmaf(0.5f, orc_sign1(r->dy), (float)ny + 0.5f)) * r->rH, 2.0f, -1.0f), -y) * r->rdy;
    const float tz = fmaf(mip_bound, fmaf((fmaf(0.5f, orc_sign1(r->dz), (float)nz + 0.5f)) * r->rH, 2.0f, -1.0f), -z) * r->rdz;
    const float tt = tc + fmaxf(0.0f, fminf(tx, fminf(ty, tz)));
    float tn = tc;
    do { tn += orc_dt(r, tn); } while (tn < tt);
    *t = tn;
    return 0;
}

/* raymarching.cu:312-480.  Rays are processed in ray order, so the slots the reference hands out by
 * arrival-order atomics (l.405-406, nondeterministic there) come out as an exclusive scan here --
 * one of the orders the reference itself can produce. */
ORC_API void orc_march_rays_train(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound,
                                  float dt_gamma, uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, uint32_t M,
                                  const float *nears, const float *fars, float *xyzs, float *dirs, float *deltas,
                                  int32_t *rays, int32_t *counter, const float *noises) {
    for (uint32_t n = 0; n < N; ++n) {
        orc_ray_t r; orc_ray_setup(&r, rays_o + 3 * n, rays_d + 3 * n, grid, bound, dt_gamma, max_steps, C, H, fars[n]);
        const float near = nears[n];
        const float t0 = fmaf(orc_dt(&r, near), noises[n], near);
        float t = t0, x, y, z, dt; uint32_t num = 0;
        while (t < r.far && num < max_steps) if (orc_probe(&r, &t, &x, &y, &z, &dt)) { ++num; t += dt; }
        const uint32_t point_index = (uint32_t)counter[0]; counter[0] += (int32_t)num;
        const uint32_t ray_index = (uint32_t)counter[1]; counter[1] += 1;
        rays[3 * ray_index] = (int32_t)n; rays[3 * ray_index + 1] = (int32_t)point_index; rays[3 * ray_index + 2] = (int32_t)num;
        if (num == 0 || point_index + num > M) continue;
        float *px = xyzs + 3 * (size_t)point_index, *pd = dirs + 3 * (size_t)point_index, *pl = deltas + 2 * (size_t)point_index;
        t = t0; float last_t = t; uint32_t step = 0;
        while (t < r.far && step < num) {
            if (!orc_probe(&r, &t, &x, &y, &z, &dt)) continue;
            px[0] = x; px[1] = y; px[2] = z; pd[0] = r.dx; pd[1] = r.dy; pd[2] = r.dz;
            t += dt; pl[0] = dt; pl[1] = t - last_t; last_t = t;
            px += 3; pd += 3; pl += 2; ++step;
        }
    }
}

/* raymarching.cu:701-805 : inference march, <= n_step samples per alive ray from rays_t[id], fixed slots */
ORC_API void orc_march_rays(uint32_t n_alive, uint32_t n_step, const int32_t *rays_alive, const float *rays_t,
                            const float *rays_o, const float *rays_d, float bound, float dt_gamma, uint32_t max_steps,
                            uint32_t C, uint32_t H, const uint8_t *grid, const float *nears, const float *fars,
                            float *xyzs, float *dirs, float *deltas, const float *noises) {
    (void)nears;
    #pragma omp parallel for schedule(dynamic, 256)
    for (int64_t n = 0; n < (int64_t)n_alive; ++n) {
        const int32_t id = rays_alive[n];
        orc_ray_t r; orc_ray_setup(&r, rays_o + 3 * (size_t)id, rays_d + 3 * (size_t)id, grid, bound, dt_gamma, max_steps, C, H, fars[id]);
        float t = rays_t[id];
        t = fmaf(orc_dt(&r, t), noises[n], t);
        float last_t = t, x, y, z, dt; uint32_t step = 0;
        float *px = xyzs + 3 * (size_t)n * n_step, *pd = dirs + 3 * (size_t)n * n_step, *pl = deltas + 2 * (size_t)n * n_step;
        while (t < r.far && step < n_step) {
            if (!orc_probe(&r, &t, &x, &y, &z, &dt)) continue;
            px[0] = x; px[1] = y; px[2] = z; pd[0] = r.dx; pd[1] = r.dy; pd[2] = r.dz;
            t += dt; pl[0] = dt; pl[1] = t - last_t; last_t = t;
            px += 3; pd += 3; pl += 2; ++step;
        }
    }
}
