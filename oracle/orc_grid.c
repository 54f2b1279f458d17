/*
 * oracle/orc_grid.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * CPU restatement of the multiresolution hash / tiled grid encoder (gridencoder/src/gridencoder.cu),
 * for D in 2..5 and any C, fp32 or fp16 tables.  fp16 tables follow the reference's rounding sequence
 * (c10::Half arithmetic: every product w*g is narrowed to half, every running sum is float-add then
 * narrowed to half, gridencoder.cu:164,187) so the forward encoding can be compared bit-for-bit where
 * the per-level scale agrees.  The scale uses exp2f(); the GPU's exp2f is a 2-ulp approximation for
 * non-integer arguments, so level scales may differ in the last bit -> trilinear results then differ
 * at the 1e-6 level, well inside the 1e-3 tolerance of BASELINE.json.
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))
#define ORC_MAXD 5

typedef _Float16 orc_half;
static inline float h2f(uint16_t h) { orc_half v; memcpy(&v, &h, 2); return (float)v; }
static inline uint16_t f2h(float f) { orc_half v = (orc_half)f; uint16_t h; memcpy(&h, &v, 2); return h; }
static inline float rh(float f) { return (float)(orc_half)f; }          /* round through binary16 */

/* gridencoder.cu:50-84 : dense index while the running stride fits, else the xor-of-primes hash */
static inline uint32_t orc_grid_index(uint32_t D, uint32_t gridtype, int align_corners, uint32_t hashmap_size,
                                      uint32_t resolution, const uint32_t *pg) {
    static const uint32_t primes[7] = {1u, 2654435761u, 805459861u, 3674653429u, 2097192037u, 1434869437u, 2165219737u};
    uint32_t stride = 1, index = 0;
    for (uint32_t d = 0; d < D && stride <= hashmap_size; ++d) {
        index += pg[d] * stride;
        stride *= align_corners ? resolution : (resolution + 1);
    }
    if (gridtype == 0 && stride > hashmap_size) {
        index = 0;
        for (uint32_t d = 0; d < D; ++d) index ^= pg[d] * primes[d];
    }
    return index % hashmap_size;
}

/* Optional override of the per-level scales (test hook): the GPU's exp2f is a 2-ulp approximation for non-integer
 * arguments, so to compare fp16 encodings BITWISE against the reference's extension the oracle can be handed the
 * scales the device computed (stored in the golden fixtures).  n == 0 restores the host exp2f. */
static float orc_scale_override[64];
static uint32_t orc_scale_override_n = 0;
ORC_API void orc_grid_set_level_scales(const float *scales, uint32_t n) {
    orc_scale_override_n = n > 64 ? 64 : n;
    for (uint32_t i = 0; i < orc_scale_override_n; ++i) orc_scale_override[i] = scales[i];
}

typedef struct { float scale; uint32_t resolution, hashmap_size; } orc_level_t;
static inline orc_level_t orc_level(const int32_t *offsets, uint32_t level, float S, uint32_t H) {
    orc_level_t l;
    l.hashmap_size = (uint32_t)(offsets[level + 1] - offsets[level]);
    l.scale = level < orc_scale_override_n ? orc_scale_override[level]
                                           : fmaf(exp2f((float)level * S), (float)H, -1.0f);   /* gridencoder.cu:138 (FMA in SASS) */
    l.resolution = (uint32_t)ceilf(l.scale) + 1;
    return l;
}
/* gridencoder.cu:146-159 */
static inline void orc_locate(uint32_t D, const float *x, float scale, int align_corners, uint32_t interp,
                              float *pos, float *pos_deriv, uint32_t *pg) {
    for (uint32_t d = 0; d < D; ++d) {
        float p = fmaf(x[d], scale, align_corners ? 0.0f : 0.5f);
        float fl = floorf(p);
        pg[d] = (uint32_t)fl;
        p -= (float)pg[d];
        if (interp == 1) { pos_deriv[d] = 6.0f * p * (1.0f - p); p = p * p * (3.0f - 2.0f * p); }
        else pos_deriv[d] = 1.0f;
        pos[d] = p;
    }
}

/* gridencoder.cu:87-245.  outputs [L,B,C]; dy_dx [B,L,D,C] (may be NULL); table/outputs are uint16 halfs
 * when is_half != 0, else floats. */
ORC_API void orc_grid_encode_forward(const float *inputs, const void *embeddings, const int32_t *offsets, void *outputs,
                                     uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, void *dy_dx,
                                     uint32_t gridtype, int align_corners, uint32_t interp, int is_half) {
    #pragma omp parallel for collapse(2) schedule(static)
    for (int64_t level = 0; level < (int64_t)L; ++level)
    for (int64_t b = 0; b < (int64_t)B; ++b) {
        const float *x = inputs + (size_t)b * D;
        const size_t obase = ((size_t)level * B + (size_t)b) * C;
        const size_t gbase = (size_t)(uint32_t)offsets[level] * C;
        const size_t dbase = (((size_t)b * L + (size_t)level) * D) * C;
        int oob = 0;
        for (uint32_t d = 0; d < D; ++d) if (x[d] < 0.0f || x[d] > 1.0f) oob = 1;
        if (oob) {
            for (uint32_t c = 0; c < C; ++c) { if (is_half) ((uint16_t *)outputs)[obase + c] = 0; else ((float *)outputs)[obase + c] = 0.0f; }
            if (dy_dx) for (uint32_t i = 0; i < D * C; ++i) { if (is_half) ((uint16_t *)dy_dx)[dbase + i] = 0; else ((float *)dy_dx)[dbase + i] = 0.0f; }
            continue;
        }
        const orc_level_t lv = orc_level(offsets, (uint32_t)level, S, H);
        float pos[ORC_MAXD], pder[ORC_MAXD]; uint32_t pg[ORC_MAXD], pl[ORC_MAXD];
        orc_locate(D, x, lv.scale, align_corners, interp, pos, pder, pg);
        float res[8] = {0};
        for (uint32_t idx = 0; idx < (1u << D); ++idx) {
            float w = 1.0f;
            for (uint32_t d = 0; d < D; ++d) {
                if ((idx >> d) & 1u) { w *= pos[d]; pl[d] = pg[d] + 1; } else { w *= 1.0f - pos[d]; pl[d] = pg[d]; }
            }
            const size_t e = gbase + (size_t)orc_grid_index(D, gridtype, align_corners, lv.hashmap_size, lv.resolution, pl) * C;
            for (uint32_t c = 0; c < C; ++c) {
                if (is_half) res[c] = rh(res[c] + rh(w * h2f(((const uint16_t *)embeddings)[e + c])));
                else         res[c] = fmaf(w, ((const float *)embeddings)[e + c], res[c]);
            }
        }
        for (uint32_t c = 0; c < C; ++c) { if (is_half) ((uint16_t *)outputs)[obase + c] = f2h(res[c]); else ((float *)outputs)[obase + c] = res[c]; }
        if (!dy_dx) continue;
        for (uint32_t gd = 0; gd < D; ++gd) {                              /* gridencoder.cu:201-243 */
            float rg[8] = {0};
            for (uint32_t idx = 0; idx < (1u << (D - 1)); ++idx) {
                float w = lv.scale;
                for (uint32_t nd = 0; nd < D - 1; ++nd) {
                    const uint32_t d = (nd >= gd) ? nd + 1 : nd;
                    if ((idx >> nd) & 1u) { w *= pos[d]; pl[d] = pg[d] + 1; } else { w *= 1.0f - pos[d]; pl[d] = pg[d]; }
                }
                pl[gd] = pg[gd];
                const size_t el = gbase + (size_t)orc_grid_index(D, gridtype, align_corners, lv.hashmap_size, lv.resolution, pl) * C;
                pl[gd] = pg[gd] + 1;
                const size_t er = gbase + (size_t)orc_grid_index(D, gridtype, align_corners, lv.hashmap_size, lv.resolution, pl) * C;
                for (uint32_t c = 0; c < C; ++c) {
                    if (is_half) {
                        const float diff = rh(h2f(((const uint16_t *)embeddings)[er + c]) - h2f(((const uint16_t *)embeddings)[el + c]));
                        rg[c] = rh(rg[c] + rh((w * diff) * pder[gd]));
                    } else {
                        rg[c] += w * (((const float *)embeddings)[er + c] - ((const float *)embeddings)[el + c]) * pder[gd];
                    }
                }
            }
            for (uint32_t c = 0; c < C; ++c) { if (is_half) ((uint16_t *)dy_dx)[dbase + gd * C + c] = f2h(rg[c]); else ((float *)dy_dx)[dbase + gd * C + c] = rg[c]; }
        }
    }
}

/* gridencoder.cu:248-340.  grad [L,B,C].  The reference scatters with (half2) atomics whose order is
 * arbitrary; the oracle accumulates every table entry in DOUBLE into grad_acc [n_entries*C] (caller-
 * zeroed) -- the exact sum the atomics approximate.  Each contribution is w*grad, narrowed to half first
 * when is_half (gridencoder.cu:329). */
ORC_API void orc_grid_encode_backward(const void *grad, const float *inputs, const int32_t *offsets, double *grad_acc,
                                      uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H,
                                      uint32_t gridtype, int align_corners, uint32_t interp, int is_half) {
    #pragma omp parallel for schedule(static)
    for (int64_t level = 0; level < (int64_t)L; ++level) {
        const orc_level_t lv = orc_level(offsets, (uint32_t)level, S, H);
        const size_t gbase = (size_t)(uint32_t)offsets[level] * C;
        for (uint32_t b = 0; b < B; ++b) {
            const float *x = inputs + (size_t)b * D;
            int oob = 0;
            for (uint32_t d = 0; d < D; ++d) if (x[d] < 0.0f || x[d] > 1.0f) oob = 1;
            if (oob) continue;
            float pos[ORC_MAXD], pder[ORC_MAXD]; uint32_t pg[ORC_MAXD], pl[ORC_MAXD];
            orc_locate(D, x, lv.scale, align_corners, interp, pos, pder, pg);
            const size_t gb = ((size_t)level * B + b) * C;
            for (uint32_t idx = 0; idx < (1u << D); ++idx) {
                float w = 1.0f;
                for (uint32_t d = 0; d < D; ++d) {
                    if ((idx >> d) & 1u) { w *= pos[d]; pl[d] = pg[d] + 1; } else { w *= 1.0f - pos[d]; pl[d] = pg[d]; }
                }
                const size_t e = gbase + (size_t)orc_grid_index(D, gridtype, align_corners, lv.hashmap_size, lv.resolution, pl) * C;
                for (uint32_t c = 0; c < C; ++c) {
                    if (is_half) grad_acc[e + c] += (double)rh(w * h2f(((const uint16_t *)grad)[gb + c]));
                    else         grad_acc[e + c] += (double)(w * ((const float *)grad)[gb + c]);
                }
            }
        }
    }
}

/* gridencoder.cu:343-369 : grad_inputs[b,d] = sum_{l,c} grad[l,b,c] * dy_dx[b,l,d,c] (scalar_t running sum) */
ORC_API void orc_grid_input_backward(const void *grad, const void *dy_dx, void *grad_inputs,
                                     uint32_t B, uint32_t D, uint32_t C, uint32_t L, int is_half) {
    #pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < (int64_t)B * D; ++t) {
        const uint32_t b = (uint32_t)(t / D), d = (uint32_t)(t % D);
        float acc = 0.0f;
        for (uint32_t l = 0; l < L; ++l) for (uint32_t c = 0; c < C; ++c) {
            const size_t gi = ((size_t)l * B + b) * C + c, di = (((size_t)b * L + l) * D + d) * C + c;
            if (is_half) acc = rh(acc + rh(h2f(((const uint16_t *)grad)[gi]) * h2f(((const uint16_t *)dy_dx)[di])));
            else         acc = fmaf(((const float *)grad)[gi], ((const float *)dy_dx)[di], acc);
        }
        if (is_half) ((uint16_t *)grad_inputs)[t] = f2h(acc); else ((float *)grad_inputs)[t] = acc;
    }
}

/* gridencoder.cu:506-610 : normalised total-variation gradient at the cells containing `inputs`
 * (fp32 only: grid.py:164 disables autocast).  Accumulates into grad_acc (double, caller-zeroed). */
ORC_API void orc_grad_total_variation(const float *inputs, const float *embeddings, double *grad_acc, const int32_t *offsets,
                                      float weight, uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H,
                                      uint32_t gridtype, int align_corners) {
    for (uint32_t level = 0; level < L; ++level) {
        const orc_level_t lv = orc_level(offsets, level, S, H);
        const size_t gbase = (size_t)(uint32_t)offsets[level] * C;
        for (uint32_t b = 0; b < B; ++b) {
            const float *x = inputs + (size_t)b * D;
            int oob = 0;
            for (uint32_t d = 0; d < D; ++d) if (x[d] < 0.0f || x[d] > 1.0f) oob = 1;
            if (oob) continue;
            uint32_t pg[ORC_MAXD];
            for (uint32_t d = 0; d < D; ++d) pg[d] = (uint32_t)floorf(fmaf(x[d], lv.scale, align_corners ? 0.0f : 0.5f));
            float res[8] = {0}, idelta[8] = {0};
            const size_t e0 = gbase + (size_t)orc_grid_index(D, gridtype, align_corners, lv.hashmap_size, lv.resolution, pg) * C;
            const float w = weight / (float)(2 * D);
            for (uint32_t d = 0; d < D; ++d) {
                const uint32_t cur = pg[d];
                if (cur < lv.resolution) {
                    pg[d] = cur + 1;
                    const size_t er = gbase + (size_t)orc_grid_index(D, gridtype, align_corners, lv.hashmap_size, lv.resolution, pg) * C;
                    for (uint32_t c = 0; c < C; ++c) { float g = embeddings[e0 + c] - embeddings[er + c]; res[c] += g; idelta[c] = fmaf(g, g, idelta[c]); }
                }
                if (cur > 0) {
                    pg[d] = cur - 1;
                    const size_t el = gbase + (size_t)orc_grid_index(D, gridtype, align_corners, lv.hashmap_size, lv.resolution, pg) * C;
                    for (uint32_t c = 0; c < C; ++c) { float g = embeddings[e0 + c] - embeddings[el + c]; res[c] += g; idelta[c] = fmaf(g, g, idelta[c]); }
                }
                pg[d] = cur;
            }
            for (uint32_t c = 0; c < C; ++c) grad_acc[e0 + c] += (double)((w * res[c]) * (1.0f / sqrtf(idelta[c] + 1e-9f)));
        }
    }
}
