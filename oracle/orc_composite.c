/*
 * oracle/orc_composite.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * CPU restatement of the front-to-back compositors of the reference (raymarching.cu) and of FOC-NeRF's
 * multi-object composition (COMBINED.py / editable.py).  The reference evaluates exp() with the GPU's
 * approximate ex2 unit (__expf); here it is expf(), so these results are compared under the 1e-3
 * relative tolerance BASELINE.json states, never bitwise.
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>

#define ORC_API __attribute__((visibility("default")))

/* raymarching.cu:501-577.  rays rows are (ray id, offset, count); outputs are indexed by ray id.
 * The sample that drives T below T_thresh is still accumulated; the loop stops after it. */
ORC_API void orc_composite_rays_train_forward(const float *sigmas, const float *rgbs, const float *deltas,
                                              const int32_t *rays, uint32_t M, uint32_t N, float T_thresh,
                                              float *weights_sum, float *depth, float *image) {
    #pragma omp parallel for schedule(dynamic, 64)
    for (int64_t n = 0; n < (int64_t)N; ++n) {
        const uint32_t id = (uint32_t)rays[3 * n], off = (uint32_t)rays[3 * n + 1], cnt = (uint32_t)rays[3 * n + 2];
        float T = 1.0f, r = 0, g = 0, b = 0, ws = 0, t = 0, d = 0;
        if (cnt != 0 && off + cnt <= M) {
            for (uint32_t s = 0; s < cnt; ++s) {
                const size_t i = (size_t)off + s;
                const float alpha = 1.0f - expf(-sigmas[i] * deltas[2 * i]);
                const float w = alpha * T;
                r = fmaf(w, rgbs[3 * i], r); g = fmaf(w, rgbs[3 * i + 1], g); b = fmaf(w, rgbs[3 * i + 2], b);
                t += deltas[2 * i + 1];
                d = fmaf(w, t, d);
                ws += w;
                T *= 1.0f - alpha;
                if (T < T_thresh) break;
            }
        }
        weights_sum[id] = ws; depth[id] = d;
        image[3 * (size_t)id] = r; image[3 * (size_t)id + 1] = g; image[3 * (size_t)id + 2] = b;
    }
}

/* raymarching.cu:602-682.  grad_depth is ignored by the reference (raymarching.py:275).
 * grad buffers are expected pre-zeroed by the caller (raymarching.py:280-281). */
ORC_API void orc_composite_rays_train_backward(const float *grad_weights_sum, const float *grad_image,
                                               const float *sigmas, const float *rgbs, const float *deltas,
                                               const int32_t *rays, const float *weights_sum, const float *image,
                                               uint32_t M, uint32_t N, float T_thresh,
                                               float *grad_sigmas, float *grad_rgbs) {
    #pragma omp parallel for schedule(dynamic, 64)
    for (int64_t n = 0; n < (int64_t)N; ++n) {
        const uint32_t id = (uint32_t)rays[3 * n], off = (uint32_t)rays[3 * n + 1], cnt = (uint32_t)rays[3 * n + 2];
        if (cnt == 0 || off + cnt > M) continue;
        const float gw = grad_weights_sum[id];
        const float gr = grad_image[3 * (size_t)id], gg = grad_image[3 * (size_t)id + 1], gb = grad_image[3 * (size_t)id + 2];
        const float rf = image[3 * (size_t)id], gf = image[3 * (size_t)id + 1], bf = image[3 * (size_t)id + 2], wsf = weights_sum[id];
        float T = 1.0f, r = 0, g = 0, b = 0;
        for (uint32_t s = 0; s < cnt; ++s) {
            const size_t i = (size_t)off + s;
            const float alpha = 1.0f - expf(-sigmas[i] * deltas[2 * i]);
            const float w = alpha * T;
            r = fmaf(w, rgbs[3 * i], r); g = fmaf(w, rgbs[3 * i + 1], g); b = fmaf(w, rgbs[3 * i + 2], b);
            T *= 1.0f - alpha;     /* T is now the transmittance AFTER this sample */
            grad_rgbs[3 * i] = gr * w; grad_rgbs[3 * i + 1] = gg * w; grad_rgbs[3 * i + 2] = gb * w;
            grad_sigmas[i] = deltas[2 * i] * (gr * (T * rgbs[3 * i] - (rf - r)) +
                                              gg * (T * rgbs[3 * i + 1] - (gf - g)) +
                                              gb * (T * rgbs[3 * i + 2] - (bf - b)) +
                                              gw * (1.0f - wsf));
            if (T < T_thresh) break;
        }
    }
}

/* raymarching.cu:819-905.  In place; T = 1 - sum(w) here (not the product form of the train path);
 * delta == 0 marks the end of a ray's samples; a ray that stops before n_step samples is killed
 * (rays_alive[n] = -1), otherwise its rays_t is advanced. */
ORC_API void orc_composite_rays(uint32_t n_alive, uint32_t n_step, float T_thresh, int32_t *rays_alive, float *rays_t,
                                const float *sigmas, const float *rgbs, const float *deltas,
                                float *weights_sum, float *depth, float *image) {
    #pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < (int64_t)n_alive; ++n) {
        const int32_t id = rays_alive[n];
        float t = rays_t[id], ws = weights_sum[id], d = depth[id];
        float r = image[3 * (size_t)id], g = image[3 * (size_t)id + 1], b = image[3 * (size_t)id + 2];
        uint32_t s = 0;
        for (; s < n_step; ++s) {
            const size_t i = (size_t)n * n_step + s;
            if (deltas[2 * i] == 0.0f) break;
            const float alpha = 1.0f - expf(-sigmas[i] * deltas[2 * i]);
            const float T = 1.0f - ws;
            const float w = alpha * T;
            ws += w;
            t += deltas[2 * i + 1];
            d = fmaf(w, t, d);
            r = fmaf(w, rgbs[3 * i], r); g = fmaf(w, rgbs[3 * i + 1], g); b = fmaf(w, rgbs[3 * i + 2], b);
            if (T < T_thresh) break;
        }
        if (s < n_step) rays_alive[n] = -1; else rays_t[id] = t;
        weights_sum[id] = ws; depth[id] = d;
        image[3 * (size_t)id] = r; image[3 * (size_t)id + 1] = g; image[3 * (size_t)id + 2] = b;
    }
}

static inline float orc_linspace01(uint32_t i, uint32_t T) {
    if (T <= 1) return 0.0f;
    const float step = 1.0f / (float)(T - 1);
    return (i < T / 2) ? step * (float)i : 1.0f - step * (float)(T - 1 - i);
}

/*
 * FOC-NeRF multi-object composite (COMBINED.py:141-200, 247-251; editable.py:443-471 only changes the
 * rays the per-object samples were taken along, not this math).
 *   sigmas [K,N,T], rgbs [K,N,T,3]: per-object densities / colours at T uniform samples of N rays.
 *   Selection (COMBINED.py:247-251): running max with strict '>' so the EARLIER object wins ties.
 *   Composite (COMBINED.py:160-177): z_i = near + (far-near) * i/(T-1); delta_i = z_{i+1}-z_i, last
 *   delta = (far-near)/T (sample_dist); alpha = 1-exp(-delta*sigma); Ti = cumprod(1-alpha+1e-15) shifted;
 *   w = alpha*Ti; rgb = sum w c; 4th channel = sum w*sigma (sic); depth = sum w*clamp((z-near)/(far-near),0,1);
 *   out4 = (rgb, a) + (1 - sum w)*bg; clamp to [0,1].
 * out_image [N,4], out_depth [N], out_wsum [N].
 */
ORC_API void orc_foc_composite(const float *sigmas, const float *rgbs, const float *nears, const float *fars,
                               uint32_t K, uint32_t N, uint32_t T, float bg,
                               float *out_image, float *out_depth, float *out_wsum) {
    #pragma omp parallel for schedule(static)
    for (int64_t n = 0; n < (int64_t)N; ++n) {
        const float near = nears[n], far = fars[n];
        const float span = far - near, sample_dist = span / (float)T;
        float Tr = 1.0f, acc[4] = {0, 0, 0, 0}, dsum = 0, ws = 0;
        for (uint32_t i = 0; i < T; ++i) {
            float best = 0.0f, c[3] = {0, 0, 0};
            for (uint32_t k = 0; k < K; ++k) {
                const size_t j = ((size_t)k * N + (size_t)n) * T + i;
                const float s = sigmas[j];
                if (k == 0 || s > best) { best = s; c[0] = rgbs[3 * j]; c[1] = rgbs[3 * j + 1]; c[2] = rgbs[3 * j + 2]; }
            }
            /* torch.linspace(0,1,T) in fp32 (symmetric fill: start+step*i below the midpoint,
             * end-step*(T-1-i) above it), then near + span*u as two separate roundings (COMBINED.py:151-153) */
            const float u0 = orc_linspace01(i, T), u1 = orc_linspace01(i + 1, T);
            const float z0 = near + span * u0;
            const float delta = (i + 1 < T) ? (near + span * u1) - z0 : sample_dist;
            const float alpha = 1.0f - expf(-delta * best);
            const float w = alpha * Tr;
            acc[0] += w * c[0]; acc[1] += w * c[1]; acc[2] += w * c[2]; acc[3] += w * best;
            float od = (z0 - near) / span; od = fminf(1.0f, fmaxf(0.0f, od));
            dsum += w * od; ws += w;
            Tr *= (1.0f - alpha + 1e-15f);
        }
        for (int ch = 0; ch < 4; ++ch) {
            float v = acc[ch] + (1.0f - ws) * bg;
            out_image[4 * (size_t)n + ch] = fminf(1.0f, fmaxf(0.0f, v));
        }
        out_depth[n] = dsum; out_wsum[n] = ws;
    }
}
