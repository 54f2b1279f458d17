/*
 * oracle/orc_mlp.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * CPU restatement of the fully fused MLP (ffmlp/src/ffmlp.cu), the frequency encoder
 * (freqencoder/src/freqencoder.cu), trunc_exp (activation.py) and a degree-4 real spherical-harmonics
 * direction encoder.  The SH encoder's source (torch-ngp `shencoder`) is ABSENT from the reference
 * tree (nerf/network_ff.py:5 imports it) -> parity unpinned for that one op; it is the
 * public closed form also used by tiny-cuda-nn's SphericalHarmonics encoding.
 *
 * MLP numerics: storage between layers is binary16 exactly as in the reference (activations are written
 * back to shared memory as __half, ffmlp.cu:118).  The reference accumulates each 16x16x16 wmma in
 * binary16 (OUT_T=__half, ffmlp.cu:564); tensor-core internal rounding is not specified, so the oracle
 * offers two models: acc_mode 0 = exact (double) dot product rounded once to half -- what the tcgen05
 * kernel (fp32 TMEM accumulators) must match tightly; acc_mode 1 = round the running sum to half after
 * every 16-wide K block -- an approximation of the reference used to bound the fp16-accumulate gap.
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <stdlib.h>

#define ORC_API __attribute__((visibility("default")))
typedef _Float16 orc_half;
static inline float h2f(uint16_t h) { orc_half v; memcpy(&v, &h, 2); return (float)v; }
static inline uint16_t f2h(float f) { orc_half v = (orc_half)f; uint16_t h; memcpy(&h, &v, 2); return h; }
static inline float rh(float f) { return (float)(orc_half)f; }

/* y[n] = sum_k x[k] * W[n*K + k], W row-major [Nout, K] (ffmlp.cu:632) */
static inline float orc_dot(const uint16_t *x, const uint16_t *w, uint32_t K, int acc_mode) {
    if (acc_mode == 0) {
        double s = 0.0;
        for (uint32_t k = 0; k < K; ++k) s += (double)h2f(x[k]) * (double)h2f(w[k]);
        return (float)s;
    }
    float acc = 0.0f;
    for (uint32_t k0 = 0; k0 < K; k0 += 16) {
        float s = acc;
        for (uint32_t k = k0; k < k0 + 16 && k < K; ++k) s += h2f(x[k]) * h2f(w[k]);
        acc = rh(s);
    }
    return acc;
}

/* utils.h:424-470 (forward activations); 0 relu, 1 exp, 2 sine, 3 sigmoid, 4 squareplus, 5 softplus, else none */
static inline float orc_act(uint32_t a, float v) {
    const float K_ACT = 10.0f;
    switch (a) {
        case 0: return v > 0.0f ? v : 0.0f;
        case 1: return expf(v);
        case 2: return sinf(v);
        case 3: return 1.0f / (1.0f + expf(-v));
        case 4: { float x = v * K_ACT; return 0.5f * (x + sqrtf(x * x + 4.0f)) / K_ACT; }
        case 5: return logf(expf(v * K_ACT) + 1.0f) / K_ACT;
        default: return v;
    }
}
/* utils.h:537-582 : derivative expressed through the stored POST-activation value */
static inline float orc_act_bwd(uint32_t a, float g, float fwd) {
    const float K_ACT = 10.0f;
    switch (a) {
        case 0: return fwd > 0.0f ? g : 0.0f;
        case 1: return g * fwd;
        case 3: return g * rh(fwd * rh(1.0f - fwd));
        case 4: { float y = fwd * K_ACT; return g * rh(y * y / (y * y + 1.0f)); }
        case 5: return g * rh(1.0f - expf(-fwd * K_ACT));
        default: return g;
    }
}

/* ffmlp.cu:331-407, 525-577, 635-671.  inputs [B,in] half, weights flat half
 * ([hidden,in] + (num_layers-1)*[hidden,hidden] + [out_pad16,hidden]), forward_buffer [num_layers,B,hidden]
 * half (may be NULL = inference), outputs [B,output_dim] half with output_dim <= 16 rows of W_out used. */
ORC_API void orc_ffmlp_forward(const uint16_t *inputs, const uint16_t *weights, uint32_t B, uint32_t input_dim,
                               uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers, uint32_t activation,
                               uint32_t output_activation, uint16_t *forward_buffer, uint16_t *outputs, int acc_mode) {
    #pragma omp parallel
    {
        uint16_t *cur = (uint16_t *)malloc(2 * hidden_dim), *nxt = (uint16_t *)malloc(2 * hidden_dim);
        #pragma omp for schedule(static)
        for (int64_t b = 0; b < (int64_t)B; ++b) {
            const uint16_t *w = weights;
            for (uint32_t n = 0; n < hidden_dim; ++n)
                cur[n] = f2h(orc_act(activation, rh(orc_dot(inputs + (size_t)b * input_dim, w + (size_t)n * input_dim, input_dim, acc_mode))));
            if (forward_buffer) memcpy(forward_buffer + (size_t)b * hidden_dim, cur, 2 * hidden_dim);
            w += (size_t)hidden_dim * input_dim;
            for (uint32_t k = 1; k < num_layers; ++k) {
                for (uint32_t n = 0; n < hidden_dim; ++n)
                    nxt[n] = f2h(orc_act(activation, rh(orc_dot(cur, w + (size_t)n * hidden_dim, hidden_dim, acc_mode))));
                uint16_t *t = cur; cur = nxt; nxt = t;
                if (forward_buffer) memcpy(forward_buffer + ((size_t)k * B + (size_t)b) * hidden_dim, cur, 2 * hidden_dim);
                w += (size_t)hidden_dim * hidden_dim;
            }
            for (uint32_t n = 0; n < output_dim; ++n)
                outputs[(size_t)b * output_dim + n] = f2h(orc_act(output_activation, rh(orc_dot(cur, w + (size_t)n * hidden_dim, hidden_dim, acc_mode))));
        }
        free(cur); free(nxt);
    }
}

/* ffmlp.cu:410-518 (dL/dact chain), 749-895 (weight gradients, optional dL/dinput).
 * grad [B,output_dim(16)] half.  backward_buffer [num_layers,B,hidden] half (slot 0 = gradient at the LAST
 * hidden layer's pre-activation, slot k = k layers earlier).  grad_weights: DOUBLE, flat, same layout as
 * weights, caller-zeroed (the exact sums the reference's fp16 split-K GEMMs approximate).
 * grad_inputs [B,in] half or NULL. */
ORC_API void orc_ffmlp_backward(const uint16_t *grad, const uint16_t *inputs, const uint16_t *weights,
                                const uint16_t *forward_buffer, uint32_t B, uint32_t input_dim, uint32_t output_dim,
                                uint32_t hidden_dim, uint32_t num_layers, uint32_t activation,
                                uint16_t *backward_buffer, uint16_t *grad_inputs, double *grad_weights, int acc_mode) {
    const size_t in_sz = (size_t)hidden_dim * input_dim, hh = (size_t)hidden_dim * hidden_dim;
    const uint16_t *W_out = weights + in_sz + (size_t)(num_layers - 1) * hh;
    /* pass 1: activation-gradient chain, row by row */
    #pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < (int64_t)B; ++b) {
        const uint16_t *g = grad + (size_t)b * output_dim;
        uint16_t *bb0 = backward_buffer + (size_t)b * hidden_dim;
        const uint16_t *f_last = forward_buffer + ((size_t)(num_layers - 1) * B + (size_t)b) * hidden_dim;
        for (uint32_t j = 0; j < hidden_dim; ++j) {
            double s = 0.0;
            for (uint32_t o = 0; o < output_dim; ++o) s += (double)h2f(g[o]) * (double)h2f(W_out[(size_t)o * hidden_dim + j]);
            bb0[j] = f2h(orc_act_bwd(activation, rh((float)s), h2f(f_last[j])));
        }
        for (uint32_t k = 1; k < num_layers; ++k) {
            const uint32_t m = num_layers - 1 - k;                 /* W_h[m]: fwd[m] -> fwd[m+1] */
            const uint16_t *W = weights + in_sz + (size_t)m * hh;
            const uint16_t *prev = backward_buffer + ((size_t)(k - 1) * B + (size_t)b) * hidden_dim;
            uint16_t *out = backward_buffer + ((size_t)k * B + (size_t)b) * hidden_dim;
            const uint16_t *f = forward_buffer + ((size_t)m * B + (size_t)b) * hidden_dim;
            for (uint32_t j = 0; j < hidden_dim; ++j) {
                double s = 0.0;
                for (uint32_t o = 0; o < hidden_dim; ++o) s += (double)h2f(prev[o]) * (double)h2f(W[(size_t)o * hidden_dim + j]);
                out[j] = f2h(orc_act_bwd(activation, rh((float)s), h2f(f[j])));
            }
        }
        if (grad_inputs) {
            const uint16_t *last = backward_buffer + ((size_t)(num_layers - 1) * B + (size_t)b) * hidden_dim;
            for (uint32_t i = 0; i < input_dim; ++i) {
                double s = 0.0;
                for (uint32_t o = 0; o < hidden_dim; ++o) s += (double)h2f(last[o]) * (double)h2f(weights[(size_t)o * input_dim + i]);
                grad_inputs[(size_t)b * input_dim + i] = f2h((float)s);
            }
        }
    }
    (void)acc_mode;
    /* pass 2: weight gradients dW = dPre^T * Act_prev (exact double sums).  Row-major walk over the batch: each thread
     * accumulates outer products of contiguous rows into a private copy, reduced at the end. */
    const size_t n_params = in_sz + (size_t)(num_layers - 1) * hh + (size_t)output_dim * hidden_dim;
    #pragma omp parallel
    {
        double *acc = (double *)calloc(n_params, sizeof(double));
        float *dv = (float *)malloc(sizeof(float) * hidden_dim), *av = (float *)malloc(sizeof(float) * (hidden_dim > input_dim ? hidden_dim : input_dim));
        #pragma omp for schedule(static)
        for (int64_t b = 0; b < (int64_t)B; ++b) {
            /* output layer: dW_out[o][j] += grad[b][o] * A_{L-1}[b][j] */
            const uint16_t *fl = forward_buffer + ((size_t)(num_layers - 1) * B + (size_t)b) * hidden_dim;
            for (uint32_t j = 0; j < hidden_dim; ++j) av[j] = h2f(fl[j]);
            for (uint32_t o = 0; o < output_dim; ++o) {
                const double gv = (double)h2f(grad[(size_t)b * output_dim + o]);
                double *row = acc + in_sz + (size_t)(num_layers - 1) * hh + (size_t)o * hidden_dim;
                if (gv != 0.0) for (uint32_t j = 0; j < hidden_dim; ++j) row[j] += gv * (double)av[j];
            }
            /* hidden matrices m: dW[o][j] += dZ_{m+1}[b][o] * A_m[b][j];  dZ_{m+1} is backward_buffer slot num_layers-2-m */
            for (uint32_t m = 0; m + 1 < num_layers; ++m) {
                const uint16_t *d = backward_buffer + ((size_t)(num_layers - 2 - m) * B + (size_t)b) * hidden_dim;
                const uint16_t *f = forward_buffer + ((size_t)m * B + (size_t)b) * hidden_dim;
                for (uint32_t j = 0; j < hidden_dim; ++j) av[j] = h2f(f[j]);
                for (uint32_t o = 0; o < hidden_dim; ++o) {
                    const double gv = (double)h2f(d[o]);
                    double *row = acc + in_sz + (size_t)m * hh + (size_t)o * hidden_dim;
                    if (gv != 0.0) for (uint32_t j = 0; j < hidden_dim; ++j) row[j] += gv * (double)av[j];
                }
            }
            /* input matrix: dW_in[o][i] += dZ_0[b][o] * x[b][i] */
            const uint16_t *d0 = backward_buffer + ((size_t)(num_layers - 1) * B + (size_t)b) * hidden_dim;
            for (uint32_t i = 0; i < input_dim; ++i) av[i] = h2f(inputs[(size_t)b * input_dim + i]);
            for (uint32_t o = 0; o < hidden_dim; ++o) {
                const double gv = (double)h2f(d0[o]);
                double *row = acc + (size_t)o * input_dim;
                if (gv != 0.0) for (uint32_t i = 0; i < input_dim; ++i) row[i] += gv * (double)av[i];
            }
        }
        #pragma omp critical
        for (size_t k = 0; k < n_params; ++k) grad_weights[k] += acc[k];
        free(acc); free(dv); free(av);
    }
}

/* freqencoder.cu:30-58 : out = [x, sin(2^f x), cos(2^f x) as sin(.+pi/2)]_f ; C = D + 2*D*deg */
ORC_API void orc_freq_encode_forward(const float *inputs, uint32_t B, uint32_t D, uint32_t deg, uint32_t C, float *outputs) {
    const float HALF_PI = 3.141592653589793f / 2;
    #pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < (int64_t)B * C; ++t) {
        const uint32_t b = (uint32_t)(t / C), c = (uint32_t)(t % C);
        if (c < D) { outputs[t] = inputs[(size_t)b * D + c]; continue; }
        const uint32_t col = c / D - 1, d = c % D, f = col / 2;
        outputs[t] = sinf(scalbnf(inputs[(size_t)b * D + d], (int)f) + (float)(col % 2) * HALF_PI);
    }
}
/* freqencoder.cu:63-94 */
ORC_API void orc_freq_encode_backward(const float *grad, const float *outputs, uint32_t B, uint32_t D, uint32_t deg, uint32_t C, float *grad_inputs) {
    #pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < (int64_t)B * D; ++t) {
        const uint32_t b = (uint32_t)(t / D), d = (uint32_t)(t % D);
        const float *g = grad + (size_t)b * C, *o = outputs + (size_t)b * C;
        float r = g[d]; g += D; o += D;
        for (uint32_t f = 0; f < deg; ++f) { r += scalbnf(1.0f, (int)f) * (g[d] * o[D + d] - g[D + d] * o[d]); g += 2 * D; o += 2 * D; }
        grad_inputs[t] = r;
    }
}

/* Real spherical harmonics up to degree 4 (16 coefficients) of a unit direction; the closed form used by
 * torch-ngp's shencoder / tiny-cuda-nn's SphericalHarmonics (source absent from the reference). */
ORC_API void orc_sh4_encode(const float *dirs, uint32_t B, float *out) {
    #pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < (int64_t)B; ++b) {
        const float x = dirs[3 * b], y = dirs[3 * b + 1], z = dirs[3 * b + 2];
        const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
        float *o = out + 16 * (size_t)b;
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
}
