"""CPU restatement of ONE training step of the hot path, chained from the oracle's C functions.  TEST / BENCH
INFRASTRUCTURE ONLY (bench.py's `cpu_baseline` leg and its `--impl reference` fallback); never imported by the product.

Follows the same call sequence as the reference's cuda_ray train step (legacy/nerf/renderer.py:268-319 +
legacy/nerf/utils.py:510-560, 863-868; network wiring nerf/network_ff.py:51-74): near/far -> march_rays_train -> hash grid
-> sigma FFMLP -> trunc_exp -> SH-4 + geo features -> colour FFMLP -> sigmoid -> composite_rays_train -> MSE ->
backward of all of it -> Adam(betas=(0.9,0.99), eps=1e-15).  The reference itself has no CPU path (every operator is a
CUDA extension), so this is a port ("kind": "port"), parallelised with OpenMP inside the C functions.
"""
import numpy as np

import oracle


class CpuState:
    def __init__(self, bound=1.0, seed=0, lr=1e-2):
        g = np.random.RandomState(seed)
        self.bound = bound
        self.offsets, self.pls = oracle.grid_offsets(desired_resolution=2048 * bound)
        n = int(self.offsets[-1])
        self.table = g.uniform(-1e-4, 1e-4, (n, 2)).astype(np.float32)
        std = np.sqrt(3 / 64)
        self.w_sigma = g.uniform(-std, std, oracle.ffmlp_num_params(32, 16, 64, 2)).astype(np.float32)
        self.w_color = g.uniform(-std, std, oracle.ffmlp_num_params(32, 16, 64, 3)).astype(np.float32)
        self.params = [self.table, self.w_sigma, self.w_color]
        self.m = [np.zeros_like(p) for p in self.params]
        self.v = [np.zeros_like(p) for p in self.params]
        self.t, self.lr = 0, lr

    def adam(self, grads):
        self.t += 1
        b1, b2, eps = 0.9, 0.99, 1e-15
        for p, m, v, g in zip(self.params, self.m, self.v, grads):
            g = g.astype(np.float32).reshape(p.shape)
            m *= b1; m += (1 - b1) * g
            v *= b2; v += (1 - b2) * g * g
            p -= self.lr * (m / (1 - b1 ** self.t)) / (np.sqrt(v / (1 - b2 ** self.t)) + eps)


def pad128(a):
    pad = 128 - a.shape[0] % 128
    return np.concatenate([a, np.zeros((pad,) + a.shape[1:], a.dtype)], 0)


def train_step(st, rays_o, rays_d, gt, bitfield, cascade, noises, bg=1.0, H=128, dt_gamma=0.0, max_steps=1024):
    """-> (loss, n_samples).  fp16 storage where the reference uses fp16 (table, MLP activations)."""
    b = st.bound
    aabb = np.array([-b, -b, -b, b, b, b], np.float32)
    N = rays_o.shape[0]
    nears, fars = oracle.near_far_from_aabb(rays_o, rays_d, aabb, 0.2)
    xyzs, dirs, deltas, rays, counter = oracle.march_rays_train(rays_o, rays_d, b, bitfield, cascade, H, nears, fars, noises,
                                                                N * max_steps, dt_gamma, max_steps)
    M = int(counter[0])
    M += 128 - M % 128
    xyzs, dirs, deltas = xyzs[:M], dirs[:M], deltas[:M]
    x01 = (xyzs + b) / (2 * b)
    enc_lbc, _ = oracle.grid_encode_forward(x01, st.table, st.offsets, st.pls, 16, False, half=True)
    enc = np.ascontiguousarray(enc_lbc.transpose(1, 0, 2)).reshape(M, 32)
    ws16, wc16 = st.w_sigma.astype(np.float16), st.w_color.astype(np.float16)
    enc_p = pad128(enc)
    h, fb_s = oracle.ffmlp_forward(enc_p, ws16, 32, 16, 64, 2)
    h0 = h[:M, 0].astype(np.float32)
    sigma = oracle.trunc_exp(h0)
    cin = np.concatenate([oracle.sh4_encode(dirs).astype(np.float16), h[:M, 1:16], np.zeros((M, 1), np.float16)], 1)
    cin_p = pad128(cin)
    hc, fb_c = oracle.ffmlp_forward(cin_p, wc16, 32, 16, 64, 3)
    hc3 = hc[:M, :3].astype(np.float32)
    rgb = 1.0 / (1.0 + np.exp(-hc3))
    wsum, depth, image = oracle.composite_rays_train_forward(sigma, rgb, deltas, rays)
    pred = image + (1 - wsum)[:, None] * bg
    diff = pred - gt
    loss = float((diff * diff).mean())
    # ---- backward
    g_pred = (2.0 / diff.size) * diff
    g_ws = -(g_pred * bg).sum(1)
    g_sig, g_rgb = oracle.composite_rays_train_backward(g_ws.astype(np.float32), g_pred.astype(np.float32), sigma, rgb, deltas, rays, wsum, image)
    g_hc = np.zeros((cin_p.shape[0], 16), np.float16)
    g_hc[:M, :3] = (g_rgb * rgb * (1 - rgb)).astype(np.float16)
    gw_c, _, gi_c = oracle.ffmlp_backward(g_hc, cin_p, wc16, fb_c, 32, 16, 64, 3, calc_grad_inputs=True)
    g_h = np.zeros((enc_p.shape[0], 16), np.float16)
    g_h[:M, 0] = oracle.trunc_exp_backward(g_sig, h0).astype(np.float16)
    g_h[:M, 1:16] = gi_c[:M, 16:31]
    gw_s, _, gi_s = oracle.ffmlp_backward(g_h, enc_p, ws16, fb_s, 32, 16, 64, 2, calc_grad_inputs=True)
    g_enc = np.ascontiguousarray(gi_s[:M].reshape(M, 16, 2).transpose(1, 0, 2))
    g_table = oracle.grid_encode_backward(g_enc, x01, st.offsets, st.pls, 16, 2, half=True)
    st.adam([g_table, gw_s, gw_c])
    return loss, M
