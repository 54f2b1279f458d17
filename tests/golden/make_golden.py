"""Generate tests/golden/*.npz from the REFERENCE's own CUDA extensions (oracle/_ref, built from /root/reference by
oracle/build_ref.py) on a B200.  The reference ships no golden vectors or tests, so these fixtures are
what pins the CPU oracle: `pytest -m "not gpu"` checks oracle/ against them on any host.

Inputs are NOT stored: they are regenerated from the seeds in tests/cases.py (GOLDEN_CASES below).  Run on the GPU box:
    python tests/golden/make_golden.py gpurun_out/golden        # then copy the .npz files into tests/golden/
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases  # noqa: E402
import ref_ext  # noqa: E402
import runners  # noqa: E402

GOLDEN_CASES = {
    "march_b1": dict(n=160, seed=31, bound=1.0, dt_gamma=0.0),
    "march_b2": dict(n=160, seed=32, bound=2.0, dt_gamma=1.0 / 128),
    "composite": dict(n=160, seed=33, bound=1.0),
    "grid": dict(B=256, seed=34),
    "mlp_sigma": dict(B=256, input_dim=32, output_dim=16, num_layers=2, seed=35),
    "mlp_color": dict(B=256, input_dim=32, output_dim=16, num_layers=3, seed=36),
    "freq": dict(B=512, seed=37, deg=4),
}


def sparse(a):
    flat = a.reshape(-1)
    idx = np.nonzero(flat)[0].astype(np.int64)
    return idx, flat[idx]


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    rm, ge, ff, fq = (ref_ext.load(p) for p in ("raymarching", "gridencoder", "ffmlp", "freqencoder"))
    assert all(m is not None for m in (rm, ge, ff, fq)), "oracle/_ref extensions missing or no CUDA device"
    for name in ("march_b1", "march_b2"):
        k = GOLDEN_CASES[name]
        c = cases.march_case(k["n"], k["seed"], k["bound"], k["dt_gamma"])
        r = runners.march_train(rm, c)
        nf = runners.near_far(rm, c)
        N = c["rays_o"].shape[0]
        import torch
        # first iteration of the inference loop (n_step = 1, no noise)
        field = lambda x, d: (torch.ones(x.shape[0], device="cuda"), torch.ones(x.shape[0], 3, device="cuda"))
        it = runners.infer_loop(rm, c, field, max_iters=1)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), counts=r["counts"], counter=r["counter"], xyzs=r["xyzs"],
                            dirs=r["dirs"], deltas=r["deltas"], nears=nf[0], fars=nf[1], infer_xyzs=it["trace"][0]["xyzs"][:N],
                            infer_deltas=it["trace"][0]["deltas"][:N])
    k = GOLDEN_CASES["composite"]
    c = cases.composite_case(k["n"], k["seed"], k["bound"])
    np.savez_compressed(os.path.join(out_dir, "composite.npz"), **runners.composite_train(rm, c))
    k = GOLDEN_CASES["grid"]
    c = cases.grid_case(k["B"], k["seed"])
    from focnerf_b200.gridencoder.grid import level_scales
    dev_scales = level_scales(c["per_level_scale"], c["H"], 16).cpu().numpy()
    for half in (True, False):
        r = runners.grid_fwd_bwd(ge, c, half=half)
        gi, gv = sparse(r["grad_emb"])
        np.savez_compressed(os.path.join(out_dir, f"grid_{'f16' if half else 'f32'}.npz"), out=r["out"], dy_dx=r["dy_dx"],
                            grad_idx=gi, grad_val=gv, grad_inputs=r["grad_inputs"], device_level_scales=dev_scales)
    for name in ("mlp_sigma", "mlp_color"):
        k = GOLDEN_CASES[name]
        c = cases.mlp_case(**k)
        ff.allocate_splitk(k["num_layers"] + 1)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **runners.mlp_fwd_bwd(ff, c))
    k = GOLDEN_CASES["freq"]
    x = np.random.RandomState(k["seed"]).uniform(-1, 1, (k["B"], 3)).astype(np.float32)
    out, gi, g = runners.freq(fq, x, k["deg"])
    np.savez_compressed(os.path.join(out_dir, "freq.npz"), out=out, grad_inputs=gi, grad=g)
    print("golden vectors written to", out_dir, sorted(os.listdir(out_dir)))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join("gpurun_out", "golden"))
