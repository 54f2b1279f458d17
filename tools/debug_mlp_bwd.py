"""Per-tile error breakdown of the fused MLP backward vs the oracle (diagnostic; run on the GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import cases, oracle, runners
from focnerf_b200.ffmlp.ffmlp import _backend as ff

B, IN, L = int(sys.argv[1]) if len(sys.argv) > 1 else 128 * 700, 32, int(sys.argv[2]) if len(sys.argv) > 2 else 3
c = cases.mlp_case(B, IN, 16, L)
x, w, g = runners.T(c["inputs"]), runners.T(c["weights"]), runners.T(c["grad"])
fb = torch.empty(L, B, 64, dtype=torch.float16, device="cuda"); out = torch.empty(B, 16, dtype=torch.float16, device="cuda")
ff.ffmlp_forward(x, w, B, IN, 16, 64, L, 0, 6, fb, out)
o_out, o_fb = oracle.ffmlp_forward(c["inputs"], c["weights"], IN, 16, 64, L)
print("fwd rel", np.abs(out.float().cpu().numpy() - o_out.astype(np.float32)).max() / np.abs(o_out.astype(np.float32)).max())
fbn = fb.cpu().numpy()
gw_o, bb_o, gi_o = oracle.ffmlp_backward(c["grad"], c["inputs"], c["weights"], fbn, IN, 16, 64, L, calc_grad_inputs=True)
for trial in range(2):
    bb = torch.zeros(L, B, 64, dtype=torch.float16, device="cuda"); gi = torch.zeros(B, IN, dtype=torch.float16, device="cuda"); gw = torch.zeros_like(w)
    ff.ffmlp_backward(g, x, w, fb, B, IN, 16, 64, L, 0, 6, True, bb, gi, gw)
    torch.cuda.synchronize()
    bbn, gin, gwn = bb.float().cpu().numpy(), gi.float().cpu().numpy(), gw.float().cpu().numpy()
    grid = min(B // 128, 296)
    sc_gi = np.abs(gi_o.astype(np.float32)).max()
    e_gi = np.abs(gin - gi_o.astype(np.float32)).reshape(B // 128, -1).max(1) / sc_gi
    bad = np.where(e_gi > 2e-3)[0]
    print(f"trial {trial}: grad_inputs bad tiles {len(bad)}/{B // 128}; first {bad[:12]}; cta {bad[:12] % grid}; iter {bad[:12] // grid}; max {e_gi.max():.3e}")
    for k in range(L):
        sc = np.abs(bb_o[k].astype(np.float32)).max()
        e = np.abs(bbn[k] - bb_o[k].astype(np.float32)).reshape(B // 128, -1).max(1) / sc
        b2 = np.where(e > 2e-3)[0]
        print(f"   backward_buffer[{k}] bad tiles {len(b2)}; first {b2[:12]}; iter {b2[:12] // grid}; max {e.max():.3e}")
    print("   grad_weights rel", np.abs(gwn - gw_o).max() / np.abs(gw_o).max())
    if len(bad):
        t = bad[0]; rows = np.abs(gin - gi_o.astype(np.float32))[t * 128:(t + 1) * 128].max(1) / sc_gi
        print("   rows of first bad tile over tol:", np.where(rows > 2e-3)[0][:40], " n=", (rows > 2e-3).sum())
        cols = np.abs(gin - gi_o.astype(np.float32))[t * 128:(t + 1) * 128].max(0) / sc_gi
        print("   cols over tol:", np.where(cols > 2e-3)[0])
