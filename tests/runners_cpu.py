"""Oracle-side counterparts of tests/runners.py (numpy, CPU)."""
import oracle
from runners import canonical_march


def march_train(c, M=None):
    N = c["rays_o"].shape[0]
    M = N * c["max_steps"] if M is None else M
    x, d, dl, rays, counter = oracle.march_rays_train(c["rays_o"], c["rays_d"], c["bound"], c["bitfield"], c["cascade"], c["H"],
                                                      c["nears"], c["fars"], c["noises"], M, c["dt_gamma"], c["max_steps"])
    out = canonical_march(rays, x, d, dl, M)
    out["counter"] = counter
    return out
