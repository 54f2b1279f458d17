"""-m gpu: the reference's FOC multi-object callers import and run on this package (SURVEY.md 8 row a19 / 8c "tinycudann shim").

nerf/network_tcnn.py and nerf/multiobjectnetwork.py (verbatim copies under oracle/_ref/py/foc, oracle/build_ref.py) `import tinycudann
as tcnn`; focnerf_b200.install_aliases() registers focnerf_b200.tinycudann under that name.  A child process (the `nerf` package name
collides with the legacy callers other tests import) builds the reference's own NeRFNetwork -- HashGrid encoding, 32-64-16 sigma net,
SH-4 direction encoding, 47(->48)-64-64-3 colour net and the 144-16-16 feature encoder, i.e. both the tcgen05 and the generic FFMLP
kernels -- evaluates density() / color() / forward(), and checks every stage against independent restatements: the C oracle for the
hash grid and SH-4, plain fp32 torch matmuls on the `.params` blobs for the MLPs.  MONeRFNetwork is then built over two checkpoints
written by the same class and its max-density selection runs."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import json, os, sys, tempfile, warnings
warnings.filterwarnings("ignore")
ROOT = sys.argv[1]
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import focnerf_b200
focnerf_b200.install_aliases()
sys.path.insert(0, os.path.join(ROOT, "oracle", "stubs"))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref", "py", "foc"))
import oracle
from nerf.network_tcnn import NeRFNetwork
from nerf.multiobjectnetwork import MONeRFNetwork
import nerf.network_tcnn as nt, tinycudann
res = {"files": [os.path.realpath(nt.__file__).endswith("oracle/_ref/py/foc/nerf/network_tcnn.py"), tinycudann is focnerf_b200.tinycudann]}
torch.manual_seed(0)
m = NeRFNetwork(bound=1, cuda_ray=True).cuda().eval()
with torch.no_grad():
    m.encoder.params.copy_(torch.randn_like(m.encoder.params) * 0.2)
N = 1000
g = torch.Generator().manual_seed(3)
x = (torch.rand(N, 3, generator=g) * 2 - 1).cuda()
d = torch.nn.functional.normalize(torch.randn(N, 3, generator=g), dim=-1).cuda()
raw = np.random.RandomState(5).rand(144).astype(np.float32)
yolo = (None, None, raw)
rel = lambda a, b: float((a.double() - b.double()).abs().max() / (b.double().abs().max() + 1e-20))

def mlp_ref(net, inp):
    """fp32 matmuls on the flat blob, fp16 storage between layers, ones in the padded input columns"""
    w = net.params.detach().half().float()
    h = torch.cat([inp.half().float(), torch.ones(inp.shape[0], net.in_pad - net.n_input_dims, device=inp.device)], 1)
    off, dims = 0, [(net.hidden, net.in_pad)] + [(net.hidden, net.hidden)] * (net.layers - 1)
    for o, i in dims:
        h = torch.relu(h @ w[off:off + o * i].view(o, i).t()).half().float(); off += o * i
    return (h @ w[off:off + 16 * net.hidden].view(16, net.hidden).t())[:, :net.n_output_dims]

with torch.no_grad():
    x01 = (x + 1) / 2
    enc = m.encoder(x01)
    e = m.encoder
    want = oracle.grid_encode_forward(x01.cpu().numpy(), e.params.detach().view(-1, 2).half().cpu().numpy(), e.offsets.cpu().numpy(),
                                      e._grid["per_level_scale"], 16, half=True)
    want = want[0] if isinstance(want, tuple) else want
    want = torch.from_numpy(np.asarray(want, np.float32)).cuda()
    if want.dim() == 3:
        want = want.permute(1, 0, 2).reshape(N, -1)
    res["encoding_vs_oracle"] = rel(enc.float(), want)
    res["sh_vs_oracle"] = rel(m.encoder_dir((d + 1) / 2).float(), torch.from_numpy(oracle.sh4_encode(d.cpu().numpy())).cuda())
    h = m.sigma_net(enc)
    res["sigma_net_vs_torch"] = rel(h.float(), mlp_ref(m.sigma_net, enc))
    dens = m.density(x)
    res["density_shapes"] = [list(dens["sigma"].shape), list(dens["geo_feat"].shape)]
    res["sigma_is_trunc_exp"] = rel(dens["sigma"].float(), torch.exp(h[:, 0].float()))
    rgb = m.color(x, d, yolo, geo_feat=dens["geo_feat"])
    obj = m.yolo_feat_encoder(torch.tensor(raw, device="cuda").unsqueeze(0))
    res["feat_encoder_vs_torch"] = rel(obj.float(), mlp_ref(m.yolo_feat_encoder, torch.tensor(raw, device="cuda").unsqueeze(0)))
    cin = torch.cat([m.encoder_dir((d + 1) / 2), dens["geo_feat"], obj.squeeze(0).repeat(N, 1)], -1)
    res["color_vs_torch"] = rel(rgb.float(), torch.sigmoid(mlp_ref(m.color_net, cin)))
    res["rgb_range"] = [float(rgb.min()), float(rgb.max())]
    mask = torch.rand(N, device="cuda") > 0.5
    rgb_m = m.color(x, d, yolo, mask=mask, geo_feat=dens["geo_feat"])
    res["masked_color"] = bool(torch.equal(rgb_m[~mask], torch.zeros_like(rgb_m[~mask]))) and rel(rgb_m[mask].float(), rgb[mask].float()) < 2e-3
# gradients flow to every parameter group through the shim (the reference trains these modules with Adam)
m.train()
with torch.autocast("cuda", dtype=torch.float16):
    s2, c2 = m(x, d, (None, None, m.yolo_feat_encoder(torch.tensor(raw, device="cuda").unsqueeze(0)).squeeze(0).repeat(N, 1).detach()))
(s2.float().mean() + c2.float().mean()).backward()
res["grads"] = {k: bool(p.grad is not None and torch.isfinite(p.grad).all() and float(p.grad.abs().max()) > 0)
                for k, p in (("encoder", m.encoder.params), ("sigma_net", m.sigma_net.params), ("color_net", m.color_net.params))}
# MONeRFNetwork over two checkpoints of that class
tmp = tempfile.mkdtemp()
ck = []
for i in range(2):
    p = os.path.join(tmp, f"obj{i}.pth"); torch.save({"model": m.state_dict()}, p); ck.append(p)
mo = MONeRFNetwork(ck, NeRFNetwork, fp16=True, nw_kwargs=dict(bound=1, cuda_ray=True))
mo.to("cuda")
out = mo.density(x)
res["mo_density_shapes"] = [list(out["sigma"].shape), list(out["geo_feat"].shape)]
res["mo_finite"] = bool(torch.isfinite(out["sigma"]).all())
print("RESULT " + json.dumps(res))
'''


def test_foc_tcnn_callers_import_and_run_on_this_package():
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "py", "foc", "nerf", "network_tcnn.py")):
        pytest.skip("oracle/_ref/py not present (built by oracle/build_ref.py in the build container)")
    p = subprocess.run([sys.executable, "-c", CHILD, ROOT], capture_output=True, text=True, timeout=600)
    line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
    assert p.returncode == 0 and line, p.stdout[-2000:] + p.stderr[-4000:]
    r = json.loads(line[0][7:])
    assert all(r["files"])
    assert r["encoding_vs_oracle"] <= 1e-3 and r["sh_vs_oracle"] <= 1e-3
    assert r["sigma_net_vs_torch"] <= 2e-3 and r["feat_encoder_vs_torch"] <= 2e-3 and r["color_vs_torch"] <= 2e-3
    assert r["sigma_is_trunc_exp"] <= 2e-3
    assert r["density_shapes"] == [[1000], [1000, 15]] and r["mo_density_shapes"] == [[1000], [1000, 15]] and r["mo_finite"]
    assert 0.0 <= r["rgb_range"][0] and r["rgb_range"][1] <= 1.0 and r["masked_color"]
    assert all(r["grads"].values()), r["grads"]
