"""The part of `tinycudann`'s PyTorch surface the reference's FOC callers use, on this repo's kernels -- so that
nerf/network_tcnn.py:447,480-546,668-681, nerf/multiobjectnetwork.py:5 and COMBINED.py:84 import and run on top of
`focnerf_b200.install_aliases()` (which registers this module as `tinycudann`).

tiny-cuda-nn is a third-party dependency that is ABSENT from the reference tree (no submodule, no pinned version in a requirements
file), so this is an API-level stand-in, not a bit-level one:

  Encoding(n_input_dims, encoding_config, dtype=torch.float16)
      "HashGrid" / "Grid"      -> gridencoder kernels (csrc/hashgrid_*.cu).  Inputs in [0, 1] as tcnn expects; `.params` is the flat
                                  fp32 table, initialised U(-1e-4, 1e-4) like tcnn.  The table LAYOUT is the reference gridencoder's
                                  ((res+1)^3 entries and stride res+1 on the dense levels, gridencoder.cu:87-130), not tcnn's (res^3):
                                  hashed levels index identically, dense levels do not, so a tcnn-trained `encoder.params` blob does
                                  not transfer; the reference's own gridencoder checkpoints do (Trainer.load_checkpoint).
      "SphericalHarmonics"     -> csrc/field.cu SH kernel (degree <= 4), inputs in [0, 1] mapped back to [-1, 1] (network_tcnn.py:573).
  Network(n_input_dims, n_output_dims, network_config)
      "FullyFusedMLP" / "CutlassMLP": `n_neurons` in {16, 32, 64, 128, 256}, `n_hidden_layers` >= 1, n_output_dims <= 16 -> the FFMLP
      kernels (tcgen05 for 64-wide nets with 32 / 64 inputs, csrc/mlp_generic.cu otherwise).  `.params` is the flat fp32 blob in tcnn's
      (= ffmlp.cu:632's) order: [n_neurons, in_pad] + (n_hidden_layers - 1) x [n_neurons, n_neurons] + [16, n_neurons], row-major; the
      input is padded to a multiple of 16 columns with ONES (tcnn wraps a bare Network in an identity encoding whose padding value is
      1, which acts as a bias) and the batch to a multiple of 128 rows, outputs come back as [B, n_output_dims] half.

Both modules always compute in fp16 and return `dtype` (tcnn does not look at autocast)."""
import math

import numpy as np
import torch
import torch.nn as nn

_ACT = {"relu": 0, "exponential": 1, "sine": 2, "sigmoid": 3, "squareplus": 4, "softplus": 5, "none": 6}


def _pad16(n):
    return -(-int(n) // 16) * 16


def _device():
    """tcnn modules are born on the GPU (their parameters are CUDA tensors from the constructor on): multiobjectnetwork.py:35-38 builds
    per-checkpoint models and calls them without a .to(device)."""
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


class Encoding(nn.Module):
    def __init__(self, n_input_dims, encoding_config, dtype=torch.float16, seed=1337):
        super().__init__()
        self.n_input_dims, self.encoding_config, self.dtype, self.seed = int(n_input_dims), dict(encoding_config), dtype, seed
        cfg = self.encoding_config
        self.otype = str(cfg.get("otype", "")).lower()
        if self.otype in ("hashgrid", "grid", "densegrid"):
            from .gridencoder.grid import _gridtype_to_id, _interp_to_id
            L, F = int(cfg.get("n_levels", 16)), int(cfg.get("n_features_per_level", 2))
            pls, H = float(cfg.get("per_level_scale", 2.0)), int(cfg.get("base_resolution", 16))
            max_params = 2 ** int(cfg.get("log2_hashmap_size", 19))
            offsets, offset = [], 0
            for i in range(L):                                          # the gridencoder's sizing rule (gridencoder/grid.py:117-128)
                res = int(np.ceil(H * pls ** i))
                n = min(max_params, (res + 1) ** self.n_input_dims)
                offsets.append(offset)
                offset += int(np.ceil(n / 8) * 8)
            offsets.append(offset)
            self._grid = dict(per_level_scale=pls, base_resolution=H, level_dim=F, gridtype_id=_gridtype_to_id["tiled" if self.otype == "densegrid" else "hash"],
                              interp_id=_interp_to_id["linear"])
            self.register_buffer("offsets", torch.from_numpy(np.array(offsets, dtype=np.int32)).to(_device()))
            gen = torch.Generator().manual_seed(seed)
            self.params = nn.Parameter(((torch.rand(offset * F, generator=gen) * 2 - 1) * 1e-4).to(_device()))
            self.n_output_dims = L * F
        elif self.otype == "sphericalharmonics":
            self.degree = int(cfg.get("degree", 4))
            if not (1 <= self.degree <= 4) or self.n_input_dims != 3:
                raise RuntimeError("tinycudann shim: SphericalHarmonics needs 3 inputs and degree <= 4")
            self.params = nn.Parameter(torch.zeros(0, device=_device()))
            self.n_output_dims = self.degree ** 2
        else:
            raise RuntimeError(f"tinycudann shim: encoding otype {cfg.get('otype')!r} is not one the reference uses")

    def forward(self, x):
        prefix = list(x.shape[:-1])
        x = x.reshape(-1, self.n_input_dims).float()
        if self.otype in ("hashgrid", "grid", "densegrid"):
            from .gridencoder.grid import grid_encode
            g = self._grid
            with torch.autocast("cuda", dtype=torch.float16):           # the fp16 path of the kernels, whatever the caller's autocast state
                out = grid_encode(x, self.params.view(-1, g["level_dim"]), self.offsets, g["per_level_scale"], g["base_resolution"],
                                  x.requires_grad, g["gridtype_id"], False, g["interp_id"], None, None)
        elif self.otype == "sphericalharmonics":
            from .shencoder import sh_encode
            out = sh_encode(x * 2.0 - 1.0, self.degree)
        else:
            raise RuntimeError("unreachable")
        return out.to(self.dtype).view(prefix + [self.n_output_dims])

    def __repr__(self):
        return f"tinycudann.Encoding(shim over focnerf_b200: {self.encoding_config})"


class Network(nn.Module):
    def __init__(self, n_input_dims, n_output_dims, network_config, seed=1337):
        super().__init__()
        self.n_input_dims, self.n_output_dims, self.network_config, self.seed = int(n_input_dims), int(n_output_dims), dict(network_config), seed
        cfg = self.network_config
        if str(cfg.get("otype", "FullyFusedMLP")).lower() not in ("fullyfusedmlp", "cutlassmlp"):
            raise RuntimeError(f"tinycudann shim: network otype {cfg.get('otype')!r}")
        self.hidden = int(cfg.get("n_neurons", 64))
        self.layers = int(cfg.get("n_hidden_layers", 2))
        self.activation = _ACT[str(cfg.get("activation", "ReLU")).lower()]
        self.output_activation = str(cfg.get("output_activation", "None")).lower()
        if self.hidden not in (16, 32, 64, 128, 256) or self.layers < 1 or not (1 <= self.n_output_dims <= 16) or self.output_activation not in _ACT:
            raise RuntimeError("tinycudann shim: n_neurons in {16, 32, 64, 128, 256}, n_hidden_layers >= 1, n_output_dims <= 16")
        self.in_pad = _pad16(self.n_input_dims)
        shapes = [(self.hidden, self.in_pad)] + [(self.hidden, self.hidden)] * (self.layers - 1) + [(16, self.hidden)]
        gen = torch.Generator().manual_seed(seed)
        blobs = []
        for k, (o, i) in enumerate(shapes):                              # tcnn's default initialisation: Xavier uniform per matrix
            fan_out = self.n_output_dims if k == len(shapes) - 1 else o
            a = math.sqrt(6.0 / (i + fan_out))
            w = (torch.rand(o, i, generator=gen) * 2 - 1) * a
            if k == len(shapes) - 1:
                w[self.n_output_dims:] = 0.0                             # padded output rows
            blobs.append(w.reshape(-1))
        self.params = nn.Parameter(torch.cat(blobs).to(_device()))
        self.dtype = torch.float16

    def forward(self, x):
        from .ffmlp.ffmlp import ffmlp_forward
        prefix = list(x.shape[:-1])
        x = x.reshape(-1, self.n_input_dims)
        B = x.shape[0]
        rows = -(-max(B, 1) // 128) * 128
        xin = x.to(torch.float16)
        if self.in_pad > self.n_input_dims:                              # identity-encoding padding: ones
            xin = torch.cat([xin, torch.ones(B, self.in_pad - self.n_input_dims, dtype=torch.float16, device=x.device)], 1)
        if rows > B:
            xin = torch.cat([xin, torch.zeros(rows - B, self.in_pad, dtype=torch.float16, device=x.device)], 0)
        with torch.autocast("cuda", dtype=torch.float16):
            out = ffmlp_forward(xin, self.params, self.in_pad, 16, self.hidden, self.layers, self.activation, 6,
                                not (torch.is_grad_enabled() and (self.params.requires_grad or x.requires_grad)), x.requires_grad)
        out = out[:B, :self.n_output_dims]
        if self.output_activation != "none":
            out = {"relu": torch.relu, "exponential": torch.exp, "sine": torch.sin, "sigmoid": torch.sigmoid,
                   "squareplus": lambda v: 0.5 * (v + torch.sqrt(v * v + 4.0)), "softplus": torch.nn.functional.softplus}[self.output_activation](out)
        return out.view(prefix + [self.n_output_dims])

    def __repr__(self):
        return f"tinycudann.Network(shim over focnerf_b200: {self.n_input_dims} -> {self.network_config} -> {self.n_output_dims})"


class NetworkWithInputEncoding(nn.Module):
    def __init__(self, n_input_dims, n_output_dims, encoding_config, network_config, seed=1337):
        super().__init__()
        self.encoding = Encoding(n_input_dims, encoding_config, seed=seed)
        self.network = Network(self.encoding.n_output_dims, n_output_dims, network_config, seed=seed)
        self.n_input_dims, self.n_output_dims = int(n_input_dims), int(n_output_dims)

    def forward(self, x):
        return self.network(self.encoding(x))
