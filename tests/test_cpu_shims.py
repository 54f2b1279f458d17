"""CPU-side checks of the host logic added in round 2 (no kernel runs): the tinycudann shim's shapes and layouts, the alias
registration, the speculative-render sizing rule and the data-parallel shard arithmetic."""
import sys

import numpy as np
import pytest
import torch


def test_install_aliases_registers_tinycudann_only_when_the_name_is_free():
    import focnerf_b200
    saved = sys.modules.pop("tinycudann", None)
    try:
        focnerf_b200.install_aliases()
        import tinycudann
        assert tinycudann is focnerf_b200.tinycudann
        sentinel = type(sys)("tinycudann")
        sys.modules["tinycudann"] = sentinel
        focnerf_b200.install_aliases()
        assert sys.modules["tinycudann"] is sentinel                 # a real tiny-cuda-nn is left alone
    finally:
        sys.modules.pop("tinycudann", None)
        if saved is not None:
            sys.modules["tinycudann"] = saved


def test_tcnn_shim_shapes_match_the_reference_call_sites():
    """nerf/network_tcnn.py:480-546: HashGrid 16 x 2 / 2^19 / base 16, SH degree 4, sigma 32->64->16, colour 47->64->64->3, feature
    encoder 144->16->16."""
    from focnerf_b200 import tinycudann as tcnn
    from focnerf_b200.gridencoder import GridEncoder
    pls = float(np.exp2(np.log2(2048 * 1 / 16) / (16 - 1)))
    enc = tcnn.Encoding(3, {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16,
                            "per_level_scale": pls})
    ref = GridEncoder(input_dim=3, num_levels=16, level_dim=2, per_level_scale=pls, base_resolution=16, log2_hashmap_size=19)
    assert enc.n_output_dims == 32 and enc.params.dim() == 1 and enc.params.numel() == ref.embeddings.numel()
    assert torch.equal(enc.offsets.cpu(), ref.offsets.cpu())
    assert float(enc.params.detach().abs().max()) <= 1e-4
    sh = tcnn.Encoding(3, {"otype": "SphericalHarmonics", "degree": 4})
    assert sh.n_output_dims == 16 and sh.params.numel() == 0
    cfg = lambda n, h: {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": n, "n_hidden_layers": h}
    sigma, colour, feat = tcnn.Network(32, 16, cfg(64, 1)), tcnn.Network(47, 3, cfg(64, 2)), tcnn.Network(144, 16, cfg(16, 1))
    assert sigma.params.numel() == 64 * 32 + 16 * 64
    assert colour.in_pad == 48 and colour.params.numel() == 64 * 48 + 64 * 64 + 16 * 64
    assert feat.params.numel() == 16 * 144 + 16 * 16
    w_out = colour.params.detach()[-16 * 64:].view(16, 64)
    assert float(w_out[3:].abs().max()) == 0.0 and float(w_out[:3].abs().max()) > 0            # padded output rows start at zero
    with pytest.raises(RuntimeError):
        tcnn.Network(32, 17, cfg(64, 2))
    with pytest.raises(RuntimeError):
        tcnn.Encoding(3, {"otype": "OneBlob"})


def test_render_capacity_rule():
    """Trainer._render_size: capacity = total + 12.5 % in 16 384-row quanta, kept while the total stays inside [cap / 1.5, cap / 1.03]."""
    from focnerf_b200.trainer import Trainer, _pad128
    t = Trainer.__new__(Trainer)
    t._render_cap = {}
    t._render_size(1000, 1_000_000)
    cap = t._render_cap[1000]
    assert cap % 16384 == 0 and 1_125_000 <= cap < 1_125_000 + 2 * 16384
    t._render_size(1000, 1_050_000)                                 # inside the band: unchanged
    assert t._render_cap[1000] == cap
    t._render_size(1000, int(cap / 1.02))                           # too close to the capacity: grows
    assert t._render_cap[1000] > cap
    t._render_size(1000, 100_000)                                   # far below: shrinks
    assert t._render_cap[1000] < cap
    assert _pad128(0) == 128 and _pad128(129) == 256


def test_generic_ffmlp_module_accepts_the_reference_shapes():
    from focnerf_b200.ffmlp import FFMLP
    for args in ((144, 16, 16, 2), (48, 3, 64, 3), (32, 16, 128, 5), (64, 8, 256, 2), (16, 1, 32, 2)):
        m = FFMLP(*args)
        i, o, h, l = args
        assert m.weights.numel() == h * (i + h * (l - 1) + 16)


def test_dp_shard_ranges_tile_the_table():
    """optim.shard_range (PeerAdam): the ranks' slices are disjoint, ordered, cover the table exactly and start on 4096-element blocks."""
    from focnerf_b200.optim import shard_range
    for n in (8, 4096, 12239728, 12239728 + 8, 50000 * 2):
        for world in (1, 2, 3, 4, 8):
            pos = 0
            for r in range(world):
                per, start, length = shard_range(n, world, r)
                assert per % 4096 == 0 and start == min(r * per, n) and start % 4096 == 0 or start == n
                assert start == pos and length >= 0
                pos += length
            assert pos == n
