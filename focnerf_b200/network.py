"""Field network with the wiring of the reference's nerf/network_ff.py:10-149 (hash grid -> FFMLP sigma net -> trunc_exp;
SH-4 directions + 15 geo features + 1 zero pad -> FFMLP colour net -> sigmoid), as a plain nn.Module (the reference
class also IS its renderer by inheritance; here rendering lives in trainer.py).  Parameter names match the reference
(`encoder.embeddings`, `sigma_net.weights`, `color_net.weights`) so state_dicts interchange."""
import torch
import torch.nn as nn

from .activation import trunc_exp
from .encoding import get_encoder
from .ffmlp import FFMLP


class NeRFNetwork(nn.Module):
    def __init__(self, encoding="hashgrid", encoding_dir="sphere_harmonics", num_layers=2, hidden_dim=64, geo_feat_dim=15,
                 num_layers_color=3, hidden_dim_color=64, bound=1, **kwargs):
        super().__init__()
        self.bound = bound
        import os
        self._field_fused = os.environ.get("FOC_FIELD_FUSED", "1") == "1"       # A/B switch for forward_fused's one-kernel field evaluation
        self.num_layers, self.hidden_dim, self.geo_feat_dim = num_layers, hidden_dim, geo_feat_dim
        self.encoder, self.in_dim = get_encoder(encoding, desired_resolution=2048 * bound)
        self.sigma_net = FFMLP(input_dim=self.in_dim, output_dim=1 + self.geo_feat_dim, hidden_dim=self.hidden_dim,
                               num_layers=self.num_layers)
        self.num_layers_color, self.hidden_dim_color = num_layers_color, hidden_dim_color
        self.encoder_dir, self.in_dim_color = get_encoder(encoding_dir)
        self.in_dim_color += self.geo_feat_dim + 1                  # padded to 32 (network_ff.py:42)
        self.color_net = FFMLP(input_dim=self.in_dim_color, output_dim=3, hidden_dim=self.hidden_dim_color,
                               num_layers=self.num_layers_color)

    def forward(self, x, d):
        x = self.encoder(x, bound=self.bound)
        h = self.sigma_net(x)
        sigma = trunc_exp(h[..., 0])
        geo_feat = h[..., 1:]
        d = self.encoder_dir(d)
        p = torch.zeros_like(geo_feat[..., :1])
        h = torch.cat([d.to(geo_feat.dtype), geo_feat, p], dim=-1)
        h = self.color_net(h)
        rgb = torch.sigmoid(h)
        return sigma, rgb

    def forward_fused(self, x, d, raw_colour=False):
        """Same function as forward() for a batch whose size is a multiple of 128 (every march output is), with the
        elementwise wiring fused (field_glue.cu) and the FFMLP kernels called without the module's zero-row padding cat
        (ffmlp.py:157-159 exists only to reach a multiple of 128).  Must run under fp16 autocast, like forward().
        raw_colour=True returns (sigma, hc [B,16] f16) for raymarching.composite_rays_train_mse_hc, which applies the sigmoid itself."""
        from .field_glue import field_mid, rgb_sigmoid
        B = x.shape[0]
        if B % 128 != 0 or self.sigma_net.hidden_dim != 64 or self.in_dim != 32 or self.in_dim_color != 32:
            if raw_colour:
                raise RuntimeError("forward_fused(raw_colour=True) needs a batch that is a multiple of 128 and the 32 -> 64 -> 16 FOC networks")
            return self.forward(x, d)
        sn, cn = self.sigma_net, self.color_net
        enc = self.encoder(x, bound=self.bound)
        if (not raw_colour and self._field_fused and torch.is_autocast_enabled() and getattr(sn, "_shadow16", None) is not None
                and getattr(cn, "_shadow16", None) is not None and sn.activation == 0 and cn.activation == 0 and cn.hidden_dim == 64
                and sn.num_layers <= 4 and cn.num_layers <= 4):
            # the whole field behind the encoder as one tcgen05 kernel (csrc/field_fwd.cu), bit-identical to the four launches below
            from .field_glue import field_eval
            return field_eval(enc, d, sn, cn, self.training and torch.is_grad_enabled())
        h = sn.call_fused(enc, enc.requires_grad)
        sigma, cin = field_mid(h, d)
        hc = cn.call_fused(cin, cin.requires_grad)
        return (sigma, hc) if raw_colour else (sigma, rgb_sigmoid(hc))

    def density(self, x):
        x = self.encoder(x, bound=self.bound)
        h = self.sigma_net(x)
        return {'sigma': trunc_exp(h[..., 0]), 'geo_feat': h[..., 1:]}

    def color(self, x, d, mask=None, geo_feat=None, **kwargs):
        """Masked colour evaluation (network_ff.py:92-123): rgb is 0 where mask is False."""
        if mask is not None:
            rgbs = torch.zeros(mask.shape[0], 3, dtype=x.dtype, device=x.device)
            if not mask.any():
                return rgbs
            x, d, geo_feat = x[mask], d[mask], geo_feat[mask]
        d = self.encoder_dir(d)
        p = torch.zeros_like(geo_feat[..., :1])
        h = torch.cat([d.to(geo_feat.dtype), geo_feat, p], dim=-1)
        h = torch.sigmoid(self.color_net(h))
        if mask is not None:
            rgbs[mask] = h.to(rgbs.dtype)
            return rgbs
        return h

    def get_params(self, lr):
        return [{'params': self.encoder.parameters(), 'lr': lr}, {'params': self.sigma_net.parameters(), 'lr': lr},
                {'params': self.encoder_dir.parameters(), 'lr': lr}, {'params': self.color_net.parameters(), 'lr': lr}]
