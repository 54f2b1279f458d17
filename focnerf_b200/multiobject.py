"""FOC-NeRF multi-object composition, host side (configs[3] / configs[4]-edit): the work of COMBINED.py's MultiTrainer
(`run` COMBINED.py:451-534 per object, `best_densities_and_colors_v3` :247-251, `image_depth_generation` :141-200) and
of editable.py's per-object ray offsets (`modify_rays_for_object` editable.py:443-471), over libfoc_b200.

The reference loads each object's checkpoint from disk per image, evaluates it in 4096-ray batches with eager PyTorch
(linspace / cumprod / boolean-mask gather-scatter for the colour net), keeps [K,N,T] densities and [K,N,T,3] colours plus
a running-max copy of both, and composites with another cumprod pass.  Here the K field networks stay resident, a
ray chunk is sampled by one kernel per object (foc_sample_uniform), each field is ONE fused evaluation (hash grid ->
tcgen05 sigma net -> glue -> tcgen05 colour net -> sigmoid) over the whole [chunk, T] block, and selection + per-object
colour masks + the 4-channel composite are one warp-per-ray kernel (foc_composite_multi).  The colour of samples the
reference masks out (`w_k <= 1e-10`) is evaluated and then dropped by the compositor: same pixels, no gather/scatter.
"""
import torch

from . import foc, raymarching


class MultiObjectRenderer:
    """K resident objects composited per sample by arg-max density.

    models  : list of field networks (network.NeRFNetwork or anything with forward(x [B,3], d [B,3]) -> (sigma [B], rgb [B,3]))
    aabb    : [6] tensor, the shared inference box (every object of the reference is trained in the same normalised box)
    offsets : per-object ray-origin offsets ((x,y,z) or None), editable.py:443-471; an offset object also gets its
              direction re-normalised, as the reference does
    """

    def __init__(self, models, aabb, offsets=None, num_steps=512, min_near=0.2, colour_mask=True, fp16=True):
        self.models = [m.eval() for m in models]
        self.K = len(self.models)
        if self.K == 0:
            raise ValueError("MultiObjectRenderer: at least one object")
        self.aabb = aabb.to(torch.float32).contiguous()
        self.offsets = list(offsets) if offsets is not None else [None] * self.K
        if len(self.offsets) != self.K:
            raise ValueError("MultiObjectRenderer: one offset (or None) per object")
        self.num_steps, self.min_near, self.colour_mask, self.fp16 = int(num_steps), float(min_near), bool(colour_mask), bool(fp16)
        self._buf = {}
        if self.fp16:
            # resident fp16 copies of the MLP weights: the objects then take forward_fused's one-kernel field evaluation (csrc/field_fwd.cu)
            # instead of four launches and three round trips per object and chunk (the copies follow later writes of the parameters)
            for m in self.models:
                for net in (getattr(m, "sigma_net", None), getattr(m, "color_net", None)):
                    if net is not None and hasattr(net, "enable_fp16_side_buffers") and getattr(net, "_shadow16", None) is None \
                            and net.weights.is_cuda:
                        net.enable_fp16_side_buffers()

    def _buffers(self, n, dev):
        key = (n, str(dev))
        if key not in self._buf:
            T, K = self.num_steps, self.K
            self._buf = {key: (torch.empty(K, n, T, dtype=torch.float32, device=dev), torch.empty(K, n, T, 3, dtype=torch.float32, device=dev),
                               torch.empty(K, n, dtype=torch.float32, device=dev), torch.empty(K, n, dtype=torch.float32, device=dev))}
        return self._buf[key]

    @torch.no_grad()
    def run(self, k, rays_o, rays_d):
        """One object over one ray chunk (COMBINED.py:451-534): -> (sigma [n,T], rgb [n,T,3], nears [n], fars [n])."""
        T = self.num_steps
        off = self.offsets[k]
        o, d, nears, fars, xyzs = foc.sample_uniform(rays_o, rays_d, self.aabb, T, self.min_near, offset=off, normalize_d=off is not None)
        n = o.shape[0]
        dirs = d.view(n, 1, 3).expand(n, T, 3).reshape(-1, 3)
        m = self.models[k]
        with torch.autocast("cuda", dtype=torch.float16, enabled=self.fp16):
            f = m.forward_fused if (self.fp16 and hasattr(m, "forward_fused")) else m
            sigma, rgb = f(xyzs.view(-1, 3), dirs)
        return sigma.view(n, T), rgb.view(n, T, 3), nears, fars

    @torch.no_grad()
    def render_chunk(self, rays_o, rays_d, bg=1.0):
        """-> (image [n,4] = rgb + sum w*sigma, clamped to [0,1]; depth [n]; weights_sum [n]) of one ray chunk."""
        rays_o, rays_d = rays_o.contiguous().view(-1, 3), rays_d.contiguous().view(-1, 3)
        n, dev = rays_o.shape[0], rays_o.device
        sig, rgb, on, of = self._buffers(n, dev)
        for k in range(self.K):
            s, c, nk, fk = self.run(k, rays_o, rays_d)
            sig[k].copy_(s); rgb[k].copy_(c); on[k].copy_(nk); of[k].copy_(fk)
        # the composite's own z_vals come from the UNMODIFIED rays (COMBINED.py:148-153)
        nears, fars = raymarching.near_far_from_aabb(rays_o, rays_d, self.aabb, self.min_near)
        if self.colour_mask:
            return foc.composite_multi(sig, rgb, nears, fars, bg=bg, obj_nears=on, obj_fars=of)
        return foc.composite_multi(sig, rgb, nears, fars, bg=bg)

    @torch.no_grad()
    def render(self, rays_o, rays_d, bg=1.0, chunk=16384):
        """A whole frame in ray chunks (the reference: 4096-ray batches, COMBINED.py:211-241).  chunk x T x K samples are resident
        at a time: 16384 x 512 x 4 = 33.5 M field evaluations, ~6 GB of activations."""
        rays_o, rays_d = rays_o.contiguous().view(-1, 3), rays_d.contiguous().view(-1, 3)
        N, dev = rays_o.shape[0], rays_o.device
        image = torch.empty(N, 4, dtype=torch.float32, device=dev)
        depth = torch.empty(N, dtype=torch.float32, device=dev)
        ws = torch.empty(N, dtype=torch.float32, device=dev)
        for h in range(0, N, chunk):
            t = min(h + chunk, N)
            i, d, w = self.render_chunk(rays_o[h:t], rays_d[h:t], bg)
            image[h:t], depth[h:t], ws[h:t] = i, d, w
        return image, depth, ws
