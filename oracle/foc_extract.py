"""Run the reference's OWN multi-object composition source lines (test / bench infrastructure; never imported by the product).

COMBINED.py and editable.py are scripts: `MultiTrainer` and its methods are defined inside `if __name__ == '__main__':`, next to
argument parsing, dataset and YOLO loading, so they cannot be imported.  This module locates the methods by name with `ast` in
the verbatim copies oracle/build_ref.py placed in oracle/_ref/py/foc/, takes their source text UNCHANGED (only dedented), and
executes it as methods of a small host class:

    COMBINED.py  MultiTrainer.run                            per-object uniform sampling + density + cumprod weights + masked colour
                 MultiTrainer.batch_run                      the 4096-ray batching of run over a frame -> [1,N,T] / [1,N,T,3] tensors
                 MultiTrainer.best_densities_and_colors_v3   running arg-max over objects (strict '>')
                 MultiTrainer.image_depth_generation         the final cumprod composite, 4th channel sum w*sigma, background, clamp
                 MultiTrainer.batch_image_depth_generation   its 4096-ray batching
    editable.py  MultiTrainer.modify_rays_for_object         the per-object origin offset + direction re-normalisation

The module-level names those bodies read (`opt`, `device`, `model`, `raymarching`, `torch`, `F`) are provided in the exec namespace;
`raymarching` is whatever stack is active in the process (the reference's own extension wrappers in the reference arm, focnerf_b200's
under install_aliases()).  The per-image driver loop (COMBINED.py:585-617, editable.py:658-692) is restated in `Compositor.render`
-- it is interleaved with checkpoint loading and metric code in the reference and is a dozen lines.
"""
import argparse
import ast
import os
import sys
import textwrap

HERE = os.path.dirname(os.path.abspath(__file__))
FOC = os.path.join(HERE, "_ref", "py", "foc")
_WANT = {"COMBINED.py": ("run", "batch_run", "best_densities_and_colors_v3", "image_depth_generation", "batch_image_depth_generation"),
         "editable.py": ("modify_rays_for_object",)}


def present():
    return all(os.path.exists(os.path.join(FOC, f)) for f in _WANT)


def method_sources(fname, names, cls="MultiTrainer"):
    """-> {name: dedented source text of `cls.name`}, located by ast (no execution of the script)."""
    path = os.path.join(FOC, fname)
    text = open(path).read()
    lines = text.splitlines(keepends=True)
    out = {}
    for node in ast.walk(ast.parse(text)):
        if isinstance(node, ast.ClassDef) and node.name == cls:
            for f in node.body:
                if isinstance(f, ast.FunctionDef) and f.name in names:
                    out[f.name] = textwrap.dedent("".join(lines[f.lineno - 1:f.end_lineno]))
    missing = [n for n in names if n not in out]
    if missing:
        raise RuntimeError(f"{fname}: methods {missing} of {cls} not found")
    return out


def build_host(num_steps=512, device="cuda"):
    """-> (Host class carrying the reference's methods, the namespace they execute in)."""
    import torch
    import raymarching                                      # the ACTIVE stack's operator package
    opt = argparse.Namespace(num_steps=num_steps, upsample_steps=0, edit_object="edited", offset_x=0.0, offset_y=0.0, offset_z=0.0)
    ns = {"torch": torch, "F": torch.nn.functional, "raymarching": raymarching, "opt": opt, "device": torch.device(device), "model": None}
    body = ""
    for fname, names in _WANT.items():
        for name, src in method_sources(fname, names).items():
            body += textwrap.indent(src, "    ") + "\n"
    exec(compile("class Host:\n" + body, "<reference COMBINED.py / editable.py methods>", "exec"), ns)
    return ns["Host"], ns


class _Object:
    """What MultiTrainer.run needs from an object network (nerf/network_tcnn.py's surface: color takes yolo_details positionally)."""

    def __init__(self, net, aabb, min_near):
        self.net, self.aabb_train, self.aabb_infer, self.min_near, self.training = net, aabb, aabb, min_near, False

    def density(self, x):
        return self.net.density(x)

    def color(self, x, d, yolo_details, mask=None, geo_feat=None, **kwargs):
        return self.net.color(x, d, mask=mask, geo_feat=geo_feat)


class Compositor:
    """K resident object networks composited exactly as the reference's evaluate loop does.
    models: objects with .density(x) -> {'sigma','geo_feat'} and .color(x, d, mask=, geo_feat=) (nerf/network_ff.py:76-123)."""

    def __init__(self, models, num_steps=512, offsets=None, min_near=0.2, aabb=None, fp16=True):
        import torch
        dev = next(models[0].parameters()).device
        if aabb is None:
            aabb = torch.tensor([-1.0, -1.0, -1.0, 1.0, 1.0, 1.0], device=dev)
        self.Host, self.ns = build_host(num_steps, dev)
        self.objects = [_Object(m, aabb, min_near) for m in models]
        self.offsets = list(offsets) if offsets is not None else [None] * len(models)
        self.fp16, self.T = fp16, num_steps

    def render(self, rays_o, rays_d, bg="white"):
        """-> image [N,4] (rgb + sum w*sigma, clamped), depth [N] of rays [N,3]"""
        import torch
        h, opt = self.Host(), self.ns["opt"]
        N = rays_o.shape[0]
        h.B, h.N, h.H, h.W, h.cuda_ray = 1, N, 1, N, False
        h.get_object_type_from_ckpt = lambda ckpt: ckpt         # editable.py:452 compares it with opt.edit_object
        max_d = max_c = None
        with torch.no_grad():
            for obj, off in zip(self.objects, self.offsets):
                h.model = obj                                   # aabb / min_near / training flag (COMBINED.py:458-460)
                self.ns["model"] = obj                          # batch_run passes the module-level `model` to run (COMBINED.py:226)
                with torch.cuda.amp.autocast(enabled=self.fp16):
                    if off is not None:
                        opt.offset_x, opt.offset_y, opt.offset_z = (float(v) for v in off)
                    o, d = h.modify_rays_for_object(rays_o, rays_d, None, "edited" if off is not None else "other")
                    res = h.batch_run(o, d, None)
                    dens, rgbs = res["densities"], res["rgbs"]
                    if max_d is None:
                        max_d, max_c = dens, rgbs
                    else:
                        max_d, max_c = h.best_densities_and_colors_v3(dens, max_d, rgbs, max_c)
            h.model = self.objects[0]
            image, depth = h.batch_image_depth_generation({"rays_o": rays_o.view(1, N, 3), "rays_d": rays_d.view(1, N, 3)}, max_d, max_c, bg)
        return image.view(N, 4), depth.view(N)
