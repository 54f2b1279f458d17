"""Fill the @@TABLE@@ placeholders of DESIGN.md.in from the committed bench lines (tools/design_tables.py) -> DESIGN.md."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = open(os.path.join(ROOT, "DESIGN.md.in")).read()
for key, arg in (("@@STATUS_TABLE@@", "status"), ("@@KERNEL_TABLE@@", "kernels"), ("@@SCALING_TABLE@@", "scaling")):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "design_tables.py"), arg], capture_output=True, text=True)
    if out.returncode != 0:
        raise SystemExit(out.stderr)
    src = src.replace(key, out.stdout.rstrip())
open(os.path.join(ROOT, "DESIGN.md"), "w").write("<!-- generated from DESIGN.md.in by tools/assemble_design.py: edit that file, the tables come from profiles/r2_*.json -->\n" + src)

# everything after the <!-- round2 --> marker is generated
b = open(os.path.join(ROOT, "BASELINE.md")).read()
head = b[:b.index("<!-- round2 -->") + len("<!-- round2 -->")]
tabs = {a: subprocess.run([sys.executable, os.path.join(ROOT, "tools", "design_tables.py"), a], capture_output=True, text=True).stdout.rstrip() for a in ("status", "scaling")}
open(os.path.join(ROOT, "BASELINE.md"), "w").write(head + """
## 6. Measured in round 2 (one B200 unless N is given; `profiles/r2_bench.json`, `profiles/r2_bench_reference.json`, `profiles/r2_scale_n*.json`)

Both arms run the reference's recipe (lr 1e-3, learnt density grid refreshed every 16 steps inside the timed region, 1 M-ray pool,
1024 pretraining steps; DESIGN.md 7.1), so a step marches about twice the samples of round 1's frozen-grid workload -- the two rounds'
rays/s are not the same workload; round 1's is kept as `frozen_bitfield`.  The reference arm is the reference's own unmodified Python
stack (its Trainer, run_cuda, L1 wrappers) on its own CUDA extensions built for sm_100a.

""" + tabs["status"] + "\n\nScaling on one 8 x B200 box (`bench.py --gpus N` under torchrun; train = weak scaling, render = strong scaling):\n\n" + tabs["scaling"] + """

The reference's GUI / OpenGL viewer is unavailable on this box and not part of the path.
""")

# profiles/r2_scaling.md from its template
sc = open(os.path.join(ROOT, "profiles", "r2_scaling.md.in")).read()
open(os.path.join(ROOT, "profiles", "r2_scaling.md"), "w").write(sc.replace("@@SCALING_TABLE@@", tabs["scaling"]))
