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
