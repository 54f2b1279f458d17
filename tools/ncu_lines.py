"""Attribute ncu pc-sampling data (SASS page of an .ncu-rep) to CUDA source lines via nvdisasm line info.
  python tools/ncu_lines.py <rep> <kernel-name-substring> <object.o> <mangled-substring>"""
import csv
import io
import re
import subprocess
import sys
import tempfile
import os


def sass_lines(obj, mangled):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, capture_output=True)
    cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
    out, on, line = [], False, None
    for l in txt.splitlines():
        if l.startswith("\t.section\t.text."):
            on = mangled in l
            continue
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            line = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
            out.append(line)
    return out


def main(rep, kname, obj, mangled):
    lines = sass_lines(obj, mangled)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kname], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h = rows[1]
    data = [r for r in rows[2:] if len(r) == len(h) and r[0].startswith("0x")]
    if len(data) >= 2 * len(lines) - 4:
        data = data[:len(data) // 2]
    isamp, iex = h.index("# Samples"), h.index("Instructions Executed")
    print(f"sass in report {len(data)}, sass with line info {len(lines)}")
    agg = {}
    for i, r in enumerate(data):
        key = lines[i] if i < len(lines) and lines[i] else ("?", 0)
        a = agg.setdefault(key, [0, 0])
        a[0] += int(r[isamp]); a[1] += int(r[iex])
    ts, te = sum(a[0] for a in agg.values()) or 1, sum(a[1] for a in agg.values()) or 1
    src = {}
    for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:28]:
        if f not in src:
            for root in ("focnerf_b200/csrc", "."):
                pth = os.path.join(root, f)
                if os.path.exists(pth):
                    src[f] = open(pth).read().splitlines(); break
            else:
                src[f] = []
        text = src[f][ln - 1].strip()[:100] if 0 < ln <= len(src[f]) else ""
        print(f"{100 * a[0] / ts:5.1f}% samp {100 * a[1] / te:5.1f}% exec  {f}:{ln}  {text}")


if __name__ == "__main__":
    main(*sys.argv[1:5])
