"""Aggregate an ncu launch-list CSV (--metrics gpu__time_duration.sum --csv) per kernel.
  python tools/ncu_launches.py gpurun_out/launches.csv STEPS > profiles/rNN_launches.md"""
import collections
import csv
import re
import sys


def main(path, steps):
    rows = list(csv.reader(open(path, errors="ignore")))
    hdr, agg, n = None, collections.OrderedDict(), 0
    for r in rows:
        if "Kernel Name" in r:
            hdr = r; continue
        if hdr is None or len(r) != len(hdr):
            continue
        d = dict(zip(hdr, r))
        if d["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", d["Kernel Name"]).replace("void ", "").replace("<unnamed>::", "")[:100]
        v = float(d["Metric Value"].replace(",", ""))
        v = v / 1000 if d["Metric Unit"] == "ns" else (v * 1000 if d["Metric Unit"] == "ms" else v)
        a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v; n += 1
    tot = sum(a[1] for a in agg.values())
    print(f"# {path}: {n} launches in {steps} step(s); {n / steps:.0f} launches and {tot / steps:.1f} us of GPU time per step")
    print("# (ncu serialises launches and flushes caches: per-launch times are cold; shares, not absolutes, compare with bench.py)\n")
    print("| us / step | share | launches / step | kernel |\n|---|---|---|---|")
    ours = 0.0
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| {a[1] / steps:.1f} | {100 * a[1] / tot:.1f} % | {a[0] / steps:g} | `{k}` |")
        if k.startswith("k_"):
            ours += a[1]
    print(f"\nlibfoc_b200 kernels: {100 * ours / tot:.1f} % of the GPU time; the rest is torch elementwise / reduction / fill kernels.")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]))
