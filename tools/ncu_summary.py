"""Summarise an .ncu-rep (ncu --set full) into the per-kernel table kept under profiles/.
  python tools/ncu_summary.py gpurun_out/prof.ncu-rep [traffic.json] > profiles/rNN_kernels.md
With a second argument also writes {kernel: {launches, time_us, dram_bytes, l2_bytes, tensor_inst, tensor_pipe_pct}} (means per launch):
bench.py reads profiles/r2_traffic.json for `roofline.traffic` and the L2 GB/s column."""
import json
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time_us"),
    ("dram__bytes_read.sum", "dram_rd_MB"),
    ("dram__bytes_write.sum", "dram_wr_MB"),
    ("lts__t_bytes.sum", "l2_MB"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "mem_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ_pct"),
    ("smsp__inst_executed.sum", "warp_inst"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor_inst"),
    ("sm__inst_executed_pipe_tensor_subpipe_hmma.sum", "tensor_hmma_inst"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_pct"),
    ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor_hmma_pct"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
]
STALL = "smsp__pcsamp_warps_issue_stalled_"


def main(path, json_out=None):
    agg = {}
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, units = rows[0], rows[1]
    col = {k: i for i, k in enumerate(h)}
    print(f"# {path}\n")
    print("| kernel | " + " | ".join(n for _, n in KEYS) + " | top stalls (pc samples) |")
    print("|---|" + "---|" * (len(KEYS) + 1))
    for r in rows[2:]:
        name = r[col["Kernel Name"]].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
        vals = []
        rec = {}
        for k, n in KEYS:
            if k not in col:
                vals.append("-"); continue
            v, u = r[col[k]].replace(",", ""), units[col[k]]
            try:
                f = float(v)
            except ValueError:
                vals.append(v); continue
            if n == "time_us":
                f = f / 1000 if u == "ns" else (f * 1000 if u == "ms" else f)
            if n.endswith("_MB"):
                f = {"byte": f / 1e6, "Kbyte": f / 1e3, "Mbyte": f, "Gbyte": f * 1e3}.get(u, f)
            rec[n] = f
            vals.append(f"{f:.1f}" if abs(f) < 1e6 else f"{f:.3g}")
        st = []
        for k, i in col.items():
            if k.startswith(STALL) and not k.endswith("_not_issued"):
                try:
                    st.append((float(r[i].replace(",", "")), k[len(STALL):]))
                except ValueError:
                    pass
        st.sort(reverse=True)
        tot = sum(v for v, _ in st) or 1
        top = ", ".join(f"{n} {100 * v / tot:.0f}%" for v, n in st[:4])
        print(f"| {name} | " + " | ".join(vals) + f" | {top} |")
        a = agg.setdefault(name, {"launches": 0, "time_us": 0.0, "dram_bytes": 0.0, "l2_bytes": 0.0, "tensor_inst": 0.0, "tensor_pipe_pct": 0.0})
        a["launches"] += 1
        a["time_us"] += rec.get("time_us", 0.0)
        a["dram_bytes"] += 1e6 * (rec.get("dram_rd_MB", 0.0) + rec.get("dram_wr_MB", 0.0))
        a["l2_bytes"] += 1e6 * rec.get("l2_MB", 0.0)
        a["tensor_inst"] += rec.get("tensor_inst", 0.0)
        a["tensor_pipe_pct"] += rec.get("tensor_pipe_pct", 0.0)
    if json_out:
        for a in agg.values():
            for k in ("time_us", "dram_bytes", "l2_bytes", "tensor_inst", "tensor_pipe_pct"):
                a[k] = round(a[k] / a["launches"], 3)
        json.dump({"source": path, "per_launch_means": agg}, open(json_out, "w"), indent=1)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)
