"""Summarise an .ncu-rep (ncu --set full) into the per-kernel table kept under profiles/.
  python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_kernels.md"""
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
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
]
STALL = "smsp__pcsamp_warps_issue_stalled_"


def main(path):
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


if __name__ == "__main__":
    main(sys.argv[1])
