"""Per-ENTRY-POINT traffic from the per-kernel means tools/ncu_summary.py writes (its second argument):
  python tools/ncu_summary.py gpurun_out/r2_prof.ncu-rep /tmp/k.json > profiles/r2_kernels.md
  python tools/ncu_traffic.py /tmp/k.json SAMPLES > profiles/r2_traffic.json
bench.py reads profiles/r2_traffic.json for `roofline.traffic` (DRAM bytes per call of the dominant entry point) and for the L2 GB/s
column of `by_entry_point`; SAMPLES = samples per step of the profiled step, so that bench.py can scale to its own step."""
import json
import sys

# kernels one call of each entry point launches (k_mlp_bwd / k_mlp_wgrad_finish run once per call, two calls per step)
ENTRY = {
    "foc_march_rays_train": ["k_march_record_const", "k_march_expand<1>"],
    "foc_grid_encode_forward": ["k_grid_fwd_tile<__half, 3, 2>"],
    "foc_field_forward": ["k_field_fwd<1>"],
    "foc_composite_mse_forward": ["k_composite_mse_fwd<0>", "k_mse_finish"],
    "foc_composite_mse_backward": ["k_composite_mse_bwd<0>"],
    "foc_ffmlp_backward_head": ["k_mlp_bwd<32, 0>", "k_mlp_wgrad_finish"],
    "foc_grid_encode_backward": ["k_grid_bwd_fused<__half, 3, 2, 1>"],
    "foc_adam_nonfinite": ["k_adam_nonfinite"],
    "foc_adam_step": ["k_adam_step<0>", "k_adam_step<0, 0>"],
}


def main(path, samples):
    k = json.load(open(path))["per_launch_means"]
    out = {}
    for ep, names in ENTRY.items():
        rows = [k[n] for n in names if n in k]
        if not rows:
            continue
        out[ep] = {"dram_bytes": round(sum(r["dram_bytes"] for r in rows)), "l2_bytes": round(sum(r["l2_bytes"] for r in rows)),
                   "ncu_time_us": round(sum(r["time_us"] for r in rows), 1), "tensor_inst": round(sum(r["tensor_inst"] for r in rows)),
                   "tensor_pipe_pct": round(max(r["tensor_pipe_pct"] for r in rows), 1), "kernels": [n for n in names if n in k]}
    json.dump({"source": "ncu --set full + lts__t_bytes / sm__inst_executed_pipe_tensor / sm__pipe_tensor_cycles_active, one eager step "
                         "(bench.py --no-graph --steps 1), per CALL of each entry point (mean over the calls of the step)",
               "samples_per_step": samples, "entry_points": out}, sys.stdout, indent=1)


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]))
