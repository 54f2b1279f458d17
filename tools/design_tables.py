"""Print the markdown tables of DESIGN.md / BASELINE.md / profiles/r2_scaling.md from the committed bench lines, so that no number in
the documents is typed by hand:  python tools/design_tables.py status|kernels|scaling"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "profiles")

KERNELS = [
    # entry point (graph-mode name), kernels, replaces, bound, algorithmic bytes / unit
    ("foc_march_rays_train_rows", "`k_march_record_const` + `k_march_expand` (raymarch.cu)", "K6 raymarching.cu:312-490", "latency (serial dependence inside a ray)", "36 B in + 12 B out / ray, 32 B out / sample"),
    ("foc_grid_encode_forward_rows", "`k_grid_fwd_tile<half,3,2>` (hashgrid_fwd.cu)", "K11 gridencoder.cu:87-245", "instruction issue / L1 gather wavefronts; table L2-resident", "12 + 512 + 64 = 588 B / sample"),
    ("foc_field_forward", "`k_field_fwd<train>` (field_fwd.cu)", "K15 ×2 ffmlp.cu:47-407 + ~10 torch kernels of network_ff.py:51-74", "HBM (stored activations)", "796 B / sample, 36 864 FLOP"),
    ("foc_composite_mse_forward", "`k_composite_mse_fwd` + `k_mse_finish` (composite.cu)", "K7 raymarching.cu:501-588 + bg mix + MSE", "latency (warp-per-ray scan)", "24 B / sample"),
    ("foc_composite_mse_backward_rows", "`k_composite_mse_bwd` (composite.cu)", "K8 raymarching.cu:602-693 + loss gradient + zero fill", "latency", "40 B / sample"),
    ("foc_ffmlp_backward_head", "`k_mlp_bwd<32,relu>` + `k_mlp_wgrad_finish`, ×2 (mlp_bwd.cu)", "K16 + K17 ffmlp.cu:410-518, 749-895 + sigmoid / trunc_exp backward", "HBM", "colour 536 B + sigma 424 B / sample, 73 728 FLOP"),
    ("foc_grid_encode_backward_rows", "`k_grid_bwd_fused` = runs + scatter side by side (hashgrid_bwd.cu)", "K12 gridencoder.cu:248-340", "L2 reduction units", "12 + 64 + 512 = 588 B / sample"),
    ("foc_adam_nonfinite", "`k_adam_nonfinite` (optim.cu)", "GradScaler.unscale_ / found_inf", "HBM / L2", "2 B / parameter"),
    ("foc_adam_step", "`k_adam_step` + `k_scaler_update` (optim.cu)", "fused Adam + `.to(half)` + `zeros_like` + GradScaler.update + LambdaLR", "HBM", "30 B / parameter, 12.26 M parameters"),
]


def load(name):
    try:
        txt = open(os.path.join(P, name)).read().strip().splitlines()[-1]
        return json.loads(txt)
    except Exception:
        return None


def f(v, n=2):
    return "—" if v is None else f"{v:,.{n}f}".replace(",", " ")


def status():
    o, r = load("r2_bench.json"), load("r2_bench_reference.json")
    rows = []
    rows.append(("train rays/s, configs[1], device-resident batches (`value`)", f"{f(r['value'], 0)} ({f(r['ms_per_step'])} ms / step)", f"**{f(o['value'], 0)}** ({f(o['ms_per_step'], 4)} ms / step)", f"{o['value'] / r['value']:.1f}×"))
    rows.append(("same, end to end from pinned host buffers (`e2e`)", f"{f(r['e2e']['value'], 0)}", f"{f(o['e2e']['value'], 0)} ({f(o['e2e']['ms_per_step'], 4)} ms / step)", f"{o['e2e']['value'] / r['e2e']['value']:.1f}×"))
    if o.get("e2e_pixels"):
        rows.append(("same, end to end from pixel indices (`e2e_pixels`, 32 KB / step up)", "—", f"{f(o['e2e_pixels']['value'], 0)} ({f(o['e2e_pixels']['ms_per_step'], 4)} ms / step)", "—"))
    q, qr = o.get("quality") or {}, r.get("quality") or {}
    rows.append(("PSNR after the 1024-step fixed-iteration run, 65 536 held-out rays, mean ± sd over seeds", f"{f(qr.get('psnr_db'))} ± {f(qr.get('psnr_sd_db'))} dB ({len(qr.get('seeds', []))} seeds)",
                 f"{f(q.get('psnr_db'))} ± {f(q.get('psnr_sd_db'))} dB ({len(q.get('seeds', []))} seeds)", f"Δ {abs(q.get('psnr_db', 0) - qr.get('psnr_db', 0)):.2f} dB"))
    ro, rr = o.get("render") or {}, r.get("render") or {}
    if ro and rr:
        rows.append(("800×800 render, configs[2], the reference arm's checkpoint in both arms", f"{f(rr['value'])} Mrays/s ({f(rr['ms_per_frame'])} ms / frame), coverage {rr.get('coverage'):.5f}",
                     f"{f(ro['value'])} Mrays/s ({f(ro['ms_per_frame'])} ms / frame), coverage {ro.get('coverage'):.5f}", f"{ro['value'] / rr['value']:.1f}×"))
    rw = o.get("render_windowed") or {}
    if rw and rr:
        rows.append(("same frame in depth windows (32 / 64 / 928 samples per ray, the reference loop's exact termination rule, speculative sizing per window: `render_windowed`)",
                     "(same reference loop)", f"{f(rw['value'])} Mrays/s ({f(rw['ms_per_frame'])} ms / frame), coverage {rw.get('coverage'):.5f}", f"{rw['value'] / rr['value']:.1f}×"))
    for key, label in (("multi_object", "4-object 640×480 composite, configs[3] (T = 512 per ray and object), full frame"), ("multi_object_edit8", "8-object composite with editable.py's offset on object 0, configs[4]")):
        mo, mr = o.get(key) or {}, r.get(key) or {}
        if mo and mr:
            rows.append((label, f"{f(mr['value'], 3)} Mrays/s ({f(mr['ms_per_frame'], 0)} ms / frame)", f"{f(mo['value'], 3)} Mrays/s ({f(mo['ms_per_frame'], 0)} ms / frame)", f"{mo['value'] / mr['value']:.1f}×"))
    t = o.get("train_262k")
    if t:
        rows.append(("configs[4] train leg, 32 768 rays per GPU and step", "—", f"{f(t['value'], 0)} rays/s ({f(t['ms_per_step'])} ms / step)", "—"))
    fr = o.get("frozen_bitfield")
    if fr:
        rows.append(("round 1's workload for continuity (frozen analytic occupancy, lr 1e-2, ≈ 157 k samples / step)", "267 609 (round 1)", f"{f(fr['value'], 0)} rays/s ({f(fr['ms_per_step'], 4)} ms / step)", "—"))
    c = o.get("cpu_baseline") or {}
    rows.append((f"CPU: oracle C port of the train step, {c.get('cores')} host cores", "—", f"{f(c.get('value'), 0)} rays/s", "—"))
    c = o.get("cpu_torch_renderer") or {}
    rows.append((f"CPU: configs[0], the PyTorch-only renderer restated on torch CPU ops, 4096 rays × 128 samples, {c.get('cores')} host threads", "—", f"{f(c.get('value'), 0)} rays/s ({f(c.get('ms_per_batch'), 0)} ms / batch)", "—"))
    print("| | reference's own Python + CUDA extensions (`bench.py --impl reference`) | this repo (`bench.py`) | ratio |\n|---|---|---|---|")
    for a in rows:
        print("| " + " | ".join(a) + " |")
    print(f"\n(samples per step: ours {f(o['config'].get('samples_per_step'), 0)}, reference arm {f((r.get('config') or {}).get('samples_per_step'), 0)}; "
          f"graph captures inside the timed region: {o['config'].get('graph_captures_in_timed_region')}; clocks {o['clocks'].get('sm_mhz')} / {o['clocks'].get('sm_max_mhz')} MHz, reasons {o['clocks'].get('reasons')})")


def kernels():
    o = load("r2_bench.json")
    rows = o["roofline"]["by_entry_point"]
    per = o["roofline"]["ms_per_step_by_entry_point"]
    print(f"(samples per step {f(o['roofline']['samples_per_step'], 0)}; timing: {o['roofline'].get('timing')})\n")
    print("| entry point → kernels | replaces (reference) | bound | alg. bytes / unit | µs / step (calls) | alg. traffic → GB/s (of HBM peak) | ncu DRAM MB / call | L2 MB / call → L2 GB/s | tensor pipe active | TFLOP/s (of burst peak) |\n|" + "---|" * 10)
    for ep, kern, rep, bound, alg in KERNELS:
        r = rows.get(ep) or rows.get(ep.replace("_rows", ""))
        if not r:
            continue
        us = per.get(ep, per.get(ep.replace("_rows", ""), 0.0)) * 1e3
        tf = f"{r['tensor_TFLOPs']:.0f} ({100 * r['tensor_frac_of_burst']:.1f} %)" if "tensor_TFLOPs" in r else "—"
        l2 = f"{r['ncu_l2_MB']:.0f} → {r['l2_GBs']:.0f}" if "l2_GBs" in r else "—"
        print(f"| `{ep}` → {kern} | {rep} | {bound} | {alg} | {us:.1f} ({r['calls_per_step']:g}) | {r.get('algorithmic_MB', 0):.0f} MB → {r['hbm_GBs']:.0f} ({100 * r['hbm_frac']:.0f} %) | "
              f"{r.get('ncu_dram_MB', float('nan')):.0f} | {l2} | {r.get('ncu_tensor_pipe_pct', 0):.1f} % | {tf} |")
    tot = sum(per.values()) * 1e3
    print(f"\nSum of the timed entry points: {tot:.0f} µs; the step itself is {o['ms_per_step'] * 1e3:.0f} µs. Not among them: the density refresh every 16 steps "
          f"({(o.get('density_refresh') or {}).get('partial_ms', 0) / 16 * 1e3:.0f} µs / step amortised), `torch.rand`, the input normalisation, the counter reset. The per-kernel "
          f"times come from a second capture of the same step with event-record nodes between the kernels, which lengthens each by the node's own latency "
          f"(≈ 1 µs) -- they are upper bounds that sum to slightly more than the untimed step.")


def scaling():
    try:
        WIN = json.load(open(os.path.join(P, 'r2_render_windowed.json')))['by_n']
    except Exception:
        WIN = {}
    base = load("r2_scale_n1.json") or load("r2_bench.json")     # the N = 1 run of the same lease when there is one
    print("| N | train rays/s (ms / step) | efficiency | e2e rays/s | render Mrays/s (ms / frame) | render speed-up | render, depth windows (ms / frame) | K = 4 composite Mrays/s (ms) | K = 8 edit (ms) | train 32 768 rays / GPU (ms / step) | dp_parity |\n|" + "---|" * 11)
    for n in (1, 2, 4, 8):
        d = base if n == 1 else load(f"r2_scale_n{n}.json")
        if not d:
            continue
        rd, m4, m8, big = d.get("render") or d.get("render_own_model") or {}, d.get("multi_object") or {}, d.get("multi_object_edit8") or {}, d.get("train_262k") or {}
        rb = base.get("render") or base.get("render_own_model") or {}
        print(f"| {n} | {f(d['value'] / 1e6)} M ({f(d['ms_per_step'], 4)}) | {100 * d['value'] / (n * base['value']):.1f} % | {f(d['e2e']['value'] / 1e6)} M | "
              f"{f(rd.get('value'), 1)} ({f(rd.get('ms_per_frame'))}) | {(rd.get('value', 0) / rb['value'] if rb else 0):.2f}× | {f((d.get('render_windowed') or {}).get('ms_per_frame') or WIN.get(str(n), {}).get('windowed_ms'))} | {f(m4.get('value'))} ({f(m4.get('ms_per_frame'), 1)}) | {f(m8.get('ms_per_frame'), 1)} | "
              f"{f((big.get('value') or 0) / 1e6)} M ({f(big.get('ms_per_step'))}) | {d.get('dp_parity')} |")
    for n in (2, 4, 8):
        d = load(f"r2_scale_n{n}.json")
        if d:
            per = d["roofline"]["ms_per_step_by_entry_point"]
            ex = {k: round(v * 1e3, 1) for k, v in per.items() if k in ("foc_adam_step_dp", "foc_peer_barrier", "foc_adam_nonfinite", "foc_dp_reduce", "foc_adam_step_peers")}
            print(f"\nN = {n}: exchange inside the step (µs / step, in-graph events): {ex}; samples / step {f(d['config'].get('samples_per_step'), 0)}")


if __name__ == "__main__":
    {"status": status, "kernels": kernels, "scaling": scaling}[sys.argv[1]]()
