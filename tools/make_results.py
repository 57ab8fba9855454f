"""Turn the bench lines of a GPU run (gpurun_out/<prefix>_bench_*.json, <prefix>_scale_*_n*.json) into the committed
summaries: profiles/r02_results.md, profiles/r02_scaling.md, and copies of the JSON lines under profiles/.
    python tools/make_results.py [prefix=r02b]"""
import glob
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")


def last_json(path):
    if not os.path.exists(path):
        return None
    for line in reversed(open(path).read().strip().splitlines()):
        line = line.strip()
        if line.startswith("{"):
            try:
                return json.loads(line)
            except ValueError:
                pass
    return None


def fmt(v, nd=0):
    return "-" if v is None else (f"{v:,.{nd}f}".replace(",", " "))


def main():
    prefix = sys.argv[1] if len(sys.argv) > 1 else "r02b"
    rows = []
    names = [("enb0hu_b64", "ENB0-HU bf16, batch 64 (BASELINE config 2, the default line)"),
             ("cfg3_n1", "RN50-HU bf16, batch 256 on one GPU (config 3 at N = 1)"),
             ("cfg4_n1", "ENB4-HU forward + Sobel loss + metrics, batch 32 (config 4 at N = 1)"),
             ("cfg5_n1", "ENB4-LR 384x384, batch 8 (config 5 per-GPU share)")]
    main_line = None
    for key, label in names:
        d = last_json(os.path.join(OUT, f"{prefix}_bench_{key}.json"))
        if d is None:
            continue
        shutil.copy(os.path.join(OUT, f"{prefix}_bench_{key}.json"), os.path.join(PROF, f"r02_bench_{key}.json"))
        if key == "enb0hu_b64":
            main_line = d
        r = d.get("roofline") or {}
        e = d.get("e2e") or {}
        rows.append((label, d["value"], d["ms_per_step"], e.get("value"), (e.get("u8_frames") or {}).get("value"),
                     r.get("frac"), r.get("exec_frac"), (r.get("hbm_kernels") or {}).get("frac")))
    ref = last_json(os.path.join(OUT, f"{prefix}_bench_reference_arm.json"))
    if ref is not None:
        shutil.copy(os.path.join(OUT, f"{prefix}_bench_reference_arm.json"), os.path.join(PROF, "r02_bench_reference_arm.json"))
    with open(os.path.join(PROF, "r02_results.md"), "w") as f:
        f.write("# Round-2 results (one B200, `tools/gpu_final_r02b.sh`; JSON lines: `profiles/r02_bench_*.json`)\n\n")
        f.write("| workload | `value` maps/s (device resident) | ms/step | `e2e` fp32 host | `e2e` uint8 frames | conv roofline frac "
                "(burst peak) | executed-MAC frac | HBM kernels frac |\n|---|---|---|---|---|---|---|---|\n")
        for label, v, ms, e1, e2, fr, ef, hf in rows:
            f.write(f"| {label} | {fmt(v)} | {ms:.3f} | {fmt(e1)} | {fmt(e2)} | {fmt(fr, 3)} | {fmt(ef, 3)} | {fmt(hf, 3)} |\n")
        if main_line is not None:
            rn = main_line.get("rn50_hu") or {}
            if rn:
                rr = rn.get("roofline") or {}
                f.write(f"| RN50-HU bf16, batch 64 (second half of the metric, same process) | {fmt(rn.get('value'))} | "
                        f"{rn.get('ms_per_step', 0):.3f} | {fmt((rn.get('e2e') or {}).get('value'))} | "
                        f"{fmt(((rn.get('e2e') or {}).get('u8_frames') or {}).get('value'))} | {fmt(rr.get('frac'), 3)} | "
                        f"{fmt(rr.get('exec_frac'), 3)} | {fmt((rr.get('hbm_kernels') or {}).get('frac'), 3)} |\n")
            lat = main_line.get("latency_b1") or {}
            if lat:
                f.write("\nBatch-1 latency (CUDA-graph replay): " + ", ".join(
                    f"{k} {v['ms_device']:.3f} ms device / {v['host_us_per_forward']:.0f} us host, {v['launches']} launches"
                    for k, v in lat.items()) + ".\n")
            cb = main_line.get("cpu_baseline") or {}
            if cb:
                legs = cb.get("legs") or {}
                f.write(f"\nCPU arm (oracle port of the reference modules, {cb.get('cores')} threads): "
                        + ", ".join(f"{k} {v['maps_per_s']:.1f} maps/s" for k, v in legs.items()) + ".\n")
            clk = main_line.get("clocks") or {}
            f.write(f"\nClocks during the timed region: {clk}.\n")
            pk = ((main_line.get('roofline') or {}).get('hbm_kernels') or {}).get('per_kernel') or {}
            if pk:
                f.write("\nBandwidth-bound kernel families (algorithmic bytes / in-plan time, ENB0-HU batch 64):\n\n"
                        "| family | GB/s | frac of HBM peak | ms per step | launches |\n|---|---|---|---|---|\n")
                for k, v in pk.items():
                    f.write(f"| {k} | {v['achieved']:.0f} | {v['frac']:.3f} | {v['ms_per_step']:.3f} | {v['launches_per_step']} |\n")
        if ref is not None:
            f.write(f"\nReference arm (`bench.py --impl reference`): {ref['value']:.1f} maps/s "
                    f"({(ref.get('cpu_baseline') or {}).get('cores')} threads, {ref['config']['workload']}).\n")
    # scaling
    sc = {}
    for path in sorted(glob.glob(os.path.join(OUT, f"{prefix}_scale_*_n*.json"))):
        d = last_json(path)
        if d is None:
            continue
        base = os.path.basename(path)[len(prefix) + 7:-5]            # <name>_n<N>
        name, n = base.rsplit("_n", 1)
        sc.setdefault(name, {})[int(n)] = d
    if sc:
        with open(os.path.join(PROF, "r02_scaling.jsonl"), "w") as fj, open(os.path.join(PROF, "r02_scaling.md"), "w") as f:
            f.write("# Round-2 multi-GPU lines (`tools/gpu_scale_r02b.sh`, one box, one process per GPU, NCCL for the timing "
                    "barrier and the 16 loss / metric doubles only)\n\n")
            f.write("| workload | N | `value` maps/s | x N=1 | `e2e` fp32 host | `e2e` uint8 frames | x N=1 (uint8 e2e) |\n|---|---|---|---|---|---|---|\n")
            one = {"enb0hu_b64": main_line,
                   "cfg4_enb4hu_b32_loss": last_json(os.path.join(OUT, f"{prefix}_bench_cfg4_n1.json")),
                   "cfg5_enb4lr_b8": last_json(os.path.join(OUT, f"{prefix}_bench_cfg5_n1.json"))}
            for name, per_n in sc.items():
                for n, d in sorted(per_n.items()):
                    fj.write(json.dumps({"name": name, **d}) + "\n")
                    b = one.get(name)
                    e = d.get("e2e") or {}
                    u8 = (e.get("u8_frames") or {}).get("value")
                    bu8 = ((b or {}).get("e2e") or {}).get("u8_frames", {}).get("value") if b else None
                    f.write(f"| {d['config']['workload'][:70]} | {n} | {fmt(d['value'])} | "
                            f"{(d['value'] / b['value']):.2f} |" if b else f"| {d['config']['workload'][:70]} | {n} | {fmt(d['value'])} | - |")
                    f.write(f" {fmt(e.get('value'))} | {fmt(u8)} | {(u8 / bu8):.2f} |\n" if (u8 and bu8) else f" {fmt(e.get('value'))} | {fmt(u8)} | - |\n")
                    rn = d.get("rn50_hu") or {}
                    if rn and b and b.get("rn50_hu"):
                        f.write(f"| rn50-hu forward, batch 64 per GPU (same process) | {n} | {fmt(rn['value'])} | "
                                f"{rn['value'] / b['rn50_hu']['value']:.2f} | {fmt((rn.get('e2e') or {}).get('value'))} | - | - |\n")
                    if d.get("loss"):
                        f.write(f"|   loss / REL / DEL1 after the all_reduce | {n} | {d['loss']} | | | | |\n")
    print(open(os.path.join(PROF, "r02_results.md")).read())
    if sc:
        print(open(os.path.join(PROF, "r02_scaling.md")).read())


if __name__ == "__main__":
    main()
