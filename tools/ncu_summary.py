"""Turn an ncu CSV launch list (gpu__time_duration.sum [+ dram__bytes_read.sum, dram__bytes_write.sum]) into a
per-kernel markdown table and, when DRAM bytes are present, into profiles/r01_traffic.json.
    python tools/ncu_summary.py gpurun_out/launches.csv profiles/r01_launches_enb0hu_b64.md --steps 2 \
        [--traffic-key enb0-hu_b64 --capture "<command line>"]"""
import collections
import csv
import json
import os
import sys


def main():
    src, dst = sys.argv[1], sys.argv[2]
    steps = int(sys.argv[sys.argv.index("--steps") + 1]) if "--steps" in sys.argv else 1
    key = sys.argv[sys.argv.index("--traffic-key") + 1] if "--traffic-key" in sys.argv else None
    capture = sys.argv[sys.argv.index("--capture") + 1] if "--capture" in sys.argv else ""
    rows = list(csv.reader(open(src)))
    start = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr = rows[start]
    ki, mi, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
    per = collections.OrderedDict()
    for r in rows[start + 1:]:
        if len(r) <= vi or not r[0].isdigit():
            continue
        d = per.setdefault(int(r[0]), {"name": r[ki]})
        d[r[mi]] = float(r[vi].replace(",", ""))
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for d in per.values():
        name = d["name"].split("(")[0].replace("void ", "").replace("ede::", "")
        name = name.split("<")[0] if name.startswith(("conv_umma_kernel", "conv_npack_kernel")) else name
        a = agg[name]
        a[0] += 1
        a[1] += d.get("gpu__time_duration.sum", 0.0)
        a[2] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    total = sum(a[1] for a in agg.values())
    unit_ns = total > 1e6          # ncu reports ns unless told otherwise
    scale = 1e-6 if unit_ns else 1e-3
    has_dram = any(a[2] for a in agg.values())
    with open(dst, "w") as f:
        f.write(f"# ncu launch list of `{capture}` (timed region only, {steps} step(s))\n")
        f.write("# --profile-from-start off --clock-control none: cold-cache, serialised -> compare SHARES, not absolutes\n")
        f.write(f"# {len(per)} launches, {total * scale:.3f} ms total\n")
        f.write("kernel | launches | ms | share" + (" | DRAM MB (read+write)" if has_dram else "") + "\n")
        for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{name} | {a[0]} | {a[1] * scale:.3f} | {100 * a[1] / total:.1f}%"
                    + (f" | {a[2] / 1e6:.1f}" if has_dram else "") + "\n")
    if key and has_dram:
        path = os.path.join(os.path.dirname(dst), "r02_traffic.json")
        t = json.load(open(path)) if os.path.exists(path) else {}
        conv = [sum(agg.get(k, [0, 0.0, 0.0])[i] for k in ("conv_umma_kernel", "conv_npack_kernel")) for i in range(3)]   # both tcgen05 conv kernels
        t[key] = {"conv_umma_dram_bytes_per_step": conv[2] / steps, "conv_umma_launches_per_step": conv[0] / steps,
                  "all_kernels_dram_bytes_per_step": sum(a[2] for a in agg.values()) / steps, "capture": capture}
        json.dump(t, open(path, "w"), indent=1)
    print(open(dst).read())


if __name__ == "__main__":
    main()
