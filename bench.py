"""Benchmark of the B200 ReSIDE forward path (contract: see the task statement / DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--arch enb0-hu] [--batch 64] [--impl ours|reference]

One "step" = one forward of `--arch` over a synthetic batch of `--batch` images PER GPU (weak scaling).
Default workload = BASELINE.json configs[1]: ENB0-HU, bf16 mode, batch 64, 228x304, one B200.
Prints ONE JSON line (rank 0).  `value` = depth maps/s with inputs resident in HBM (CUDA events, max over
ranks); `e2e` = the same through the public nn.Module call with pinned HOST input and the depth maps read
back to the host inside the timed region.  `--impl reference` times the CPU oracle port of the reference
(the reference's own Python cannot travel to the GPU box: /root/reference is absent there).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "depth maps/sec (228x304)"
CPU_SAMPLE_BATCH = 8          # the reference's own benchmark batch (ReSIDE/inference_benchmark.py:55)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--arch", default="enb0-hu")
    ap.add_argument("--batch", type=int, default=64, help="images per GPU per step")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(tensor=p["bf16_tflops_sustained"], tensor_burst=p["bf16_tflops"], hbm=p["hbm_gbs"],
                    source="measured")
    except Exception:  # noqa: BLE001
        return dict(tensor=1400.0, tensor_burst=1590.0, hbm=6650.0, source="fallback")


def recorded_traffic(arch, batch, alg_bytes=None):
    """dram__bytes_read + dram__bytes_write summed over the conv_umma_kernel launches of ONE step, from the committed
    ncu capture of this workload (profiles/r01_traffic.json, written by tools/ncu_summary.py); None if not captured."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            t = json.load(f)
        e = t.get(f"{arch}_b{batch}")
        return None if e is None else {"bytes_per_step": e["conv_umma_dram_bytes_per_step"], "unit": "B",
                                       "algorithmic_bytes_per_step": alg_bytes,
                                       "source": "ncu dram__bytes_read.sum+dram__bytes_write.sum, " + e["capture"]}
    except Exception:  # noqa: BLE001
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            if not (t0 - 0.05 <= t <= t1 + 0.15):
                continue
            f = [s.strip() for s in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except Exception:  # noqa: BLE001
                continue
            for nme, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference(arch, steps, warmup, batch=CPU_SAMPLE_BATCH):
    """The reference's CPU path for this workload: the oracle port (oracle/reside_cpu.py, pinned against the
    reference's own modules) on all host threads; each step = one batch-`batch` forward."""
    import torch
    from oracle import weights as W
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    sd = W.make_state_dict(arch, seed=1234, calib_batch=1)
    x = W.make_input(arch, batch, seed=0)
    with torch.no_grad():
        for _ in range(warmup):
            W.forward(sd, x, arch)
        t0 = time.perf_counter()
        for _ in range(steps):
            W.forward(sd, x, arch)
        dt = time.perf_counter() - t0
    return dict(value=batch * steps / dt, unit="depth maps/s", cores=threads, kind="port",
                sample=f"{steps} x batch-{batch} fp32 forwards of {arch} at {x.shape[2]}x{x.shape[3]} "
                       f"(oracle port of the reference modules, torch {torch.__version__} CPU, {threads} threads)",
                ms_per_step=dt / steps * 1e3)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    h, w = (228, 304) if args.arch.endswith("-hu") else (384, 384)
    r = cpu_reference(args.arch, max(1, args.steps), max(0, args.warmup))
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": "depth maps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": f"{args.arch} forward, batch {CPU_SAMPLE_BATCH} per step, {h}x{w}, CPU"},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "depth maps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def per_op_times(plan, reps=3):
    """CUDA-event time of every operator of the plan (single-op replays on the launching stream)."""
    import ctypes as C
    import torch
    from efficientdepthestimation_b200 import _cabi as A
    lib, st = A.lib(), torch.cuda.current_stream().cuda_stream
    base, size = C.addressof(plan.ops), C.sizeof(A.EdeOp)
    out = []
    for i in range(plan.n_ops):
        one = C.cast(base + i * size, C.POINTER(A.EdeOp))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        A.check(lib.ede_plan_run(one, 1, st))
        e0.record()
        for _ in range(reps):
            A.check(lib.ede_plan_run(one, 1, st))
        e1.record()
        e1.synchronize()
        out.append(e0.elapsed_time(e1) / reps)
    return out


def run_ours(args):
    import torch
    from efficientdepthestimation_b200 import _cabi as A
    from efficientdepthestimation_b200 import engine, factory, parallel
    rank, world, local = parallel.init_from_env()
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; efficientdepthestimation_b200 has no CPU path "
                           "(use --impl reference for the CPU baseline)")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    engine.set_precision(args.precision)
    h, w = factory.input_hw(args.arch)
    torch.manual_seed(1234)
    model = factory.build_model(args.arch).to(dev).eval()
    n = args.batch
    g = torch.Generator().manual_seed(rank)
    host = [torch.randn(n, 3, h, w, generator=g).pin_memory() for _ in range(2)]
    n_rot = 4                                   # rotate inputs so that they never sit in the 126 MB L2
    xs = [host[i % 2].to(dev) + 0.01 * i for i in range(n_rot)]
    with torch.no_grad():
        for i in range(max(3, args.warmup)):
            y = model(xs[i % n_rot])
    torch.cuda.synchronize()
    plan, _, _ = model.plan_for(xs[0])

    # ---- device-resident throughput ("value")
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    parallel.barrier()
    torch.cuda.synchronize()
    A.lib().ede_launch_count(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.profiler.start()           # `ncu --profile-from-start off` then sees exactly the timed region
    t0 = time.perf_counter()
    e0.record()
    with torch.no_grad():
        for i in range(args.steps):
            y = model(xs[i % n_rot])
    e1.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    torch.cuda.profiler.stop()
    launches = int(A.lib().ede_launch_count(0))
    parallel.barrier()
    ms = parallel.reduce_max(e0.elapsed_time(e1) / args.steps, dev)
    clocks = sampler.stop(t0, t1) if rank == 0 else None

    # ---- end to end through the public API: pinned host input -> model(x) -> depth maps back on the host.
    # Every step's H2D input copy and D2H result copy are inside the timed region; the upload of batch i+1 runs
    # on a copy stream while batch i computes (what any serving loop around model(x) does).
    out_host = torch.empty(tuple(y.shape), dtype=torch.float32).pin_memory()
    e2e_steps = max(3, min(args.steps, 10))
    main_s, copy_s = torch.cuda.current_stream(dev), torch.cuda.Stream(dev)
    dbuf = [torch.empty_like(xs[0]) for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    free = [torch.cuda.Event() for _ in range(2)]

    def upload(i):
        with torch.cuda.stream(copy_s):
            copy_s.wait_event(free[i % 2])
            dbuf[i % 2].copy_(host[i % 2], non_blocking=True)
            ready[i % 2].record(copy_s)

    def e2e_loop(k):
        for ev in free:
            ev.record(main_s)
        upload(0)
        for i in range(k):
            if i + 1 < k:
                upload(i + 1)
            main_s.wait_event(ready[i % 2])
            with torch.no_grad():
                yy = model(dbuf[i % 2])
            free[i % 2].record(main_s)
            out_host.copy_(yy, non_blocking=True)

    e2e_loop(2)
    parallel.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e2e_loop(e2e_steps)
    torch.cuda.synchronize()
    e2e_ms = parallel.reduce_max((time.perf_counter() - t0) / e2e_steps * 1e3, dev)
    parallel.barrier()
    if rank != 0:
        return

    # ---- roofline of the dominant kernel family (conv_umma_kernel) + the bandwidth-bound kernels
    peaks = measured_peaks()
    times = per_op_times(plan)
    # dense-conv family: the tcgen05 kernel, the tap gather that completes the low-resolution up-projections and
    # the CUDA-core stem conv (its FLOPs are part of plan.flops)
    conv_ms = sum(t for t, k in zip(times, plan.kinds) if k in (A.OP_CONV, A.OP_UPCONV, A.OP_STEM))
    n_conv = sum(1 for k in plan.kinds if k == A.OP_CONV)
    bw_kinds = (A.OP_DWCONV, A.OP_SE_SCALE, A.OP_RESIZE, A.OP_MAXPOOL, A.OP_ELTWISE, A.OP_TO_NHWC, A.OP_TO_NCHW)
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from profile_plan import op_bytes
    bw_ms = sum(t for t, k in zip(times, plan.kinds) if k in bw_kinds)
    bw_bytes = sum(op_bytes(k, s) for k, s in zip(plan.kinds, plan.structs) if k in bw_kinds)
    achieved = plan.flops / conv_ms / 1e9 if conv_ms > 0 else 0.0
    conv_alg_bytes = sum(op_bytes(k, s) for k, s in zip(plan.kinds, plan.structs) if k == A.OP_CONV)
    roofline = {"bound": "tensor", "kernel": "conv_umma_kernel", "achieved": round(achieved, 1),
                "peak": peaks["tensor"], "unit": "TFLOP/s", "frac": round(achieved / peaks["tensor"], 4),
                "traffic": recorded_traffic(args.arch, n, conv_alg_bytes), "peak_source": peaks["source"] + " bf16 sustained",
                "launches_per_step": n_conv, "share_of_step": round(conv_ms / sum(times), 3),
                "how": "algorithmic 2*MAC of all dense convs per step / summed CUDA-event time of the conv launches",
                "hbm_kernels": {"achieved": round(bw_bytes / bw_ms / 1e6, 1) if bw_ms else None,
                                "peak": peaks["hbm"], "unit": "GB/s",
                                "frac": round(bw_bytes / bw_ms / 1e6 / peaks["hbm"], 4) if bw_ms else None,
                                "share_of_step": round(bw_ms / sum(times), 3)}}
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        r = cpu_reference(args.arch, 3, 1)
        cpu = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
    total = n * world
    line = {"metric": METRIC, "value": round(total / ms * 1e3, 2), "unit": "depth maps/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": round(ms, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
            "config": {"workload": f"{args.arch} forward, batch {n} per GPU, {h}x{w}, random-init weights",
                       "global_batch": total, "precision": f"{args.precision} (encoder {engine.get_encoder_precision()})",
                       "parallelism": f"dp{world} (batch-sharded, no forward collective)",
                       "l2": f"inputs rotate over {n_rot} buffers ({n_rot * n * 3 * h * w * 4 >> 20} MiB) and the "
                             f"{plan.arena_bytes >> 20} MiB activation arena is rewritten every step (> 126 MB L2)"},
            "e2e": {"value": round(total / e2e_ms * 1e3, 2), "unit": "depth maps/s",
                    "h2d_bytes_per_step": n * 3 * h * w * 4, "d2h_bytes_per_step": int(y.numel()) * 4,
                    "ms_per_step": round(e2e_ms, 4)},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
            "algorithmic_tflops": round(plan.flops / ms / 1e9, 1)}
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        pass


if __name__ == "__main__":
    main()
