"""Benchmark of the B200 ReSIDE forward path (contract: see the task statement / DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--arch enb0-hu] [--batch 64] [--impl ours|reference]
                    [--with-loss] [--quick]

One "step" = one forward of `--arch` over a synthetic batch of `--batch` images PER GPU (weak scaling).
Default workload = BASELINE.json configs[1]: ENB0-HU, bf16 mode, batch 64, 228x304, one B200.
Prints ONE JSON line (rank 0):
  value      depth maps/s with inputs resident in HBM (CUDA events on the launching stream, max over ranks)
  e2e        the same through the public nn.Module call with pinned HOST input (fp32 NCHW, what the reference's
             callers feed) and the depth maps read back to the host, copies inside the timed region;
             e2e.u8_frames = the same fed with uint8 640x480 frames through ede_preprocess_frames (4x fewer H2D bytes)
  roofline   dense-conv family (conv_umma_kernel + tap gathers + stem): algorithmic and executed TFLOP/s against the
             measured bf16 peak (burst unless the clocks record shows a power cap); hbm_kernels per kernel family,
             every op timed after an L2 flush so that no entry can exceed the HBM peak
  rn50_hu    value / e2e / roofline of RN50-HU batch 64 (the metric is "ENB0-HU & RN50-HU"); default workload, N=1 only
  latency_b1 batch-1 latency of both models (BASELINE config 1's GPU counterpart), CUDA-graph replay
  cpu_baseline  the reference's CPU path (oracle port) on the box's host cores: batch 1 and batch 8, all threads and
             one thread, median of >= 10 forwards (rank 0, N=1 only)
`--with-loss` runs BASELINE config 4 as written: forward + fused resize + Sobel loss + metric sums + ONE all_reduce of the
16 device sums per step.  `--impl reference` times the CPU oracle port of the reference (the reference's own Python
cannot travel to the GPU box: /root/reference is absent there).
"""
import argparse
import json
import os
import statistics
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
    ap.add_argument("--with-loss", action="store_true",
                    help="config 4: forward + resize + Sobel loss + metrics + 1 all_reduce of the 16 sums per step")
    ap.add_argument("--quick", action="store_true", help="skip the RN50-HU, batch-1 and CPU legs (profiling runs)")
    return ap.parse_args()


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(tensor=p["bf16_tflops_sustained"], tensor_burst=p["bf16_tflops"], hbm=p["hbm_gbs"],
                    source="MEASURED_PEAKS.json")
    except Exception:  # noqa: BLE001
        return dict(tensor=1400.0, tensor_burst=1590.0, hbm=6650.0, source="fallback of B200_PROFILING.md")


def recorded_traffic(arch, batch, alg_bytes=None):
    """dram__bytes_read + dram__bytes_write summed over the conv_umma_kernel launches of ONE step, from the committed
    ncu capture of this workload (profiles/*_traffic.json, written by tools/ncu_summary.py); None if not captured."""
    for name in ("r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                t = json.load(f)
            e = t.get(f"{arch}_b{batch}")
            if e is not None:
                return {"bytes_per_step": e["conv_umma_dram_bytes_per_step"], "unit": "B",
                        "algorithmic_bytes_per_step": alg_bytes,
                        "source": f"profiles/{name}: ncu dram__bytes_read.sum+dram__bytes_write.sum, " + e["capture"]}
        except Exception:  # noqa: BLE001
            continue
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
        sm, mx, reasons, power = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.rows:
            if not (t0 - 0.05 <= t <= t1 + 0.15):
                continue
            f = [s.strip() for s in line.split(",")]
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
                power.append(float(f[2]))
            except Exception:  # noqa: BLE001
                continue
            for nme, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power) if power else None}


# ------------------------------------------------------------------------------------------------ CPU legs
def _cpu_forward_times(sd, x, arch, threads, iters, warmup):
    import torch
    from oracle import weights as W
    torch.set_num_threads(threads)
    ts = []
    with torch.no_grad():
        for _ in range(warmup):
            W.forward(sd, x, arch)
        for _ in range(iters):
            t0 = time.perf_counter()
            W.forward(sd, x, arch)
            ts.append(time.perf_counter() - t0)
    return ts


def cpu_reference(arch, steps, warmup, batch=CPU_SAMPLE_BATCH, threads=None):
    """The reference's CPU path for this workload: the oracle port (oracle/reside_cpu.py, pinned bit for bit against
    the reference's own modules) on the host cores; each step = one batch-`batch` forward; median over the steps."""
    import torch
    from oracle import weights as W
    threads = threads or os.cpu_count() or 1
    sd = W.make_state_dict(arch, seed=1234, calib_batch=1)
    x = W.make_input(arch, batch, seed=0)
    ts = _cpu_forward_times(sd, x, arch, threads, max(1, steps), warmup)
    mean = sum(ts) / len(ts)                     # the arm's line is K timed steps / their total time, like the GPU arm
    return dict(value=batch / mean, unit="depth maps/s", cores=threads, kind="port",
                sample=f"{len(ts)} batch-{batch} fp32 forwards of {arch} at {x.shape[2]}x{x.shape[3]} after "
                       f"{warmup} warm-up (oracle port of the reference modules, torch {torch.__version__} CPU, "
                       f"{threads} threads); median {statistics.median(ts) * 1e3:.1f} ms per forward",
                ms_per_step=mean * 1e3)


def cpu_baseline_legs(arch):
    """batch 1 (config 1) and batch 8 (the reference's own benchmark batch), all host threads and
    one thread, 3 warm-up + median of >= 10 forwards (the 1-thread batch-8 leg: 1 + 3, it alone would take a minute)."""
    import torch
    from oracle import weights as W
    sd = W.make_state_dict(arch, seed=1234, calib_batch=1)
    allc = os.cpu_count() or 1
    legs = {}
    for batch, threads, iters, warm in ((8, allc, 10, 3), (1, allc, 10, 3), (1, 1, 10, 2), (8, 1, 3, 1)):
        x = W.make_input(arch, batch, seed=0)
        ts = _cpu_forward_times(sd, x, arch, threads, iters, warm)
        legs[f"batch{batch}_threads{threads if threads == 1 else 'all'}"] = {
            "maps_per_s": round(batch / statistics.median(ts), 3), "ms_per_forward": round(statistics.median(ts) * 1e3, 2),
            "threads": threads, "iters": iters}
    torch.set_num_threads(allc)
    head = legs["batch8_threadsall"]
    return {"value": head["maps_per_s"], "unit": "depth maps/s", "cores": allc, "kind": "port",
            "sample": f"{arch} fp32 forward at 228x304 on the host cores (oracle port of the reference modules, bit-identical "
                      f"to them; torch {torch.__version__}): value = batch 8 on {allc} threads, median of 10 after 3 warm-up",
            "legs": legs}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    h, w = (228, 304) if args.arch.endswith("-hu") else (384, 384)
    r = cpu_reference(args.arch, max(1, args.steps), max(0, args.warmup))
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": "depth maps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.arch} forward, batch {CPU_SAMPLE_BATCH} per step, {h}x{w}, CPU"},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "depth maps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU legs
_FLUSH = {}


def flush_l2(dev):
    """Write a buffer larger than the 126 MB L2, so the next kernel finds its operands in HBM."""
    import torch
    buf = _FLUSH.get(str(dev))
    if buf is None:
        buf = _FLUSH[str(dev)] = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    buf.fill_(1)


def per_op_times(plan, dev, reps=3):
    """CUDA-event time of every operator of the plan, each launch on a cold L2 (flush before every timed launch):
    a bandwidth-bound op then moves its algorithmic bytes from/to HBM, so bytes / time cannot exceed the HBM peak.
    (In the plan itself part of an op's input is still L2-resident from its producer, so these times are upper
    bounds; the sum is reported next to the whole-step time.)"""
    import ctypes as C
    import torch
    from efficientdepthestimation_b200 import _cabi as A
    lib, st = A.lib(), torch.cuda.current_stream().cuda_stream
    base, size = C.addressof(plan.ops), C.sizeof(A.EdeOp)
    # the plan's input / output pointers are patched per forward: point them at tensors that live through this function
    keep = []
    for inp in plan.inputs:
        t = torch.randn(inp["shape"], device=dev)
        keep.append(t)
        for i, field in inp["targets"]:
            setattr(getattr(plan.ops[i].u, A.OP_FIELD[plan.kinds[i]]), field, t.data_ptr())
    for i, shape in plan.outputs:
        t = torch.empty(shape, device=dev)
        keep.append(t)
        plan.ops[i].u.layout.dst = t.data_ptr()
    A.check(lib.ede_plan_run(plan.ops, plan.n_ops, st))      # every intermediate buffer holds real activations
    out = []
    for i in range(plan.n_ops):
        one = C.cast(base + i * size, C.POINTER(A.EdeOp))
        ts = []
        for _ in range(reps):
            flush_l2(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            A.check(lib.ede_plan_run(one, 1, st))
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        out.append(min(ts))
    return out


KIND_NAMES = {2: "dwconv_tma_kernel / dwconv_tile_kernel", 4: "se_scale_kernel", 5: "resize_rows_kernel / resize_kernel", 6: "maxpool_kernel", 7: "eltwise_kernel",
              8: "nchw_to_nhwc_kernel", 9: "nhwc_to_nchw_kernel", 11: "upconv_rows/cols_kernel", 12: "stem_conv_kernel"}


def roofline_of(plan, arch, n, dev, clocks, step_ms):
    """Dense-conv family against the tensor peak + every bandwidth-bound kernel family against the HBM peak."""
    from efficientdepthestimation_b200 import _cabi as A
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from profile_plan import op_bytes
    peaks = measured_peaks()
    times = per_op_times(plan, dev)
    capped = bool(clocks and "sw_power_cap" in (clocks.get("reasons") or []))
    peak = peaks["tensor"] if capped else peaks["tensor_burst"]
    conv_kinds = (A.OP_CONV, A.OP_UPCONV, A.OP_STEM, A.OP_TAPSUM)
    # Cold-L2 single launches are serialised upper bounds (in the plan part of an op's input is still on chip, and
    # programmatic dependent launch overlaps prologues): like an ncu launch list they give each kernel's SHARE of the
    # step; the in-plan time of a family = share x the measured step time.
    total_cold = sum(times)
    k_plan = step_ms / total_cold if total_cold > 0 else 1.0
    conv_cold_ms = sum(t for t, k in zip(times, plan.kinds) if k in conv_kinds)
    conv_ms = conv_cold_ms * k_plan
    umma_ms = sum(t for t, k in zip(times, plan.kinds) if k == A.OP_CONV) * k_plan
    achieved = plan.conv_flops / conv_ms / 1e9 if conv_ms > 0 else 0.0
    executed = plan.conv_exec_flops / conv_ms / 1e9 if conv_ms > 0 else 0.0
    achieved_cold = plan.conv_flops / conv_cold_ms / 1e9 if conv_cold_ms > 0 else 0.0
    conv_alg_bytes = sum(op_bytes(k, s) for k, s in zip(plan.kinds, plan.structs) if k == A.OP_CONV)
    fam = {}
    for t, k, s in zip(times, plan.kinds, plan.structs):
        if k in KIND_NAMES and k != A.OP_STEM:
            e = fam.setdefault(KIND_NAMES[k], {"ms": 0.0, "bytes": 0, "launches": 0})
            e["ms"] += t * k_plan
            e["bytes"] += op_bytes(k, s)
            e["launches"] += 1
    hbm = {}
    for name, e in fam.items():
        gbs = e["bytes"] / e["ms"] / 1e6 if e["ms"] > 0 else 0.0
        hbm[name] = {"achieved": round(gbs, 1), "frac": round(gbs / peaks["hbm"], 4), "ms_per_step": round(e["ms"], 4),
                     "achieved_cold_l2": round(gbs * k_plan, 1),
                     "launches_per_step": e["launches"], "algorithmic_bytes_per_step": e["bytes"]}
    bw_ms = sum(e["ms"] for e in fam.values())
    bw_bytes = sum(e["bytes"] for e in fam.values())
    total = sum(times)
    return {"bound": "tensor", "kernel": "conv_umma_kernel / conv_npack_kernel (+ tap gathers + stem: the reference's dense-conv FLOPs)",
            "achieved": round(achieved, 1), "peak": peak, "unit": "TFLOP/s", "frac": round(achieved / peak, 4),
            "peak_kind": "bf16 sustained (sw_power_cap seen)" if capped else "bf16 burst (no power cap in the clocks record)",
            "peak_source": peaks["source"],
            "frac_vs_sustained": round(achieved / peaks["tensor"], 4),
            "achieved_cold_l2": round(achieved_cold, 1), "frac_cold_l2": round(achieved_cold / peak, 4),
            "exec_tflops": round(executed, 1), "exec_frac": round(executed / peak, 4),
            "traffic": recorded_traffic(arch, n, conv_alg_bytes),
            "launches_per_step": sum(1 for k in plan.kinds if k == A.OP_CONV),
            "ms_per_step": round(conv_ms, 4), "conv_umma_ms_per_step": round(umma_ms, 4),
            "share_of_step": round(conv_cold_ms / total, 3),
            "sum_of_cold_l2_op_times_ms": round(total, 4), "step_ms": round(step_ms, 4),
            "how": "algorithmic 2*MAC of the dense convs per step (plan.conv_flops; exec_* = MACs actually issued, the "
                   "low-resolution up-projection path skips 45 % of ENB0-HU's) / in-plan time of those launches = their "
                   "share of the summed single-launch CUDA-event times (each op timed alone after an L2 flush) x the "
                   "measured step time; *_cold_l2 = the same against the serialised cold-L2 times themselves",
            "hbm_kernels": {"achieved": round(bw_bytes / bw_ms / 1e6, 1) if bw_ms else None, "peak": peaks["hbm"],
                            "unit": "GB/s", "frac": round(bw_bytes / bw_ms / 1e6 / peaks["hbm"], 4) if bw_ms else None,
                            "share_of_step": round(bw_ms / total, 3), "per_kernel": hbm,
                            "how": "algorithmic bytes (input + output elements x dtype size) / in-plan time (share of the "
                                   "cold-L2 single-launch times x step time); achieved_cold_l2 never exceeds the HBM peak"}}


def build_on(arch, dev, precision):
    import torch
    from efficientdepthestimation_b200 import engine, factory
    engine.set_precision(precision)
    torch.manual_seed(1234)
    return factory.build_model(arch).to(dev).eval()


def measure(arch, n, args, dev, rank, world, local, with_loss=False, sample_clocks=True):
    """value (device-resident), e2e (pinned fp32 host input -> host depth maps) and e2e from uint8 frames for one model."""
    import torch
    from efficientdepthestimation_b200 import _cabi as A
    from efficientdepthestimation_b200 import factory, loss, parallel, preprocess
    h, w = factory.input_hw(arch)
    model = build_on(arch, dev, args.precision)
    g = torch.Generator().manual_seed(rank)
    host = [torch.randn(n, 3, h, w, generator=g).pin_memory() for _ in range(2)]
    n_rot = 4                                   # rotate inputs so that they never sit in the 126 MB L2
    xs = [host[i % 2].to(dev) + 0.01 * i for i in range(n_rot)]
    labels = sums_dev = None
    if with_loss:                               # config 4: labels at 228x304 (metres, NYU range), resident like the inputs
        labels = [(0.5 + 9.5 * torch.rand(n, 1, h, w, generator=g)).to(dev) for _ in range(n_rot)]
        sums_dev = torch.zeros(16, dtype=torch.float64, device=dev)

    def step(i):
        y = model(xs[i % n_rot])
        if with_loss:
            sums_dev.zero_()
            loss.depth_epilogue(y, label=labels[i % n_rot], want_sums=True, sums=sums_dev)
            parallel.all_reduce_sums(sums_dev)                 # the ONE collective: 128 bytes over NCCL
        return y

    with torch.no_grad():
        for i in range(max(3, args.warmup)):
            y = step(i)
    torch.cuda.synchronize()
    plan, _, _ = model.plan_for(xs[0])

    sampler = ClockSampler(local)
    if rank == 0 and sample_clocks:
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
            y = step(i)
    e1.record()
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    torch.cuda.profiler.stop()
    launches = int(A.lib().ede_launch_count(0))
    parallel.barrier()
    ms = parallel.reduce_max(e0.elapsed_time(e1) / args.steps, dev)
    clocks = sampler.stop(t0, t1) if (rank == 0 and sample_clocks) else None
    loss_out = None
    if with_loss:
        s = sums_dev.cpu().tolist()
        loss_out = {"loss": loss.loss_from_sums(s)["loss"], "abs_rel": loss.metrics_from_sums(s)["abs_rel"],
                    "delta1": loss.metrics_from_sums(s)["delta1"], "pixels_all_ranks": s[12]}

    # ---- end to end through the public API: pinned host input -> model(x) -> depth maps back on the host.
    # Every step's H2D input copy and D2H result copy are inside the timed region; the upload of batch i+1 runs
    # on a copy stream while batch i computes (what any serving loop around model(x) does).
    out_host = torch.empty(tuple(y.shape), dtype=torch.float32).pin_memory()
    e2e_steps = max(3, min(args.steps, 10))
    main_s, copy_s = torch.cuda.current_stream(dev), torch.cuda.Stream(dev)
    ready = [torch.cuda.Event() for _ in range(2)]
    free = [torch.cuda.Event() for _ in range(2)]

    def e2e_loop(k, src, dbuf, pre):
        for ev in free:
            ev.record(main_s)

        def upload(i):
            with torch.cuda.stream(copy_s):
                copy_s.wait_event(free[i % 2])
                dbuf[i % 2].copy_(src[i % 2], non_blocking=True)
                ready[i % 2].record(copy_s)
        upload(0)
        for i in range(k):
            if i + 1 < k:
                upload(i + 1)
            main_s.wait_event(ready[i % 2])
            with torch.no_grad():
                xin = pre(dbuf[i % 2]) if pre is not None else dbuf[i % 2]
                yy = model(xin)
                if with_loss:
                    sums_dev.zero_()
                    loss.depth_epilogue(yy, label=labels[i % n_rot], want_sums=True, sums=sums_dev)
                    parallel.all_reduce_sums(sums_dev)
            free[i % 2].record(main_s)
            out_host.copy_(yy, non_blocking=True)

    def timed(src, dbuf, pre):
        e2e_loop(2, src, dbuf, pre)
        parallel.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e2e_loop(e2e_steps, src, dbuf, pre)
        torch.cuda.synchronize()
        v = parallel.reduce_max((time.perf_counter() - t0) / e2e_steps * 1e3, dev)
        parallel.barrier()
        return v

    e2e_ms = timed(host, [torch.empty_like(xs[0]) for _ in range(2)], None)
    e2e = {"value": round(n * world / e2e_ms * 1e3, 2), "unit": "depth maps/s", "h2d_bytes_per_step": n * 3 * h * w * 4,
           "d2h_bytes_per_step": int(y.numel()) * 4, "ms_per_step": round(e2e_ms, 4),
           "input": "pinned fp32 NCHW (what the reference's callers feed), upload of batch i+1 overlapped on a copy stream"}
    if arch.endswith("-hu"):
        # the frames -> maps pipeline: uint8 640x480 video frames -> ede_preprocess_frames (Scale 320x240, CenterCrop
        # 304x228, ToTensor, Normalize; bit-exact vs Pillow) -> forward: 4x fewer host->device bytes
        fr = [torch.randint(0, 256, (n, 480, 640, 3), dtype=torch.uint8, generator=g).pin_memory() for _ in range(2)]
        pre = preprocess.FramePreprocessor()
        u8_ms = timed(fr, [torch.empty(n, 480, 640, 3, dtype=torch.uint8, device=dev) for _ in range(2)], pre)
        e2e["u8_frames"] = {"value": round(n * world / u8_ms * 1e3, 2), "unit": "depth maps/s",
                            "h2d_bytes_per_step": n * 480 * 640 * 3, "d2h_bytes_per_step": int(y.numel()) * 4,
                            "ms_per_step": round(u8_ms, 4),
                            "input": "pinned uint8 640x480x3 frames -> ede_preprocess_frames -> forward"}
    return dict(model=model, plan=plan, ms=ms, launches=launches, clocks=clocks, e2e=e2e, y=y, hw=(h, w), n_rot=n_rot,
                loss=loss_out)


def latency_b1(arch, dev, precision, iters=200):
    """Batch-1 latency (BASELINE config 1's GPU counterpart): resident input -> depth map, CUDA-graph replay."""
    import torch
    from efficientdepthestimation_b200 import factory
    h, w = factory.input_hw(arch)
    model = build_on(arch, dev, precision)
    model.freeze_plans()
    x = torch.randn(1, 3, h, w, device=dev)
    with torch.no_grad():
        for _ in range(5):
            model(x)
        torch.cuda.synchronize()
        plan, _, _ = model.plan_for(x)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(iters):
            model(x)
        e1.record()
        t_host = (time.perf_counter() - t0) / iters * 1e6          # host time to enqueue one forward
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / iters * 1e3
    return {"ms_device": round(e0.elapsed_time(e1) / iters, 4), "ms_wall": round(wall, 4),
            "host_us_per_forward": round(t_host, 1), "cuda_graph": plan._cg is not None, "launches": plan.n_ops}


def run_ours(args):
    import torch
    from efficientdepthestimation_b200 import engine, parallel
    rank, world, local = parallel.init_from_env()
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; efficientdepthestimation_b200 has no CPU path "
                           "(use --impl reference for the CPU baseline)")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    n = args.batch
    r = measure(args.arch, n, args, dev, rank, world, local, with_loss=args.with_loss)
    default_workload = args.arch == "enb0-hu" and n == 64 and not args.with_loss and args.precision == "bf16"
    extra = {}
    if default_workload and not args.quick:
        # the metric is "ENB0-HU & RN50-HU": the second model rides on the same line (every rank runs it: weak scaling)
        r2 = measure("rn50-hu", 64, args, dev, rank, world, local, sample_clocks=False)
        if rank == 0:
            extra["rn50_hu"] = {"value": round(64 * world / r2["ms"] * 1e3, 2), "unit": "depth maps/s",
                                "ms_per_step": round(r2["ms"], 4), "e2e": r2["e2e"],
                                "config": {"workload": "rn50-hu forward, batch 64 per GPU, 228x304, random-init weights"},
                                "roofline": roofline_of(r2["plan"], "rn50-hu", 64, dev, r["clocks"], r2["ms"])}
        del r2
        torch.cuda.empty_cache()
    if rank != 0:
        return
    if default_workload and not args.quick and world == 1:
        extra["latency_b1"] = {a: latency_b1(a, dev, args.precision) for a in ("enb0-hu", "rn50-hu")}
    plan, ms, h, w = r["plan"], r["ms"], r["hw"][0], r["hw"][1]
    roofline = roofline_of(plan, args.arch, n, dev, r["clocks"], ms)
    cpu = None
    if world == 1 and not args.no_cpu_baseline and not args.quick:
        cpu = cpu_baseline_legs(args.arch)
    total = n * world
    what = "forward + Sobel loss + metrics, 1 all_reduce/step" if args.with_loss else "forward"
    line = {"metric": METRIC, "value": round(total / ms * 1e3, 2), "unit": "depth maps/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": round(ms, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
            "config": {"workload": f"{args.arch} {what}, batch {n} per GPU, {h}x{w}, random-init weights",
                       "global_batch": total, "precision": f"{args.precision} (encoder {engine.get_encoder_precision()})",
                       "parallelism": f"dp{world} (batch-sharded, no forward collective"
                                      + (", one 128-byte all_reduce of the loss/metric sums per step)" if args.with_loss else ")"),
                       "l2": f"inputs rotate over {r['n_rot']} buffers ({r['n_rot'] * n * 3 * h * w * 4 >> 20} MiB) and the "
                             f"{plan.arena_bytes >> 20} MiB activation arena is rewritten every step (> 126 MB L2)"},
            "e2e": r["e2e"], "gpu_launches": r["launches"], "clocks": r["clocks"], "roofline": roofline,
            "algorithmic_tflops": round(plan.flops / ms / 1e9, 1)}
    if r["loss"] is not None:
        line["loss"] = r["loss"]
    line.update(extra)
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
