"""Throughput of the two kernels either side of the forward path against the HBM roofline,
next to the reference's CPU code for the same step on this box's host cores.

    python tools/bench_pipeline.py [--batch 64] [--steps 20] [--warmup 5] [--no-cpu]

One JSON line per kernel: algorithmic bytes (read-once / write-once) per launch / CUDA-event time per launch, inputs
rotating over buffers larger than the 126 MB L2.  Also the whole frames -> depth_mm pipeline (H2D of uint8 frames,
preprocess, ENB0-HU forward, fused epilogue, D2H of the 16-bit maps) per step, inside the timed region.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from efficientdepthestimation_b200 import factory, loss as L, preprocess as PP  # noqa: E402


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback"


def timed(fn, steps, warmup):
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def raw_call(fn_name, struct):
    """Time the kernel, not the Python wrapper: prebuilt argument struct, bare ctypes call on the current stream."""
    import ctypes
    from efficientdepthestimation_b200 import _cabi as A
    f, st = getattr(A.lib(), fn_name), torch.cuda.current_stream().cuda_stream
    ref = ctypes.byref(struct)
    return lambda: f(ref, st)


def cpu_preprocess(frames, seconds=8.0):
    """The reference's per-frame chain (demo_transform.py: PIL resize / crop, ToTensor, per-channel Normalize)."""
    from PIL import Image
    mean, std = PP.IMAGENET_MEAN, PP.IMAGENET_STD
    t0, k = time.perf_counter(), 0
    while time.perf_counter() - t0 < seconds:
        img = Image.fromarray(frames[k % len(frames)]).resize((320, 240), Image.BILINEAR).crop((8, 6, 312, 234))
        t = torch.from_numpy(np.asarray(img).copy()).permute(2, 0, 1).contiguous().float().div(255)
        for c, m, s in zip(t, mean, std):
            c.sub_(m).div_(s)
        k += 1
    return k / (time.perf_counter() - t0)


def cpu_epilogue(pred, label, seconds=8.0):
    """test.py:49-54 + test_nyu.py:85-88 on CPU tensors, all host threads (oracle port)."""
    from oracle import reside_cpu as rc
    t0, k = time.perf_counter(), 0
    while time.perf_counter() - t0 < seconds:
        r = rc.post_resize(pred, label.shape[2:])
        rc.metric_sums(r, label)
        rc.depth_to_mm_u16(r)
        k += pred.shape[0]
    return k / (time.perf_counter() - t0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--band", type=int, default=0)
    ap.add_argument("--kernels-only", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    peak, src = peaks()
    n, rot = a.batch, 4
    g = torch.Generator().manual_seed(0)

    # ---- ede_preprocess_frames: 640x480 uint8 frames -> 3x228x304 fp32
    host_frames = torch.randint(0, 256, (n, 480, 640, 3), dtype=torch.uint8, generator=g).pin_memory()
    frames = [host_frames.to(dev).roll(i, 0) for i in range(rot)]
    pre = PP.FramePreprocessor()
    outs = [None]
    import ctypes
    from efficientdepthestimation_b200 import _cabi as A
    band = a.band or pre._pick_band(n)
    (xb, xk, yb, yk), xks, yks, span = pre._get_tables(480, 640, dev, band)
    y_pre = torch.empty(n, 3, 228, 304, device=dev)
    calls = [raw_call("ede_preprocess_frames", A.EdePreprocess(
        frames=f.data_ptr(), y=y_pre.data_ptr(), xb=xb.data_ptr(), xk=xk.data_ptr(), yb=yb.data_ptr(), yk=yk.data_ptr(),
        n=n, ih=480, iw=640, ch=228, cw=304, xks=xks, yks=yks, band=band, span=span,
        mean=(ctypes.c_float * 3)(*pre.mean), std=(ctypes.c_float * 3)(*pre.std))) for f in frames]
    ms = timed(lambda i: calls[i % rot](), a.steps, a.warmup)
    assert torch.equal(y_pre, pre(frames[(a.steps - 1) % rot]))
    by = n * (480 * 640 * 3 + 3 * 228 * 304 * 4)
    line = {"kernel": "preprocess_kernel", "band": band, "workload": f"{n} frames 640x480x3 u8 -> Scale(320x240) CenterCrop(304x228) normalise -> fp32 NCHW",
            "ms_per_launch": round(ms, 4), "frames_per_s": round(n / ms * 1e3, 1),
            "roofline": {"bound": "hbm", "achieved": round(by / ms / 1e6, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(by / ms / 1e6 / peak, 4), "peak_source": src, "algorithmic_bytes": by}}
    if not a.no_cpu:
        line["cpu_baseline"] = {"value": round(cpu_preprocess(host_frames.numpy()[:8]), 1), "unit": "frames/s", "cores": 1,
                                "kind": "port", "sample": "8 s of the reference's per-frame PIL + torch chain (single thread, as in its DataLoader with num_workers=0)"}
    print(json.dumps(line), flush=True)

    # ---- ede_depth_epilogue, test.py configuration: 114x152 -> 228x304, label, metric/loss sums + edge counts only
    preds = [(0.3 + 9.0 * torch.rand(n, 1, 114, 152, generator=g)).to(dev) for _ in range(rot)]
    labels = [(0.5 + 9.5 * torch.rand(n, 1, 228, 304, generator=g)).to(dev) for _ in range(rot)]
    big = [torch.empty(n, 1, 480, 640, device=dev) for _ in range(rot)]          # pushes L2 out between launches
    sums, edges = torch.zeros(16, dtype=torch.float64, device=dev), torch.zeros(4, dtype=torch.float64, device=dev)
    calls = [raw_call("ede_depth_epilogue", A.EdeDepthEpilogue(
        pred=preds[i].data_ptr(), label=labels[i].data_ptr(), resized=None, depth_mm=None, sums16=sums.data_ptr(),
        edges4=edges.data_ptr(), n=n, h=114, w=152, oh=228, ow=304, edge_thr=0.25, cap=10.0, mm_scale=1000.0)) for i in range(rot)]
    ms = timed(lambda i: calls[i % rot](), a.steps, a.warmup)
    by = n * (114 * 152 + 228 * 304) * 4
    line = {"kernel": "depth_epilogue_kernel", "workload": f"{n} maps 114x152 -> 228x304 + label: 13 loss/metric sums + 4 edge counts (test.py:49-76)",
            "ms_per_launch": round(ms, 4), "maps_per_s": round(n / ms * 1e3, 1),
            "roofline": {"bound": "hbm", "achieved": round(by / ms / 1e6, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(by / ms / 1e6 / peak, 4), "peak_source": src, "algorithmic_bytes": by,
                         "note": f"{by / 1e6:.1f} MB per launch fits the 126 MB L2 with the rotation; latency/launch bound at this size"}}
    if not a.no_cpu:
        line["cpu_baseline"] = {"value": round(cpu_epilogue(preds[0].cpu()[:8], labels[0].cpu()[:8]), 1), "unit": "maps/s",
                                "cores": torch.get_num_threads(), "kind": "port", "sample": "8 s of resize + metric sums + uint16 conversion on batches of 8 (oracle port, ATen CPU)"}
    print(json.dumps(line), flush=True)

    # ---- test_nyu.py configuration: 114x152 -> 480x640 uint16 millimetres
    mm = [torch.empty(n, 480, 640, dtype=torch.uint16, device=dev) for _ in range(rot)]
    calls = [raw_call("ede_depth_epilogue", A.EdeDepthEpilogue(
        pred=preds[i].data_ptr(), label=None, resized=None, depth_mm=mm[i].data_ptr(), sums16=None, edges4=None,
        n=n, h=114, w=152, oh=480, ow=640, edge_thr=0.25, cap=10.0, mm_scale=1000.0)) for i in range(rot)]
    ms = timed(lambda i: calls[i % rot](), a.steps, a.warmup)
    by = n * (114 * 152 * 4 + 480 * 640 * 2)
    print(json.dumps({"kernel": "depth_epilogue_kernel", "workload": f"{n} maps 114x152 -> 480x640 uint16 mm (test_nyu.py:85-88)",
                      "ms_per_launch": round(ms, 4), "maps_per_s": round(n / ms * 1e3, 1),
                      "roofline": {"bound": "hbm", "achieved": round(by / ms / 1e6, 1), "peak": peak, "unit": "GB/s",
                                   "frac": round(by / ms / 1e6 / peak, 4), "peak_source": src, "algorithmic_bytes": by}}), flush=True)
    del big
    if a.kernels_only:
        return

    # ---- whole pipeline from host frames to host 16-bit maps (inference_benchmark.py / test_nyu.py shape of work)
    model = factory.build_model("enb0-hu").to(dev).eval()
    out_host = torch.empty(n, 228, 304, dtype=torch.uint16).pin_memory()
    dbuf = torch.empty_like(frames[0])

    def step(i):
        dbuf.copy_(host_frames, non_blocking=True)
        with torch.no_grad():
            y = model(pre(dbuf))
        out_host.copy_(L.depth_epilogue(y, size=(228, 304), want_depth_mm=True)["depth_mm"], non_blocking=True)

    for i in range(3):
        step(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    k = max(3, min(a.steps, 10))
    for i in range(k):
        step(i)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / k * 1e3
    print(json.dumps({"pipeline": "uint8 frames (pinned host) -> H2D -> preprocess -> enb0-hu forward -> epilogue (228x304 uint16 mm) -> D2H",
                      "batch": n, "ms_per_step": round(ms, 3), "frames_per_s": round(n / ms * 1e3, 1),
                      "h2d_bytes_per_step": host_frames.numel(), "d2h_bytes_per_step": out_host.numel() * 2,
                      "note": "single stream, no copy/compute overlap"}), flush=True)


if __name__ == "__main__":
    main()
