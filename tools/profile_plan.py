"""Per-operator timing of a model's forward plan on the GPU (CUDA events around single-op replays).
    python tools/profile_plan.py rn50-hu 32 [--top 25] [--json out.json]
Every op is replayed `reps` times back to back between two events; activations of one op are larger
than L2 only for big batches, so use the bench batch sizes for numbers that mean something.
"""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from efficientdepthestimation_b200 import _cabi as A           # noqa: E402
from efficientdepthestimation_b200 import engine, factory      # noqa: E402


def op_bytes(kind, st):
    es = {A.F32: 4, A.BF16: 2, A.BF16X2: 4, A.F32_TO_BF16X2: 4}
    if kind == A.OP_CONV:
        # x_dtype is plain bf16 with doubled lanes for split inputs, so cin_p * 2 B is right either way
        return st.n * st.h * st.w * st.cin_p * es[st.x_dtype] + st.n * st.oh * st.ow * st.cout_p * es[st.y_dtype] * (
            2 if st.res else 1)
    if kind == A.OP_DWCONV:
        return st.n * (st.h * st.w_ + st.oh * st.ow) * st.c_p * es[st.dtype]
    if kind == A.OP_SE_SCALE:
        return 2 * st.n * st.hw * st.c_p * es[st.dtype]
    if kind == A.OP_RESIZE:
        return st.n * st.c_p * (st.h * st.w * es[st.x_dtype] + st.oh * st.ow * es[st.y_dtype])
    if kind == A.OP_MAXPOOL:
        return st.n * st.c_p * (st.h * st.w + st.oh * st.ow) * es[st.dtype]
    if kind == A.OP_ELTWISE:
        return st.n_pix * st.c_p * es[st.dtype] * (3 if st.x2 else 2)
    if kind in (A.OP_TO_NHWC, A.OP_TO_NCHW):
        return st.n * st.h * st.w * (st.c * 4 + st.c_p * es[st.dtype])
    if kind == A.OP_STEM:
        return st.n * (st.cin * st.h * st.w_ * 4 + st.oh * st.ow * st.cout_p * es[st.y_dtype])
    if kind == A.OP_TAPSUM:
        return st.n * st.h * st.w * (st.kh * st.kw * 4 + 4)
    if kind == A.OP_UPCONV:
        return st.n * st.cgrp * (st.h * st.w * st.kh * st.kw * 4 + st.oh * st.ow * es[st.y_dtype])
    return 0


def main():
    arch = sys.argv[1]
    n = int(sys.argv[2])
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 30
    reps = 5
    dev = torch.device("cuda:0")
    m = factory.build_model(arch).to(dev).eval()
    h, w = factory.input_hw(arch)
    x = torch.randn(n, 3, h, w, device=dev)
    with torch.no_grad():
        for _ in range(2):
            m(x)
    torch.cuda.synchronize()
    plan, _, _ = m.plan_for(x)
    lib = A.lib()
    st = torch.cuda.current_stream().cuda_stream
    rows = []
    op_size = C.sizeof(A.EdeOp)
    base = C.addressof(plan.ops)
    for i in range(plan.n_ops):
        one = C.cast(base + i * op_size, C.POINTER(A.EdeOp))
        A.check(lib.ede_plan_run(one, 1, st))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            A.check(lib.ede_plan_run(one, 1, st))
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        kind, s = plan.kinds[i], plan.structs[i]
        fl = 0.0
        desc = ""
        if kind == A.OP_CONV:
            cin_log = s.cin_p
            fl = 2.0 * s.n * s.oh * s.ow * s.cout_p * s.cin_p * s.kh * s.kw * max(1, s.w_passes)
            cfg = A.umma_config(s)
            desc = (f"{s.kh}x{s.kw}s{s.sh} {s.cin_p}->{s.cout_p}@{s.oh}x{s.ow} p{s.w_passes} "
                    f"{'res' if cfg.a_mode == 0 else 'tap'}{'/swap' if cfg.swap else ''} ck{cfg.ck} {cfg.bh}x{cfg.bw} mt{cfg.m_tiles} n{cfg.n_pad}x{cfg.n_tiles}"
                    if cfg.eligible else f"{s.kh}x{s.kw} DIRECT {s.cin_p}->{s.cout_p}")
        if kind == A.OP_STEM:
            fl = 2.0 * s.n * s.oh * s.ow * s.cout_p * s.cin * s.k * s.k
            desc = f"{s.k}x{s.k}s{s.s} {s.cin}->{s.cout_p}@{s.oh}x{s.ow} CUDA-core stem from NCHW fp32"
        if kind == A.OP_DWCONV:
            desc = f"{s.k}x{s.k}s{s.s} c{s.c_p} {s.h}x{s.w_}->{s.oh}x{s.ow} {'+pool' if s.pool_sum else ''} dt{s.dtype}"
        if kind == A.OP_UPCONV:
            desc = f"{s.kh}x{s.kw} taps, {s.cgrp} lanes, {s.h}x{s.w}->{s.oh}x{s.ow}"
        if kind == A.OP_SE_SCALE:
            desc = f"c{s.c_p} hw{s.hw} dt{s.dtype}"
        if kind == A.OP_RESIZE:
            desc = f"c{s.c_p} {s.h}x{s.w}->{s.oh}x{s.ow}"
        by = op_bytes(kind, s)
        rows.append(dict(i=i, tag=plan.tags[i], kind=A.OP_FIELD[kind], ms=ms, tflops=fl / ms / 1e9 if fl else 0.0,
                         gbs=by / ms / 1e6, desc=desc, flops=fl, bytes=by))
    total = sum(r["ms"] for r in rows)
    # whole-plan time
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        A.check(lib.ede_plan_run(plan.ops, plan.n_ops, st))
    e1.record()
    torch.cuda.synchronize()
    whole = e0.elapsed_time(e1) / reps
    print(f"{arch} N={n} precision={engine.get_precision()}/{engine.get_encoder_precision()}: sum of ops {total:.3f} ms, "
          f"whole plan {whole:.3f} ms = {n / whole * 1e3:.1f} img/s; issued tensor TFLOP/s overall "
          f"{sum(r['flops'] for r in rows) / whole / 1e9:.1f}; algorithmic {plan.flops / whole / 1e9:.1f}")
    bykind = {}
    for r in rows:
        bykind[r["kind"]] = bykind.get(r["kind"], 0.0) + r["ms"]
    print("  by kind:", {k: round(v, 3) for k, v in sorted(bykind.items(), key=lambda kv: -kv[1])})
    for r in sorted(rows, key=lambda r: -r["ms"])[:top]:
        print(f"  #{r['i']:3d} {r['tag']:26s} {r['kind']:8s} {r['ms']:8.3f} ms {100 * r['ms'] / total:5.1f}%  "
              f"{r['tflops']:7.1f} TF/s {r['gbs']:7.0f} GB/s  {r['desc']}")
    if "--json" in sys.argv:
        with open(sys.argv[sys.argv.index("--json") + 1], "w") as f:
            json.dump(dict(arch=arch, n=n, whole_ms=whole, rows=rows), f)


if __name__ == "__main__":
    main()
