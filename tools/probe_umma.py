"""GPU probe for the tcgen05 convolution kernel: sweeps operand-feeding modes, swizzle widths
and descriptor conventions on small random problems and prints one JSON line per case.

    python tools/probe_umma.py            # driver: one subprocess per configuration
    python tools/probe_umma.py --child    # runs the cases under the current EDE_UMMA_* env
"""
import ctypes as C
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = [
    # name, n, h, w, cin, cout, k, s, pad(t,l,b,r), res, act
    ("1x1_64_64", 2, 20, 24, 64, 64, 1, 1, (0, 0, 0, 0), False, 1),
    ("1x1_256_128_res", 3, 15, 19, 256, 128, 1, 1, (0, 0, 0, 0), True, 1),
    ("1x1_24_144", 2, 57, 76, 24, 144, 1, 1, (0, 0, 0, 0), False, 2),
    ("1x1_2048_1024", 2, 8, 10, 2048, 1024, 1, 1, (0, 0, 0, 0), False, 1),
    ("3x3_64_64", 2, 29, 38, 64, 64, 3, 1, (1, 1, 1, 1), False, 1),
    ("3x3_16_16_res", 1, 114, 152, 16, 16, 3, 1, (1, 1, 1, 1), True, 1),
    ("5x5_128_128", 1, 57, 76, 128, 128, 5, 1, (2, 2, 2, 2), False, 1),
    ("5x5_74_74", 1, 114, 152, 74, 74, 5, 1, (2, 2, 2, 2), False, 1),
    ("5x5_320_32", 1, 114, 152, 320, 32, 5, 1, (2, 2, 2, 2), False, 1),
    ("5x5_128_1_bias", 2, 114, 152, 128, 1, 5, 1, (2, 2, 2, 2), False, 0),
    ("5x5_1024_1024", 1, 15, 19, 1024, 1024, 5, 1, (2, 2, 2, 2), False, 1),
    ("3x3s2_128_128", 2, 57, 76, 128, 128, 3, 2, (1, 1, 1, 1), False, 1),
    ("1x1s2_256_512", 2, 57, 76, 256, 512, 1, 2, (0, 0, 0, 0), False, 0),
    ("7x7s2_3_64", 2, 228, 304, 3, 64, 7, 2, (3, 3, 3, 3), False, 1),
    ("3x3s2_asym_3_32", 2, 228, 304, 3, 32, 3, 2, (0, 0, 1, 1), False, 0),
]

CONFIGS = [
    {"EDE_UMMA_MODE": "1"},
    {"EDE_UMMA_MODE": "1", "EDE_UMMA_CK": "32"},
    {"EDE_UMMA_MODE": "1", "EDE_UMMA_CK": "16"},
    {"EDE_UMMA_MODE": "0", "EDE_UMMA_BASE_OFFSET": "0"},
    {"EDE_UMMA_MODE": "0", "EDE_UMMA_BASE_OFFSET": "1"},
    {"EDE_UMMA_MODE": "0", "EDE_UMMA_BASE_OFFSET": "0", "EDE_UMMA_CK": "32"},
    {"EDE_UMMA_MODE": "0", "EDE_UMMA_BASE_OFFSET": "0", "EDE_UMMA_CK": "16"},
    {"EDE_UMMA_MODE": "0", "EDE_UMMA_BASE_OFFSET": "1", "EDE_UMMA_CK": "32"},
]


def child():
    import torch
    import torch.nn.functional as F
    from efficientdepthestimation_b200 import _cabi as A
    from efficientdepthestimation_b200 import prepack as P

    lib = A.lib()
    dev = torch.device("cuda:0")
    only = os.environ.get("PROBE_ONLY", "")
    for (name, n, h, w, cin, cout, k, s, pad, use_res, act) in CASES:
        if only and only not in name:
            continue
        g = torch.Generator().manual_seed(hash(name) % 1000)
        pt, pl, pb, pr = pad
        oh = (h + pt + pb - k) // s + 1
        ow = (w + pl + pr - k) // s + 1
        cin_p, cout_p = P.ceil8(cin), P.ceil8(cout)
        x = torch.randn(n, h, w, cin_p, generator=g).to(dev)
        x[..., cin:] = 0
        xb = x.to(torch.bfloat16).contiguous()
        wt = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).to(dev)
        scale = (0.5 + torch.rand(cout, generator=g)).to(dev)
        shift = (0.2 * torch.randn(cout, generator=g)).to(dev)
        full = P.scatter_weight(wt, P.identity_map(cin, cin_p), P.identity_map(cout, cout_p))
        w_direct = P.pack_direct(full)
        w_umma, co_p16, ci_p64 = P.pack_umma(full)
        sc = P.place_lanes(scale, P.identity_map(cout, cout_p))
        sf = P.place_lanes(shift, P.identity_map(cout, cout_p))
        res = None
        if use_res:
            res = torch.randn(n, oh, ow, cout_p, generator=g).to(dev).to(torch.bfloat16)
            res[..., cout:] = 0
        # reference: fp32 conv on bf16-rounded operands
        xr = xb.float()[..., :cin].permute(0, 3, 1, 2)
        wr = wt.to(torch.bfloat16).float()
        ref = F.conv2d(F.pad(xr, (pl, pr, pt, pb)), wr, None, s)
        ref = ref * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)
        if use_res:
            ref = ref + res.float()[..., :cout].permute(0, 3, 1, 2)
        if act == 1:
            ref = F.relu(ref)
        elif act == 2:
            ref = ref * torch.sigmoid(ref)
        ref = ref.permute(0, 2, 3, 1).contiguous()

        def run(force_direct):
            y = torch.full((n, oh, ow, cout_p), float("nan"), device=dev, dtype=torch.bfloat16)
            a = A.EdeConv(x=xb.data_ptr(), y=y.data_ptr(), res=res.data_ptr() if use_res else None,
                          scale=sc.data_ptr(), shift=sf.data_ptr(), w_umma=w_umma.data_ptr(),
                          w_direct=w_direct.data_ptr(), n=n, h=h, w=w, cin_p=cin_p, x_cs=cin_p, x_co=0,
                          oh=oh, ow=ow, cout_p=cout_p, y_cs=cout_p, y_co=0, res_cs=cout_p, res_co=0,
                          kh=k, kw=k, sh=s, sw=s, pt=pt, pl=pl, act=act, act_c_end=cout_p,
                          x_dtype=A.BF16, y_dtype=A.BF16, x_layout=A.NHWC, relu_in=0,
                          w_umma_cout_p=co_p16, w_umma_cin_p=ci_p64, force_direct=force_direct)
            cfg = A.umma_config(a)
            st = torch.cuda.current_stream().cuda_stream
            A.check(lib.ede_conv2d(C.byref(a), st), "ede_conv2d")
            torch.cuda.synchronize()
            t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
            t0.record()
            for _ in range(5):
                A.check(lib.ede_conv2d(C.byref(a), st), "ede_conv2d")
            t1.record(); torch.cuda.synchronize()
            return y.float(), cfg, t0.elapsed_time(t1) / 5

        rec = {"case": name, "env": {k_: v for k_, v in os.environ.items() if k_.startswith("EDE_UMMA")}}
        try:
            y, cfg, ms = run(0)
            err = (y[..., :cout] - ref).abs()
            tol = 0.02 + 0.01 * ref.abs()
            rec.update(ok=bool((err <= tol).all()) and not bool(torch.isnan(y).any()),
                       max_err=float(err.max()), nan=int(torch.isnan(y).sum()), pad_abs=float(y[..., cout:].abs().sum()),
                       eligible=cfg.eligible, a_mode=cfg.a_mode, ck=cfg.ck, bh=cfg.bh, bw=cfg.bw, pitch=cfg.pitch,
                       m_tiles=cfg.m_tiles, n_pad=cfg.n_pad, n_tiles=cfg.n_tiles, a_stages=cfg.a_stages,
                       b_stages=cfg.b_stages, smem=cfg.smem_bytes, tmem=cfg.tmem_cols, frac=round(cfg.useful_frac, 3),
                       ms=round(ms, 4), gflops=round(2.0 * n * oh * ow * cout * cin * k * k / ms / 1e6, 1))
            if os.environ.get("PROBE_DIRECT", "0") == "1":
                yd, _, msd = run(1)
                errd = (yd[..., :cout] - ref).abs()
                rec.update(direct_ok=bool((errd <= tol).all()), direct_max_err=float(errd.max()), direct_ms=round(msd, 4))
        except Exception as e:  # noqa: BLE001
            rec.update(ok=False, error=str(e)[:300])
            print(json.dumps(rec), flush=True)
            break               # the CUDA context is likely poisoned
        print(json.dumps(rec), flush=True)


def main():
    if "--child" in sys.argv:
        return child()
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    log = open(os.path.join(out, "probe_umma.jsonl"), "a")
    for i, cfg in enumerate(CONFIGS):
        env = dict(os.environ)
        env.update(cfg)
        env["PROBE_DIRECT"] = "1" if i == 0 else "0"
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"], env=env, capture_output=True,
                               text=True, timeout=240)
            txt, err, code = r.stdout, r.stderr[-2000:], r.returncode
        except subprocess.TimeoutExpired as e:
            txt = (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or "")
            err, code = "TIMEOUT", -9
        log.write(txt)
        log.write(json.dumps({"config": cfg, "returncode": code, "secs": round(time.time() - t0, 1),
                              "stderr": err if code != 0 else ""}) + "\n")
        log.flush()
        print(f"== config {cfg} rc={code}")
        for line in txt.splitlines():
            try:
                d = json.loads(line)
                print("  ", d["case"], "OK" if d.get("ok") else "FAIL", {k: d.get(k) for k in
                      ("max_err", "nan", "a_mode", "ck", "bh", "bw", "m_tiles", "n_pad", "ms", "gflops", "frac",
                       "direct_ok", "direct_ms", "error")})
            except Exception:  # noqa: BLE001
                print("  ", line[:200])
        if code != 0:
            print(err[-800:])


if __name__ == "__main__":
    main()
