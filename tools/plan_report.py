"""Host-only report of a model's forward plan: operator list, conv path (tcgen05 / direct), tile config,
useful-MAC fraction and arena size.  Needs no GPU (dry compile).
    python tools/plan_report.py enb0-hu 64 [--all]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from efficientdepthestimation_b200 import _cabi as A          # noqa: E402
from efficientdepthestimation_b200 import factory             # noqa: E402
from efficientdepthestimation_b200.engine import compile_module  # noqa: E402


def main():
    arch = sys.argv[1] if len(sys.argv) > 1 else "enb0-hu"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    m = factory.build_model(arch).eval()
    h, w = factory.input_hw(arch)
    x = torch.empty(n, 3, h, w, device="meta")
    plan, _, _ = compile_module(m, (x,), torch.device("cpu"), "bf16")
    print(f"{arch} N={n}: {plan.n_ops} ops, arena {plan.arena_bytes / 2**20:.1f} MiB, "
          f"algorithmic {plan.flops / n / 1e9:.2f} GFLOP/img")
    tot = 0.0
    for tag, st, cfg in plan.conv_configs():
        fl = 2.0 * st.n * st.oh * st.ow * st.cout_p * st.cin_p * st.kh * st.kw
        tot += fl
        if "--all" in sys.argv or fl / n > 2e8:
            path = ("umma/" + ("resident" if cfg.a_mode == 0 else "per-tap") + ("/swap" if cfg.swap else "")) if cfg.eligible else "DIRECT"
            print(f"  {tag:28s} {st.kh}x{st.kw}s{st.sh} {st.cin_p:5d}->{st.cout_p:5d} @{st.oh}x{st.ow} {path:14s} "
                  f"ck={cfg.ck} tile={cfg.bh}x{cfg.bw} mt={cfg.m_tiles} n_pad={cfg.n_pad}x{cfg.n_tiles} "
                  f"st={cfg.a_stages}/{cfg.b_stages} smem={cfg.smem_bytes // 1024}K frac={cfg.useful_frac:.2f} "
                  f"GF/img={fl / n / 1e9:.2f}")
    kinds = {}
    for k in plan.kinds:
        kinds[k] = kinds.get(k, 0) + 1
    print("  op kinds:", {A.OP_FIELD[k] + str(k): v for k, v in kinds.items()})


if __name__ == "__main__":
    main()
