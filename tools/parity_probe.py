"""Per-architecture parity (max over images of the relative L1 of the depth map vs the CPU oracle) under the current
environment knobs:  python tools/parity_probe.py [batch] [archs...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from efficientdepthestimation_b200 import factory      # noqa: E402
from oracle import reside_cpu as rc                      # noqa: E402
from oracle import weights as W                          # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    archs = sys.argv[2:] or ["enb0-hu", "rn50-hu", "enb4-hu", "enb4-lr"]
    for arch in archs:
        sd = W.make_state_dict(arch, seed=1234)
        x = W.make_input(arch, n, seed=0)
        m = factory.build_model(arch)
        m.load_state_dict(sd, strict=True)
        m = m.to("cuda:0").eval()
        with torch.no_grad():
            y = m(x.to("cuda:0")).cpu()
            ref = W.forward(sd, x, arch)
        errs = [rc.relative_l1(y[i:i + 1], ref[i:i + 1]) for i in range(n)]
        print(f"{arch} batch {n}: rel L1 max {max(errs):.3e} mean {sum(errs) / n:.3e}", flush=True)


if __name__ == "__main__":
    main()
