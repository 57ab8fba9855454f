"""Where does the bf16-mode error come from on a given input distribution?  (diagnostic, needs a GPU)
    python tests/precision_probe.py enb0-hu rand      # ImageNet-normalised uniform noise (Benchmark/reside_model.py)
    python tests/precision_probe.py enb4-hu randn 32"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from efficientdepthestimation_b200 import engine, factory      # noqa: E402
from oracle import weights as W                                # noqa: E402


def main():
    arch, dist = sys.argv[1], sys.argv[2]
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 4
    sd = W.make_state_dict(arch, seed=1234)
    if dist == "rand":
        g = torch.Generator().manual_seed(3)
        x = torch.rand(n, 3, 228, 304, generator=g)
        x = (x - torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)) / torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    else:
        x = W.make_input(arch, n, seed=0)
    with torch.no_grad():
        ref = W.forward(sd, x, arch)
    print(f"{arch} {dist} n={n}: ref mean {float(ref.mean()):.3f} std {float(ref.std()):.3f} min {float(ref.min()):.3f}")
    for mode, enc in (("fp32", None), ("bf16", "bf16x2"), ("bf16", "bf16")):
        engine.set_precision(mode, encoder=enc)
        m = factory.build_model(arch)
        m.load_state_dict(sd, strict=True)
        m = m.cuda().eval()
        with torch.no_grad():
            y = m(x.cuda()).cpu()
        e = (y - ref).abs().flatten(1).sum(1) / ref.abs().flatten(1).sum(1)
        print(f"  mode {mode}/{enc}: per-image rel L1 max {float(e.max()):.3e} mean {float(e.mean()):.3e}")
    engine.set_precision("bf16", encoder="bf16x2")


if __name__ == "__main__":
    main()
