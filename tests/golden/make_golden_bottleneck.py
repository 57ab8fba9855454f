"""Golden fixture for the LR decoder built with `block_type=BottleneckBlock` (ReSIDE/models/lasinger2019.py:166-211,
252-262), produced by the REAL reference modules.  Build container only (needs /root/reference):
    python tests/golden/make_golden_bottleneck.py
Loads the oracle's seeded state_dict STRICTLY into the reference MidasNet (pins the key layout of the bottleneck
decoder), runs the reference forward and asserts that oracle.reside_cpu reproduces it."""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle", "_shim"), "/root/reference", ROOT]

from oracle import weights as W              # noqa: E402

from ReSIDE.models.lasinger2019 import BottleneckBlock, Decoder, Encoder, MidasNet       # noqa: E402


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    for arch in ("enb0-lr", "rn50-lr"):
        sd = W.make_state_dict(arch, seed=4321, lr_block="bottleneck")
        h, w = W.input_hw(arch)
        e = Encoder(W.encoder_name(arch), pretrained=False)
        ref = MidasNet(e, Decoder(e.block_out_channels, num_features="auto", block_type=BottleneckBlock),
                       output_size=(w // 2, h // 2), input_size=(w, h))
        ref.load_state_dict(sd, strict=True)
        assert list(ref.state_dict().keys()) == list(sd.keys()), arch
        ref.eval()
        x = W.make_input(arch, 1, seed=0)
        with torch.no_grad():
            y_ref = ref(x)
            y_or = W.forward(sd, x, arch, lr_output_hw=(h // 2, w // 2))
        err = float((y_ref - y_or).abs().max())
        assert err <= 1e-5 * float(y_ref.abs().max()), (arch, err)
        torch.save({"arch": arch, "lr_block": "bottleneck", "seed": 4321, "input_seed": 0, "output": y_ref.clone(),
                    "n_entries": len(sd)}, os.path.join(HERE, f"{arch}-bottleneck.pt"))
        print(arch, "bottleneck decoder:", tuple(y_ref.shape), float(y_ref.min()), float(y_ref.mean()),
              float(y_ref.max()), "oracle max abs diff", err)


if __name__ == "__main__":
    main()
