"""CPU emulation of reduced-precision operand rounding on the oracle graph (TEST INFRASTRUCTURE).

Answers, without GPU time, "which stages can run with bf16 tensor-core operands and still meet the
1e-2 relative-L1 bar?" by rounding every stored activation / dense-conv weight of a stage to bf16 or
tf32 (fp32 accumulate, as tcgen05 does) inside oracle.reside_cpu.hu_forward.
    python -m oracle.precision_study enb0-hu
"""
import itertools
import sys

import torch

from oracle import reside_cpu as rc
from oracle import weights as W


def _round(x, mode):
    if mode == "bf16":
        return x.to(torch.bfloat16).float()
    if mode == "tf32":      # keep 10 explicit mantissa bits (round to nearest even on the dropped 13)
        i = x.contiguous().view(torch.int32)
        r = ((i >> 13) & 1) + 0x0FFF
        return ((i + r) & ~0x1FFF).view(torch.float32)
    return x


class QCtx(rc.Ctx):
    def __init__(self, sd, prec):
        super().__init__(sd)
        self.prec = prec          # stage -> (activation mode, weight mode)

    def aq(self, x):
        return _round(x, self.prec.get(self.stage, ("fp32", "fp32"))[0])

    def wq(self, w):
        return _round(w, self.prec.get(self.stage, ("fp32", "fp32"))[1])


def study(arch="enb0-hu", batch=1):
    sd = W.make_state_dict(arch, seed=1234)
    x = W.make_input(arch, batch, seed=0)
    enc = W.encoder_name(arch)
    with torch.no_grad():
        ref = rc.hu_forward(sd, x, enc)
        stages = ("E", "D", "MFF", "R")
        rows = []
        for modes in itertools.product(("bf16", "tf32"), repeat=4):
            prec = {s: (m, m) for s, m in zip(stages, modes)}
            y = rc.hu_forward(sd, x, enc, ctx=QCtx(sd, prec))
            rows.append((rc.relative_l1(y, ref), modes))
        for only in stages:
            y = rc.hu_forward(sd, x, enc, ctx=QCtx(sd, {only: ("bf16", "bf16")}))
            print(f"only {only:4s} in bf16 (rest fp32): rel_l1 = {rc.relative_l1(y, ref):.3e}")
        for err, modes in sorted(rows):
            print(" ".join(f"{s}={m}" for s, m in zip(stages, modes)), f"rel_l1={err:.3e}")


if __name__ == "__main__":
    study(sys.argv[1] if len(sys.argv) > 1 else "enb0-hu")
