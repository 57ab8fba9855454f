"""Is a forward plan deterministic?  Runs the same plan twice on the same input and reports where the two runs differ:
whole forward at full speed, then op by op with a device synchronize after every launch, hashing each op's output
buffer.  `EDE_PDL=0` in the environment disables programmatic dependent launch.
    python tests/replay_check.py enb0-hu 64
Lives under tests/ because it loads the oracle's conditioned weights (default-init EfficientNets collapse to a constant
output, which hides every difference)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from efficientdepthestimation_b200 import _cabi as A           # noqa: E402
from efficientdepthestimation_b200 import factory              # noqa: E402
from oracle import weights as W                                # noqa: E402


def rel(a, b):
    return float((a - b).abs().sum() / b.abs().sum())


def main():
    arch, n = sys.argv[1], int(sys.argv[2])
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m = factory.build_model(arch)
    m.load_state_dict(W.make_state_dict(arch, seed=1234, calib_batch=1), strict=True)
    m = m.to(dev).eval()
    x = W.make_input(arch, n, seed=0).to(dev)
    with torch.no_grad():
        ys = [m(x).clone() for _ in range(4)]
    torch.cuda.synchronize()
    print(f"{arch} N={n} PDL={os.environ.get('EDE_PDL', '1')}: full-speed replays vs first:",
          [f"{rel(y, ys[0]):.2e}" for y in ys[1:]])
    plan, _, _ = m.plan_for(x)
    lib, st = A.lib(), torch.cuda.current_stream().cuda_stream
    base, size = C.addressof(plan.ops), C.sizeof(A.EdeOp)
    arena = plan.arena
    a0 = plan.arena_base - arena.data_ptr()

    def run_ops(sync):
        sums = []
        for i in range(plan.n_ops):
            one = C.cast(base + i * size, C.POINTER(A.EdeOp))
            A.check(lib.ede_plan_run(one, 1, st))
            if sync:
                torch.cuda.synchronize()
                outs = [(f, o, nb) for f, o, nb in plan.op_bufs[i] if f in ("y", "dst", "rows")]
                sums.append([int(arena[a0 + o:a0 + o + nb].view(torch.int16).to(torch.int64).sum()) for _, o, nb in outs])
        torch.cuda.synchronize()
        return sums

    # point input / output at fixed tensors
    out = torch.empty(plan.outputs[0][1], device=dev)
    for inp in plan.inputs:
        for i, field in inp["targets"]:
            setattr(getattr(plan.ops[i].u, A.OP_FIELD[plan.kinds[i]]), field, x.data_ptr())
    plan.ops[plan.outputs[0][0]].u.layout.dst = out.data_ptr()
    a = run_ops(True)
    o1 = out.clone()
    b = run_ops(True)
    o2 = out.clone()
    print(f"  op-by-op with sync: outputs differ by {rel(o2, o1):.2e}")
    bad = [i for i in range(plan.n_ops) if a[i] != b[i]]
    print(f"  ops whose output checksum differs between the two synchronous runs: {len(bad)} of {plan.n_ops}")
    for i in bad[:12]:
        print(f"    #{i} {plan.tags[i]} ({A.OP_FIELD[plan.kinds[i]]})")
    run_ops(False)
    o3 = out.clone()
    print(f"  op-by-op without sync vs synchronous: {rel(o3, o1):.2e}")


if __name__ == "__main__":
    main()
