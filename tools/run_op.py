"""Replay ONE operator of a model plan (for ncu captures of a single kernel):
    python tools/run_op.py enb0-hu 64 D.up4.conv1+conv2 [reps]
Runs the whole plan twice first so that the op's inputs hold real activations."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from efficientdepthestimation_b200 import _cabi as A      # noqa: E402
from efficientdepthestimation_b200 import factory        # noqa: E402


def main():
    arch, n, tag = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
    dev = torch.device("cuda:0")
    m = factory.build_model(arch).to(dev).eval()
    h, w = factory.input_hw(arch)
    x = torch.randn(n, 3, h, w, device=dev)
    with torch.no_grad():
        for _ in range(2):
            m(x)
    torch.cuda.synchronize()
    plan, _, _ = m.plan_for(x)
    lib, st = A.lib(), torch.cuda.current_stream().cuda_stream
    base, size = C.addressof(plan.ops), C.sizeof(A.EdeOp)
    i = int(tag[1:]) if tag.startswith('#') else plan.tags.index(tag)      # '#21' = op number 21 of the plan
    one = C.cast(base + i * size, C.POINTER(A.EdeOp))
    torch.cuda.profiler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        A.check(lib.ede_plan_run(one, 1, st))
    e1.record()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print(f"{tag}: {e0.elapsed_time(e1) / reps * 1e3:.1f} us per launch")


if __name__ == "__main__":
    main()
