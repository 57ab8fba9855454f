"""Where does a tcgen05 conv CTA spend its time?  Runs ONE conv op of a model plan with EDE_UMMA_PROF=1 and
summarises the per-CTA globaltimer stamps.   python tools/cta_timeline.py rn50-hu 32 R.conv0"""
import ctypes as C
import os
import sys

os.environ["EDE_UMMA_PROF"] = "1"
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from efficientdepthestimation_b200 import _cabi as A      # noqa: E402
from efficientdepthestimation_b200 import factory        # noqa: E402


def main():
    arch, n, tags = sys.argv[1], int(sys.argv[2]), sys.argv[3:]
    dev = torch.device("cuda:0")
    m = factory.build_model(arch).to(dev).eval()
    h, w = factory.input_hw(arch)
    x = torch.randn(n, 3, h, w, device=dev)
    with torch.no_grad():
        m(x)
    torch.cuda.synchronize()
    plan, _, _ = m.plan_for(x)
    lib, st = A.lib(), torch.cuda.current_stream().cuda_stream
    base, size = C.addressof(plan.ops), C.sizeof(A.EdeOp)
    for tag in tags:
        i = plan.tags.index(tag)
        one = C.cast(base + i * size, C.POINTER(A.EdeOp))
        for _ in range(2):
            A.check(lib.ede_plan_run(one, 1, st))
        torch.cuda.synchronize()
        buf = np.zeros((4096, 16), dtype=np.uint64)
        k = lib.ede_debug_umma_prof(buf.ctypes.data, 4096)
        t = buf[:k].astype(np.int64)
        cfg = A.umma_config(plan.structs[i])
        span = (t[:, 3].max() - t[:, 0].min()) / 1e3
        mma = (t[:, 2] - t[:, 1]) / 1e3
        clk = float(np.median(t[:, 12] / np.maximum(t[:, 2] - t[:, 1], 1)) * 1e3)   # SM cycles per us while this kernel ran
        print(f"   SM clock while the kernel ran: {clk:.0f} MHz; issuer {np.median(t[:, 12]) / max(1.0, float(np.median(t[:, 7]))):.0f} clk per tile")
        print(f"{tag}: {k} persistent CTAs x {np.median(t[:, 7]):.0f} tiles, kernel span {span:.1f} us; "
              f"mode={'res' if cfg.a_mode == 0 else 'tap'} ck={cfg.ck} tile={cfg.bh}x{cfg.bw} mt={cfg.m_tiles} "
              f"n_pad={cfg.n_pad}x{cfg.n_tiles} stages={cfg.a_stages}/{cfg.b_stages} smem={cfg.smem_bytes >> 10}K "
              f"tmem={cfg.tmem_cols}")
        print(f"   MMA warp busy span       median {np.median(mma):9.1f} us")
        print(f"   issuer waiting for weight/pixel stages median {np.median(t[:, 4]) / clk:9.1f} us")
        print(f"   issuer waiting for the resident pixel tile median {np.median(t[:, 13]) / clk:9.1f} us")
        print(f"   producer waiting for free weight stages median {np.median(t[:, 5]) / clk:9.1f} us")
        print(f"   issuer waiting for a drained accumulator median {np.median(t[:, 8]) / clk:9.1f} us")
        print(f"   swapped epilogue (warp 2): barrier {np.median(t[:, 9]) / clk:.1f} us, finish+store {np.median(t[:, 10]) / clk:.1f} us, "
              f"ld wait+stage {np.median(t[:, 11]) / clk:.1f} us")
        print(f"   epilogue (warp 2): tile decode + scale/shift {np.median(t[:, 14]) / clk:.1f} us, accumulator -> stores -> release {np.median(t[:, 15]) / clk:.1f} us")
        print(f"   epilogue waiting for MMA median {np.median(t[:, 6]) / clk:9.1f} us")
        print(f"   epilogue end - MMA end   median {np.median(t[:, 3] - t[:, 2]) / 1e3:9.1f} us")


if __name__ == "__main__":
    main()
