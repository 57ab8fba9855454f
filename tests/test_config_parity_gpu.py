"""-m gpu: parity at the batch sizes that are BENCHMARKED (BASELINE.json configs 2-5), not only at batch 1.

At batch 64 a CTA of conv_umma_kernel walks dozens of tiles (operand-ring phases flip many times, the two TMEM
accumulator stages alternate), dwconv blocks walk several images, and the arena is reused by liveness under
programmatic dependent launch -- none of which a batch-1 run exercises.  Every case compares the CUDA path with the
CPU oracle on the same seeded inputs, per image (max over n of sum|y-ref| / sum|ref|), tolerance 1e-2 in bf16 mode
(BASELINE.json north_star), and checks that row i of the big-batch output equals the batch-1 output of image i up
to the bf16 noise floor (the tile picker chooses other tile shapes per batch size), and that a replay is bit-identical.
"""
import pytest
import torch

from efficientdepthestimation_b200 import engine, factory
from oracle import reside_cpu as rc
from oracle import weights as W

pytestmark = pytest.mark.gpu
TOL_BF16 = 1e-2
_SD = {}


def state_dict(arch):
    if arch not in _SD:
        _SD[arch] = W.make_state_dict(arch, seed=1234)
    return _SD[arch]


def build(arch, lr_out_hw=None):
    m = factory.build_model(arch, lr_output_wh=(lr_out_hw[1], lr_out_hw[0]) if lr_out_hw else None)
    m.load_state_dict(state_dict(arch), strict=True)
    return m.cuda().eval()


def per_image_rel_l1(y, ref):
    num = (y - ref).abs().flatten(1).sum(1)
    den = ref.abs().flatten(1).sum(1)
    return num / den


@pytest.mark.parametrize("arch,batch", [("enb0-hu", 64), ("rn50-hu", 32), ("rn50-hu", 64), ("enb4-hu", 32),
                                        ("enb4-lr", 8)])
def test_benchmark_batch_matches_oracle_per_image(cuda, lib, arch, batch):
    engine.set_precision("bf16")
    x = W.make_input(arch, batch, seed=0)
    m = build(arch)
    with torch.no_grad():
        y = m(x.cuda())
        y_again = m(x.cuda())                    # second replay of the same plan (arena reuse, SE pools re-zeroed)
        y1 = torch.cat([m(x[i:i + 1].cuda()) for i in (0, batch // 2, batch - 1)])
    torch.cuda.synchronize()
    y, y_again, y1 = y.cpu(), y_again.cpu(), y1.cpu()
    with torch.no_grad():
        ref = W.forward(state_dict(arch), x, arch)
    assert y.shape == ref.shape
    assert torch.isfinite(y).all()
    err = per_image_rel_l1(y, ref)
    mt = rc.metrics_from_sums(rc.metric_sums(y, ref))
    print(f"\n{arch} batch {batch}: per-image rel L1 max {float(err.max()):.3e} mean {float(err.mean()):.3e}  "
          f"REL {mt['abs_rel']:.3e} DEL1 {mt['delta1']:.4f}")
    assert float(err.max()) <= TOL_BF16, (arch, batch, err)
    # the forward is bit-reproducible: the squeeze-excite sums are written per column strip and added in a fixed order
    # (with fp32 atomics there, two replays of ENB0-HU differed by 3.9e-3 -- a random net decorrelates its bf16 rounding)
    assert torch.equal(y_again, y), float(per_image_rel_l1(y_again, y).max())
    # batch independence: images do not leak into each other (eval-mode BN, per-image SE)
    rows = y[[0, batch // 2, batch - 1]]
    leak = per_image_rel_l1(rows, y1)
    print(f"    rows of the batch-{batch} output vs batch-1 outputs: rel L1 {leak.tolist()}")
    # the tile picker chooses other tile shapes / tap packings per batch size: fp32 sums in another order, which a
    # random net amplifies until its bf16 rounding is decorrelated (~sqrt(2) x the error against the oracle)
    assert float(leak.max()) <= TOL_BF16, leak
