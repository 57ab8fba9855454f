"""-m gpu: whole-model parity of the B200 path against the CPU oracle.

Same seeded, conditioned state_dict is loaded (strict) into the product modules and fed to the
oracle; the parity metric is max_n sum|new-ref| / sum|ref| over the depth map.
Tolerances (BASELINE.json north_star): 1e-3 in fp32 mode, 1e-2 in bf16 mode.
The golden fixtures (tests/golden/*.pt, produced by the REAL reference modules) pin the oracle.
"""
import os

import pytest
import torch

from efficientdepthestimation_b200 import engine, factory
from oracle import reside_cpu as rc
from oracle import weights as W

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
TOL = {"fp32": 1e-3, "bf16": 1e-2}
_SD = {}


def state_dict(arch):
    if arch not in _SD:
        _SD[arch] = W.make_state_dict(arch, seed=1234)
    return _SD[arch]


def run_product(arch, sd, x, precision, lr_out_hw=None):
    m = factory.build_model(arch, lr_output_wh=(lr_out_hw[1], lr_out_hw[0]) if lr_out_hw else None)
    m.load_state_dict(sd, strict=True)
    m = m.cuda().eval()
    engine.set_precision(precision)
    try:
        with torch.no_grad():
            y = m(x.cuda())
        torch.cuda.synchronize()
    finally:
        engine.set_precision("bf16")
    return y.cpu()


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("arch", ["enb0-hu", "rn50-hu", "enb4-hu", "enb4-lr", "enb0-lr", "rn50-lr"])
def test_model_matches_oracle_and_golden(cuda, lib, arch, precision):
    sd = state_dict(arch)
    x = W.make_input(arch, 1, seed=0)
    y = run_product(arch, sd, x, precision)
    with torch.no_grad():
        y_ref = W.forward(sd, x, arch)
    assert y.shape == y_ref.shape and y.dtype == torch.float32 and y.is_contiguous()
    err = rc.relative_l1(y, y_ref)
    m = rc.metrics_from_sums(rc.metric_sums(y, y_ref))
    print(f"\n{arch} {precision}: rel_l1={err:.3e} REL={m['abs_rel']:.3e} DEL1={m['delta1']:.4f} "
          f"max_abs={float((y - y_ref).abs().max()):.3e}")
    assert err <= TOL[precision], (arch, precision, err)
    gold = torch.load(os.path.join(GOLDEN, f"{arch}.pt"))
    assert rc.relative_l1(y, gold["output"]) <= TOL[precision] * 1.05


@pytest.mark.parametrize("arch,n", [("enb0-hu", 3), ("rn50-hu", 2)])
def test_batch_and_odd_shapes(cuda, lib, arch, n):
    """Batch > 1 and the 224x320 shape of Benchmark/benchmark.py:682-700 (kernels are shape-generic)."""
    sd = state_dict(arch)
    x = W.make_input(arch, n, seed=3, hw=(224, 320))
    y = run_product(arch, sd, x, "bf16")
    with torch.no_grad():
        y_ref = W.forward(sd, x, arch)
    assert y.shape == y_ref.shape
    assert rc.relative_l1(y, y_ref) <= TOL["bf16"]


def test_unconditioned_default_init_smoke(cuda, lib):
    """Degenerate default-like init: output collapses to ~R.conv2.bias; must still agree."""
    sd = W.make_state_dict("enb0-hu", seed=5, condition=False)
    x = W.make_input("enb0-hu", 1, seed=1)
    y = run_product("enb0-hu", sd, x, "bf16")
    with torch.no_grad():
        y_ref = W.forward(sd, x, "enb0-hu")
    assert float((y - y_ref).abs().max()) <= 1e-2 * max(1.0, float(y_ref.abs().max()))


def test_forward_errors(cuda, lib):
    m = factory.build_model("enb0-hu")
    with pytest.raises(RuntimeError):
        m.train()(torch.zeros(1, 3, 64, 64, device="cuda"))          # training mode: no fallback
    with pytest.raises(RuntimeError):
        m.eval()(torch.zeros(1, 3, 64, 64))                          # CPU tensor: no fallback


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("arch", ["enb0-lr", "rn50-lr"])
def test_bottleneck_decoder_matches_oracle_and_golden(cuda, lib, arch, precision):
    """Decoder(block_type=BottleneckBlock), lasinger2019.py:166-211,252-262; golden from the real reference modules."""
    gold = torch.load(os.path.join(GOLDEN, f"{arch}-bottleneck.pt"))
    sd = W.make_state_dict(arch, seed=gold["seed"], lr_block="bottleneck")
    x = W.make_input(arch, 1, seed=gold["input_seed"])
    m = factory.build_model(arch, lr_block="bottleneck")
    m.load_state_dict(sd, strict=True)
    m = m.cuda().eval()
    engine.set_precision(precision)
    try:
        with torch.no_grad():
            y = m(x.cuda()).cpu()
    finally:
        engine.set_precision("bf16")
    err = rc.relative_l1(y, gold["output"])
    print(f"\n{arch} bottleneck decoder {precision}: rel_l1 vs reference golden {err:.3e}")
    assert err <= TOL[precision], (arch, precision, err)
