"""not gpu: the CPU oracle against the golden vectors produced by the REAL reference modules
(tests/golden/make_golden.py, run in the build container where /root/reference exists)."""
import os

import pytest
import torch

from oracle import reside_cpu as rc
from oracle import weights as W

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
# README.md:51-56 of the reference (checkpoint sizes) pin the structure; counts from SURVEY.md section 4
N_PARAMS = {"rn50-hu": 67569473, "enb0-hu": 5333583, "enb4-hu": 19635607,
            "rn50-lr": 40692641, "enb0-lr": 3834157, "enb4-lr": 17149801}
N_ENTRIES = {"rn50-hu": 488, "enb0-hu": 522, "enb4-hu": 868, "rn50-lr": 446, "enb0-lr": 480, "enb4-lr": 826}


@pytest.mark.parametrize("arch", ["enb0-hu", "enb0-lr", "rn50-hu"])
def test_oracle_forward_matches_reference_output(arch):
    g = torch.load(os.path.join(GOLDEN, f"{arch}.pt"))
    sd = W.make_state_dict(arch, seed=g["seed"])
    x = W.make_input(arch, 1, seed=g["input_seed"])
    with torch.no_grad():
        y = W.forward(sd, x, arch)
    assert y.shape == g["output"].shape
    # weights are regenerated here (BN calibration = one CPU forward): thread-count dependent summation order
    # moves the running stats in the last bits, hence 1e-4 rather than bit equality
    assert rc.relative_l1(y, g["output"]) <= 1e-4
    assert g["n_entries"] == len(sd) == N_ENTRIES[arch]
    assert g["n_params"] == N_PARAMS[arch]


@pytest.mark.parametrize("arch", sorted(N_PARAMS))
def test_state_dict_layout_counts(arch):
    lay = W.layout(arch)
    assert len(lay) == N_ENTRIES[arch]
    n = 0
    for key, shape, kind in lay:
        if kind in ("conv", "bias", "gamma", "beta"):
            k = 1
            for s in shape:
                k *= s
            n += k
    assert n == N_PARAMS[arch]


@pytest.mark.parametrize("holes", [False, True])
def test_sobel_loss_metrics_golden(holes):
    g = torch.load(os.path.join(GOLDEN, f"sobel_loss_metrics{'_holes' if holes else ''}.pt"))
    out, lab = g["output"], g["labels"]
    assert torch.allclose(rc.sobel(out), g["sobel"], atol=1e-5)
    if not holes:
        mine = rc.sobel_loss_terms(out, lab)
        for k, v in g["loss"].items():
            assert abs(float(mine[k]) - v) <= 1e-5 * max(1.0, abs(v)), k
    mine = rc.metrics_from_sums(rc.metric_sums(out, lab))
    for k, v in g["metrics"].items():
        assert (mine[k] != mine[k] and v != v) or abs(mine[k] - v) <= 2e-5 * max(1.0, abs(v)), k


def test_sobel_analytic():
    """Closed-form pins from ReSIDE/sobel.py:9-10: a horizontal ramp x has dx = -8 (interior), dy = 0."""
    ramp = torch.arange(12.0).repeat(9, 1).view(1, 1, 9, 12)
    g = rc.sobel(ramp)
    assert torch.all(g[0, 0, 1:-1, 1:-1] == -8.0) and torch.all(g[0, 1, 1:-1, 1:-1] == 0.0)
    const = torch.full((1, 1, 5, 5), 3.0)
    assert torch.all(rc.sobel(const)[0, :, 1:-1, 1:-1] == 0)


def test_effnet_spec_tables():
    from oracle import effnet_spec as spec
    stem, blocks, res = spec.block_specs("efficientnet-b0")
    assert (stem, len(blocks), res) == (32, 16, 224)
    stem, blocks, res = spec.block_specs("efficientnet-b4")
    assert (stem, len(blocks), res) == (48, 32, 380)
    assert spec.static_same_pad(3, 2, 224) == (0, 1) and spec.static_same_pad(5, 2, 380) == (1, 2)
    assert spec.static_same_pad(3, 1, 224) == (1, 1) and spec.static_same_pad(5, 1, 224) == (2, 2)
    with pytest.raises(RuntimeError):
        spec.split_indices("efficientnet-b9")


def test_oracle_bottleneck_decoder_matches_reference_output():
    """Decoder(block_type=BottleneckBlock) (lasinger2019.py:166-211): fixture from the real reference modules
    (tests/golden/make_golden_bottleneck.py)."""
    arch = "enb0-lr"
    gold = torch.load(os.path.join(GOLDEN, f"{arch}-bottleneck.pt"))
    sd = W.make_state_dict(arch, seed=gold["seed"], lr_block="bottleneck")
    assert len(sd) == gold["n_entries"] and "decoder.blocks.0.res_block1.conv3.weight" in sd
    with torch.no_grad():
        y = W.forward(sd, W.make_input(arch, 1, seed=gold["input_seed"]), arch)
    assert torch.allclose(y, gold["output"], rtol=1e-5, atol=1e-5)
