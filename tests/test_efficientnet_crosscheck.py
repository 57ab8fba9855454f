"""not gpu: the MBConv arithmetic of the oracle (oracle/reside_cpu.py::mbconv, restated from the published
efficientnet-pytorch 0.6.3 algorithm -- the package itself is not available offline) checked
against an INDEPENDENT implementation that is: torchvision.models.efficientnet.MBConv.  Same weights in both, stride-1
blocks (where torchvision's symmetric padding equals 0.6.3's static same padding), BatchNorm eps forced to 1e-3.
What this pins: expand 1x1 -> BN -> swish -> depthwise -> BN -> swish -> squeeze-excite (mean, reduce, swish, expand,
sigmoid, multiply; C_sq from the block INPUT channels) -> project 1x1 -> BN -> identity skip."""
from functools import partial

import pytest
import torch

from oracle import effnet_spec as spec
from oracle import reside_cpu as rc

tv = pytest.importorskip("torchvision.models.efficientnet")


def _randomise(m, g):
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            mod.weight.data = 0.5 + torch.rand(mod.num_features, generator=g)
            mod.bias.data = 0.2 * torch.randn(mod.num_features, generator=g)
            mod.running_mean = 0.1 * torch.randn(mod.num_features, generator=g)
            mod.running_var = 0.5 + torch.rand(mod.num_features, generator=g)
        elif isinstance(mod, torch.nn.Conv2d):
            mod.weight.data = torch.randn(mod.weight.shape, generator=g) / (mod.weight[0].numel() ** 0.5)
            if mod.bias is not None:
                mod.bias.data = 0.1 * torch.randn(mod.bias.shape, generator=g)


@pytest.mark.parametrize("kernel,expand,cin,cout", [(3, 1, 32, 16),      # block 0 of EfficientNet-B0: no expand conv
                                                    (3, 6, 24, 24),      # identity skip
                                                    (5, 6, 40, 40),
                                                    (5, 6, 80, 112),     # stage-first stride-1 block: no skip
                                                    (3, 6, 192, 320)])
def test_oracle_mbconv_equals_torchvision_mbconv(kernel, expand, cin, cout):
    g = torch.Generator().manual_seed(kernel * 100 + cin)
    cnf = tv.MBConvConfig(expand, kernel, 1, cin, cout, 1, 1.0, 1.0)
    ref = tv.MBConv(cnf, 0.0, partial(torch.nn.BatchNorm2d, eps=spec.BN_EPS)).eval()
    _randomise(ref, g)
    layers = list(ref.block.children())
    names = (["_expand_conv", "_bn0"] if expand != 1 else []) + ["_depthwise_conv", "_bn1"]
    sd, i = {}, 0
    for j in range(0, len(names), 2):                                   # Conv2dNormActivation = (conv, bn, SiLU)
        conv, bn = layers[i][0], layers[i][1]
        sd[f"b.{names[j]}.weight"] = conv.weight.data
        for k in ("weight", "bias", "running_mean", "running_var"):
            sd[f"b.{names[j + 1]}.{k}"] = getattr(bn, k).data
        i += 1
    se = layers[i]
    assert se.fc1.out_channels == max(1, cin // 4)                       # C_sq from the block input, as in 0.6.3
    sd["b._se_reduce.weight"], sd["b._se_reduce.bias"] = se.fc1.weight.data, se.fc1.bias.data
    sd["b._se_expand.weight"], sd["b._se_expand.bias"] = se.fc2.weight.data, se.fc2.bias.data
    proj = layers[i + 1]
    sd["b._project_conv.weight"] = proj[0].weight.data
    for k in ("weight", "bias", "running_mean", "running_var"):
        sd[f"b._bn2.{k}"] = getattr(proj[1], k).data
    # stage_first mirrors 0.6.3's `[1] == 1` quirk: the first block of a stage never takes the skip.  torchvision
    # takes it whenever stride == 1 and cin == cout, so the comparison uses non-stage-first blocks where cin == cout.
    b = spec.BlockSpec(kernel=kernel, stride=1, expand=expand, cin=cin, cout=cout, stage_first=False)
    x = torch.randn(2, cin, 19, 23, generator=g)
    with torch.no_grad():
        want = ref(x)
        got = rc.mbconv(rc.Ctx(sd), x, "b.", b, 224)
    assert got.shape == want.shape
    assert torch.allclose(got, want, rtol=1e-5, atol=1e-5), float((got - want).abs().max())
    assert ref.use_res_connect == b.skip
