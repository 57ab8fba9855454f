"""Seeded, conditioned random state_dicts with the reference's key layout.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  The key layouts restate what the
reference constructors register: ``ReSIDE/models/net.py:18-24``,
``modules.py:16-31,46-56,185-196,205-226,240-257,272-285``, ``resnet.py:59-73,98-138``,
``lasinger2019.py:107-111,138-149,225-226,266-277``.  ``tests/golden/make_golden.py``
loads them *strictly* into the real reference modules to pin the layout.

Conditioning: BN gamma~U(0.5,1.5), beta~N(0,0.2), running stats = batch
statistics of one calibration forward, head bias +3 so that depth maps are positive.
"""
import math
from typing import Dict, List, Tuple

import torch

from oracle import effnet_spec as spec
from oracle import reside_cpu as rc

ARCHS = ("rn50-hu", "enb0-hu", "enb4-hu", "rn50-lr", "enb0-lr", "enb4-lr")
_ENC = {"rn50": "resnet50", "enb0": "efficientnet-b0", "enb4": "efficientnet-b4"}


def encoder_name(arch: str) -> str:
    return _ENC[arch.split("-")[0]]


class _Layout:
    def __init__(self):
        self.entries: List[Tuple[str, tuple, str]] = []

    def conv(self, key, cout, cin, k, bias=False, groups=1):
        self.entries.append((key + ".weight", (cout, cin // groups, k, k), "conv"))
        if bias:
            self.entries.append((key + ".bias", (cout,), "bias"))

    def bn(self, key, c):
        for suffix, kind in (("weight", "gamma"), ("bias", "beta"), ("running_mean", "mean"),
                             ("running_var", "var")):
            self.entries.append((f"{key}.{suffix}", (c,), kind))
        self.entries.append((key + ".num_batches_tracked", (), "nbt"))


def _bottlenecks(L: _Layout, prefix_of_layer):
    inplanes = 64
    for k, (planes, blocks) in enumerate(zip((64, 128, 256, 512), (3, 4, 6, 3)), start=1):
        for i in range(blocks):
            p = f"{prefix_of_layer(k)}{i}."
            L.conv(p + "conv1", planes, inplanes, 1); L.bn(p + "bn1", planes)
            L.conv(p + "conv2", planes, planes, 3); L.bn(p + "bn2", planes)
            L.conv(p + "conv3", planes * 4, planes, 1); L.bn(p + "bn3", planes * 4)
            if i == 0:
                L.conv(p + "downsample.0", planes * 4, inplanes, 1); L.bn(p + "downsample.1", planes * 4)
            inplanes = planes * 4


def _effnet(L: _Layout, name: str, blocks_prefix: str) -> List[int]:
    stem, blocks, _ = spec.block_specs(name)
    L.conv(blocks_prefix + "0.0", stem, 3, 3); L.bn(blocks_prefix + "0.1", stem)
    for p, b in zip(rc.effnet_block_prefixes(name, blocks_prefix), blocks):
        if b.expand != 1:
            L.conv(p + "_expand_conv", b.cexp, b.cin, 1); L.bn(p + "_bn0", b.cexp)
        L.conv(p + "_depthwise_conv", b.cexp, b.cexp, b.kernel, groups=b.cexp); L.bn(p + "_bn1", b.cexp)
        L.conv(p + "_se_reduce", b.csq, b.cexp, 1, bias=True)
        L.conv(p + "_se_expand", b.cexp, b.csq, 1, bias=True)
        L.conv(p + "_project_conv", b.cout, b.cexp, 1); L.bn(p + "_bn2", b.cout)
    idx = spec.split_indices(name)
    return [blocks[i - 1].cout for i in idx[1:]]


def _up(L: _Layout, p: str, cin: int, cout: int):
    L.conv(p + "conv1", cout, cin, 5); L.bn(p + "bn1", cout)
    L.conv(p + "conv1_2", cout, cout, 3); L.bn(p + "bn1_2", cout)
    L.conv(p + "conv2", cout, cin, 5); L.bn(p + "bn2", cout)


def layout(arch: str, lr_block: str = "residual") -> List[Tuple[str, tuple, str]]:
    enc, dec = arch.split("-")
    L = _Layout()
    if dec == "hu":
        if enc == "rn50":
            L.conv("E.conv1", 64, 3, 7); L.bn("E.bn1", 64)
            _bottlenecks(L, lambda k: f"E.layer{k}.")
            ch = [256, 512, 1024, 2048]
        else:
            ch = _effnet(L, _ENC[enc], "E.blocks.")
        nf = ch[3]
        L.conv("D.conv", nf // 2, nf, 1); L.bn("D.bn", nf // 2)
        nf //= 2
        for i in range(1, 5):
            _up(L, f"D.up{i}.", nf, nf // 2)
            nf //= 2
        for i in range(4):
            _up(L, f"MFF.up{i + 1}.", ch[i], 16)
        L.conv("MFF.conv", 64, 64, 5); L.bn("MFF.bn", 64)
        r = 64 + ch[3] // 32
        L.conv("R.conv0", r, r, 5); L.bn("R.bn0", r)
        L.conv("R.conv1", r, r, 5); L.bn("R.bn1", r)
        L.conv("R.conv2", 1, r, 5, bias=True)
    else:
        if enc == "rn50":
            L.conv("encoder.blocks.0.0", 64, 3, 7); L.bn("encoder.blocks.0.1", 64)
            _bottlenecks(L, lambda k: "encoder.blocks.0.4." if k == 1 else f"encoder.blocks.{k - 1}.0.")
            ch = [256, 512, 1024, 2048]
        else:
            ch = _effnet(L, _ENC[enc], "encoder.blocks.")
        f = ch[0]
        for i, cin in enumerate(reversed(ch)):
            for rb, c_in in (("res_block1", cin), ("res_block2", f)):
                p = f"decoder.blocks.{i}.{rb}."
                if lr_block == "bottleneck":                    # BottleneckBlock, lasinger2019.py:169-195
                    mid = max(1, f // 4)
                    L.conv(p + "conv1", mid, c_in, 1); L.bn(p + "bn1", mid)
                    L.conv(p + "conv2", mid, mid, 3); L.bn(p + "bn2", mid)
                    L.conv(p + "conv3", f, mid, 1); L.bn(p + "bn3", f)
                else:
                    L.conv(p + "conv1", f, c_in, 3); L.bn(p + "bn1", f)
                    L.conv(p + "conv2", f, f, 3); L.bn(p + "bn2", f)
                if c_in != f:
                    L.conv(p + "down_sample.0", f, c_in, 1); L.bn(p + "down_sample.1", f)
        L.conv("decoder.conv1", 128, f, 3); L.bn("decoder.bn1", 128)
        L.conv("decoder.conv2", 32, 128, 3); L.bn("decoder.bn2", 32)
        L.conv("decoder.conv3", 1, 32, 1, bias=True)
    return L.entries


def head_bias_key(arch: str) -> str:
    return "R.conv2.bias" if arch.endswith("hu") else "decoder.conv3.bias"


def input_hw(arch: str) -> Tuple[int, int]:
    return (228, 304) if arch.endswith("hu") else (384, 384)


def forward(sd, x, arch: str, lr_output_hw=None, calibrate=False, taps=None):
    """Oracle forward for `arch` (dispatch to reside_cpu)."""
    if arch.endswith("hu"):
        return rc.hu_forward(sd, x, encoder_name(arch), calibrate, taps)
    if lr_output_hw is None:
        lr_output_hw = (x.shape[2] // 2, x.shape[3] // 2)
    return rc.lr_forward(sd, x, encoder_name(arch), lr_output_hw, calibrate, taps)


def make_state_dict(arch: str, seed: int = 1234, condition: bool = True,
                    calib_hw=None, calib_batch: int = 2, lr_block: str = "residual") -> Dict[str, torch.Tensor]:
    """Random weights for `arch`; deterministic for (arch, seed) on CPU.

    condition=False gives the degenerate default-like init (identity BN statistics)."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, torch.Tensor] = {}
    for key, shape, kind in layout(arch, lr_block):
        if kind == "conv":
            fan_in = shape[1] * shape[2] * shape[3]
            sd[key] = torch.randn(shape, generator=g) * math.sqrt(2.0 / fan_in)
        elif kind == "bias":
            sd[key] = torch.randn(shape, generator=g) * 0.1
        elif kind == "gamma":
            sd[key] = (0.5 + torch.rand(shape, generator=g)) if condition else torch.ones(shape)
        elif kind == "beta":
            sd[key] = (0.2 * torch.randn(shape, generator=g)) if condition else torch.zeros(shape)
        elif kind == "mean":
            sd[key] = torch.zeros(shape)
        elif kind == "var":
            sd[key] = torch.ones(shape)
        else:
            sd[key] = torch.tensor(0, dtype=torch.long)
    if condition:
        h, w = calib_hw or input_hw(arch)
        x = torch.randn(calib_batch, 3, h, w, generator=g)
        with torch.no_grad():
            y = forward(sd, x, arch, calibrate=True)
            # depth-like positive range: head bias +3, head weight scaled to std ~0.5
            hk = head_bias_key(arch)
            wk = hk[:-4] + "weight"
            sd[wk] = sd[wk] * float(0.5 / max(float(y.std()), 1e-6))
            y = forward(sd, x, arch)
            sd[hk] = sd[hk] + (3.0 - float(y.mean()))
    return sd


def make_input(arch: str, batch: int, seed: int = 0, hw=None) -> torch.Tensor:
    h, w = hw or input_hw(arch)
    return torch.randn(batch, 3, h, w, generator=torch.Generator().manual_seed(seed))


def make_labels(batch: int, h: int, w: int, seed: int = 1, holes: bool = False) -> torch.Tensor:
    """depth = 0.5 + 9.5*rand (metres, NYU range); holes=True injects ~1% NaN and ~1% zeros."""
    g = torch.Generator().manual_seed(seed)
    d = 0.5 + 9.5 * torch.rand(batch, 1, h, w, generator=g)
    if holes:
        r = torch.rand(batch, 1, h, w, generator=g)
        d[r < 0.01] = float("nan")
        d[(r >= 0.01) & (r < 0.02)] = 0.0
    return d
