"""Import shim standing in for efficientnet-pytorch==0.6.3 (absent offline).

TEST INFRASTRUCTURE -- lets the real reference modules under /root/reference be
imported by tests/golden/make_golden.py.  Plain PyTorch CPU ops; module registration
order follows SURVEY.md Appendix A.2 so that the reference's `children()[:2]` slice
(ReSIDE/models/modules.py:185) and `list(block.modules())[-2].num_features`
(modules.py:198) resolve as they do with the real package.  "parity unpinned"
against upstream (package source unavailable).
"""
import torch
from torch import nn
from torch.nn import functional as F

from oracle import effnet_spec as spec


class MemoryEfficientSwish(nn.Module):
    def forward(self, x):
        return x * torch.sigmoid(x)


class Conv2dStaticSamePadding(nn.Conv2d):
    def __init__(self, cin, cout, kernel_size, image_size, stride=1, groups=1, bias=True):
        super().__init__(cin, cout, kernel_size, stride=stride, groups=groups, bias=bias)
        lo, hi = spec.static_same_pad(kernel_size, stride, image_size)
        self.static_padding = nn.ZeroPad2d((lo, hi, lo, hi)) if lo + hi > 0 else nn.Identity()

    def forward(self, x):
        x = self.static_padding(x)
        return F.conv2d(x, self.weight, self.bias, self.stride, 0, self.dilation, self.groups)


class MBConvBlock(nn.Module):
    def __init__(self, b: spec.BlockSpec, image_size: int):
        super().__init__()
        self._spec = b
        bn = dict(momentum=spec.BN_MOMENTUM, eps=spec.BN_EPS)
        if b.expand != 1:
            self._expand_conv = Conv2dStaticSamePadding(b.cin, b.cexp, 1, image_size, bias=False)
            self._bn0 = nn.BatchNorm2d(b.cexp, **bn)
        self._depthwise_conv = Conv2dStaticSamePadding(b.cexp, b.cexp, b.kernel, image_size,
                                                       stride=b.stride, groups=b.cexp, bias=False)
        self._bn1 = nn.BatchNorm2d(b.cexp, **bn)
        self._se_reduce = Conv2dStaticSamePadding(b.cexp, b.csq, 1, image_size)
        self._se_expand = Conv2dStaticSamePadding(b.csq, b.cexp, 1, image_size)
        self._project_conv = Conv2dStaticSamePadding(b.cexp, b.cout, 1, image_size, bias=False)
        self._bn2 = nn.BatchNorm2d(b.cout, **bn)
        self._swish = MemoryEfficientSwish()

    def forward(self, inputs, drop_connect_rate=None):
        h = inputs
        if self._spec.expand != 1:
            h = self._swish(self._bn0(self._expand_conv(h)))
        h = self._swish(self._bn1(self._depthwise_conv(h)))
        s = F.adaptive_avg_pool2d(h, 1)
        s = self._se_expand(self._swish(self._se_reduce(s)))
        h = torch.sigmoid(s) * h
        h = self._bn2(self._project_conv(h))
        if self._spec.skip:
            h = h + inputs
        return h


class EfficientNet(nn.Module):
    def __init__(self, name):
        super().__init__()
        stem, blocks, res = spec.block_specs(name)
        bn = dict(momentum=spec.BN_MOMENTUM, eps=spec.BN_EPS)
        self._conv_stem = Conv2dStaticSamePadding(3, stem, 3, res, stride=2, bias=False)
        self._bn0 = nn.BatchNorm2d(stem, **bn)
        self._blocks = nn.ModuleList([MBConvBlock(b, res) for b in blocks])
        head = spec.round_filters(1280, spec.MODEL_PARAMS[name][0])
        self._conv_head = Conv2dStaticSamePadding(blocks[-1].cout, head, 1, res, bias=False)
        self._bn1 = nn.BatchNorm2d(head, **bn)
        self._avg_pooling = nn.AdaptiveAvgPool2d(1)
        self._dropout = nn.Dropout(spec.MODEL_PARAMS[name][3])
        self._fc = nn.Linear(head, 1000)
        self._swish = MemoryEfficientSwish()

    @classmethod
    def from_name(cls, name, override_params=None):
        if name not in spec.MODEL_PARAMS:
            raise ValueError(f"model_name should be one of: {', '.join(spec.MODEL_PARAMS)}")
        return cls(name)

    @classmethod
    def from_pretrained(cls, name, **_kw):
        raise RuntimeError("efficientnet_pytorch shim: pretrained weights need the network")
