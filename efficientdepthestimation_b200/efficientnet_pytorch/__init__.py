"""B200 stand-in for ``efficientnet-pytorch==0.6.3`` as the reference uses it.

The reference imports the third-party package (``ReSIDE/models/modules.py:8``,
``ReSIDE/models/lasinger2019.py:4``; pinned at ``requirements.txt:8``) and only touches:
``EfficientNet.from_name / from_pretrained``, ``net._blocks``, ``net.children()[:2]`` and
``list(block.modules())[-2].num_features`` (modules.py:166-198).  This module re-provides
exactly that surface with the SAME parameter names and shapes -- so released ``.pth`` files
load unchanged -- but the blocks carry no torch arithmetic: they *emit* sm_100a operators
(``_emit``) into the forward plan (engine.py).  Architecture rules follow the published
EfficientNet scaling rules of the 0.6.3 release: block strings, width /
depth rounding, BN eps 1e-3 / momentum 0.01, and the 0.6.x STATIC same-padding computed from
the nominal image size (asymmetric (0,1)/(1,2) pads on stride-2 convs).
"""
import math
import re

from torch import nn

from .. import _cabi as A
from ..engine import EdeModule

# width, depth, nominal resolution, dropout
_COEFFS = {f"efficientnet-b{i}": v for i, v in enumerate([
    (1.0, 1.0, 224, 0.2), (1.0, 1.1, 240, 0.2), (1.1, 1.2, 260, 0.3), (1.2, 1.4, 300, 0.3),
    (1.4, 1.8, 380, 0.4), (1.6, 2.2, 456, 0.4), (1.8, 2.6, 528, 0.5), (2.0, 3.1, 600, 0.5)])}
_BLOCK_STRINGS = ["r1_k3_s11_e1_i32_o16_se0.25", "r2_k3_s22_e6_i16_o24_se0.25", "r2_k5_s22_e6_i24_o40_se0.25",
                  "r3_k3_s22_e6_i40_o80_se0.25", "r3_k5_s11_e6_i80_o112_se0.25", "r4_k5_s22_e6_i112_o192_se0.25",
                  "r1_k3_s11_e6_i192_o320_se0.25"]
_BN = dict(eps=1e-3, momentum=0.01)


def _decode(s: str) -> dict:
    f = {m.group(1): m.group(2) for m in (re.match(r"([a-z]+)([0-9.]+)", p) for p in s.split("_"))}
    return dict(repeats=int(f["r"]), k=int(f["k"]), s=int(f["s"][0]), e=int(f["e"]), i=int(f["i"]), o=int(f["o"]),
                se=float(f["se"]))


def round_filters(filters, width, divisor=8):
    filters *= width
    new = max(divisor, int(filters + divisor / 2) // divisor * divisor)
    return int(new + divisor if new < 0.9 * filters else new)


def round_repeats(repeats, depth):
    return int(math.ceil(depth * repeats))


def _same_pad(k, s, image_size):
    out = math.ceil(image_size / s)
    total = max((out - 1) * s + (k - 1) + 1 - image_size, 0)
    return total // 2, total - total // 2


class Conv2dStaticSamePadding(nn.Conv2d):
    """Parameter holder; `static_pad` = (left, right, top, bottom) is consumed by the plan
    compiler as predicated halo loads / TMA zero fill -- no padded copy is ever made."""

    def __init__(self, cin, cout, kernel_size, image_size, stride=1, groups=1, bias=True):
        super().__init__(cin, cout, kernel_size, stride=stride, groups=groups, bias=bias)
        lo, hi = _same_pad(kernel_size, stride, image_size)
        self.static_pad = (lo, hi, lo, hi)
        self.static_padding = nn.ZeroPad2d(self.static_pad) if lo + hi else nn.Identity()

    def forward(self, x):
        raise RuntimeError("Conv2dStaticSamePadding is executed through its parent block's B200 plan")


class MemoryEfficientSwish(nn.Module):
    def forward(self, x):
        raise RuntimeError("swish is fused into the producing convolution on the B200 path")


class MBConvBlock(EdeModule, nn.Module):
    """expand 1x1+BN+swish -> depthwise kxk+BN+swish (+SE squeeze) -> SE FCs -> gate -> project 1x1+BN (+skip)."""

    def __init__(self, k, stride, expand, cin, cout, se_ratio, stage_first, image_size):
        super().__init__()
        self._cfg = dict(k=k, stride=stride, expand=expand, cin=cin, cout=cout)
        cexp = cin * expand
        if expand != 1:
            self._expand_conv = Conv2dStaticSamePadding(cin, cexp, 1, image_size, bias=False)
            self._bn0 = nn.BatchNorm2d(cexp, **_BN)
        self._depthwise_conv = Conv2dStaticSamePadding(cexp, cexp, k, image_size, stride=stride, groups=cexp,
                                                       bias=False)
        self._bn1 = nn.BatchNorm2d(cexp, **_BN)
        csq = max(1, int(cin * se_ratio))
        self._se_reduce = Conv2dStaticSamePadding(cexp, csq, 1, image_size)
        self._se_expand = Conv2dStaticSamePadding(csq, cexp, 1, image_size)
        self._project_conv = Conv2dStaticSamePadding(cexp, cout, 1, image_size, bias=False)
        self._bn2 = nn.BatchNorm2d(cout, **_BN)
        self._swish = MemoryEfficientSwish()
        # upstream decodes the stride of a stage's first block as the list [s]; `[1] == 1` is False,
        # so those blocks never take the identity skip (they always change width anyway)
        self._id_skip = (not stage_first) and stride == 1 and cin == cout

    def _emit(self, g, x, tag="mbconv"):
        h = x
        if self._cfg["expand"] != 1:
            # the expanded tensor only feeds the depthwise kernel (CUDA cores): in the split-precision scope it is written
            # as plain fp32 -- the same 4 bytes per lane -- so the conv epilogue does not split and the depthwise
            # kernel does not recombine (EDE_F32_TO_BF16X2)
            h = g.conv_bn_act(h, self._expand_conv, self._bn0, A.ACT_SWISH, tag=tag + ".expand",
                              out_dtype=A.F32 if g.expand_f32() else None, hi_lo=g.expand_f32())
        # where the gate is a separate multiply (late blocks), the depthwise output stays fp32 up to it as well
        dw = self._depthwise_conv
        s, k = dw.stride[0], dw.kernel_size[0]
        pl, pr, pt, pb = dw.static_pad
        oh, ow = (h.h + pt + pb - k) // s + 1, (h.w + pl + pr - k) // s + 1
        fold = g.se_fold_pays_dims(oh, ow, h.c_p, g.act_dtype, self._cfg["cout"])
        h, pool = g.dwconv_bn_act(h, dw, self._bn1, A.ACT_SWISH, pool=True, tag=tag + ".dw",
                                  out_f32=g.expand_f32() and not fold and h.dtype == A.F32)
        res = x if self._id_skip else None
        if fold:
            # (h * gate) . W == h . (gate * W): no pass over the expanded tensor, the gate rides in per-image weights
            h, fold = g.se_gate(h, pool, self._se_reduce, self._se_expand, tag=tag + ".se", fold=True)
            return g.conv_bn_act(h, self._project_conv, self._bn2, A.ACT_NONE, res=res, tag=tag + ".project", se_fold=fold)
        h = g.se_gate(h, pool, self._se_reduce, self._se_expand, tag=tag + ".se")
        return g.conv_bn_act(h, self._project_conv, self._bn2, A.ACT_NONE, res=res, tag=tag + ".project")


class EfficientNet(EdeModule, nn.Module):
    def __init__(self, name):
        super().__init__()
        width, depth, res, dropout = _COEFFS[name]
        stem = round_filters(32, width)
        self._conv_stem = Conv2dStaticSamePadding(3, stem, 3, res, stride=2, bias=False)
        self._bn0 = nn.BatchNorm2d(stem, **_BN)
        blocks = []
        for a in map(_decode, _BLOCK_STRINGS):
            i, o, r = round_filters(a["i"], width), round_filters(a["o"], width), round_repeats(a["repeats"], depth)
            blocks.append(MBConvBlock(a["k"], a["s"], a["e"], i, o, a["se"], True, res))
            blocks += [MBConvBlock(a["k"], 1, a["e"], o, o, a["se"], False, res) for _ in range(r - 1)]
        self._blocks = nn.ModuleList(blocks)
        head = round_filters(1280, width)
        self._conv_head = Conv2dStaticSamePadding(blocks[-1]._cfg["cout"], head, 1, res, bias=False)
        self._bn1 = nn.BatchNorm2d(head, **_BN)
        self._avg_pooling = nn.AdaptiveAvgPool2d(1)
        self._dropout = nn.Dropout(dropout)
        self._fc = nn.Linear(head, 1000)
        self._swish = MemoryEfficientSwish()

    @classmethod
    def from_name(cls, model_name, override_params=None):
        if model_name not in _COEFFS:
            raise ValueError("model_name should be one of: " + ", ".join(_COEFFS))
        return cls(model_name)

    @classmethod
    def from_pretrained(cls, model_name, **_kw):
        """Upstream downloads ImageNet weights here.  Every reference caller overwrites them right away
        with `load_state_dict` (test.py:18-22, reside_model.py:35-45), so offline this returns the
        random-init network instead of failing."""
        return cls.from_name(model_name)

    def _emit(self, g, x):
        raise RuntimeError("the classifier head of EfficientNet is outside the depth-estimation path; "
                           "use E_efficientnet / Encoder")
