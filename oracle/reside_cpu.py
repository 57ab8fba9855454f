"""Functional fp32 CPU restatement of the ReSIDE forward path (the parity oracle).

TEST INFRASTRUCTURE (see ``oracle/__init__.py``) -- state_dict in, tensor out, plain
``torch.nn.functional`` on CPU.  Every function cites the reference lines it follows
(paths relative to /root/reference).  Pinned against the reference's own modules by
``tests/golden/make_golden.py`` (run in the build container, where /root/reference
exists); the EfficientNet encoder arithmetic follows SURVEY.md Appendix A
(efficientnet-pytorch 0.6.3, source unavailable -> "parity unpinned" upstream).
"""
import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

from oracle import effnet_spec as spec

SD = Dict[str, torch.Tensor]


class Ctx:
    """state_dict accessor.  With ``calibrate=True`` every batch-norm uses batch
    statistics and writes them back as running stats (one train()-mode forward with
    momentum=None, SURVEY.md 8c weight conditioning)."""

    def __init__(self, sd: SD, calibrate: bool = False, taps: Optional[dict] = None):
        self.sd, self.calibrate, self.taps = sd, calibrate, taps

    def tap(self, name, t):
        if self.taps is not None:
            self.taps[name] = t.detach().clone()

    # Precision-study hooks (oracle/precision_study.py): identity in the oracle proper.  `aq` is applied
    # to every stored activation where it is consumed, `wq` to dense-conv weights.
    stage = ""

    def aq(self, x):
        return x

    def wq(self, w):
        return w


def _interp(x, size):
    # F.interpolate(..., mode='bilinear', align_corners=True): modules.py:34, lasinger2019.py:244,289
    return F.interpolate(x, size=tuple(int(s) for s in size), mode="bilinear", align_corners=True)


def _conv(c: Ctx, x, key, stride=1, padding=0, groups=1):
    w = c.sd[key + ".weight"]
    return F.conv2d(c.aq(x), c.wq(w) if groups == 1 else w, c.sd.get(key + ".bias"), stride, padding, 1, groups)


def _bn(c: Ctx, x, key, eps=1e-5):
    sd = c.sd
    if c.calibrate:
        mean = x.mean(dim=(0, 2, 3))
        var_b = x.var(dim=(0, 2, 3), unbiased=False)
        n = x.numel() / x.shape[1]
        sd[key + ".running_mean"] = mean.detach().clone()
        sd[key + ".running_var"] = (var_b * n / max(n - 1, 1)).detach().clone()
        return F.batch_norm(x, None, None, sd[key + ".weight"], sd[key + ".bias"], True, 0.0, eps)
    return F.batch_norm(x, sd[key + ".running_mean"], sd[key + ".running_var"],
                        sd[key + ".weight"], sd[key + ".bias"], False, 0.0, eps)


# ----------------------------------------------------------------------------- ResNet-50
def bottleneck(c: Ctx, x, p: str, stride: int, has_ds: bool):
    """ReSIDE/models/resnet.py:75-95 (stride on the 3x3, resnet.py:66)."""
    out = F.relu(_bn(c, _conv(c, x, p + "conv1"), p + "bn1"))
    out = F.relu(_bn(c, _conv(c, out, p + "conv2", stride, 1), p + "bn2"))
    out = _bn(c, _conv(c, out, p + "conv3"), p + "bn3")
    res = _bn(c, _conv(c, x, p + "downsample.0", stride), p + "downsample.1") if has_ds else x
    return F.relu(out + c.aq(res))


def resnet50_stages(c: Ctx, x, stem: str, layer_fmt: str) -> List[torch.Tensor]:
    """E_resnet.forward, ReSIDE/models/modules.py:58-69 over resnet.py:98-138.
    `stem` is the prefix of conv1/bn1, `layer_fmt.format(k)` the prefix of layer k."""
    x = F.relu(_bn(c, _conv(c, x, stem + "conv1", 2, 3), stem + "bn1"))
    x = F.max_pool2d(x, 3, 2, 1)                                      # resnet.py:107
    outs = []
    for k, blocks in enumerate([3, 4, 6, 3], start=1):
        for i in range(blocks):
            x = bottleneck(c, x, f"{layer_fmt.format(k)}{i}.", 2 if (i == 0 and k > 1) else 1, i == 0)
        outs.append(x)
        c.tap(f"enc{k}", x)
    return outs


# ----------------------------------------------------------------------------- EfficientNet
def _swish(x):
    return x * torch.sigmoid(x)


def _same_conv(c: Ctx, x, key, k, s, res, groups=1):
    """Conv2dStaticSamePadding of efficientnet-pytorch 0.6.3 (Appendix A.4)."""
    lo, hi = spec.static_same_pad(k, s, res)
    if lo + hi:
        x = F.pad(x, (lo, hi, lo, hi))
    return _conv(c, x, key, s, 0, groups)


def mbconv(c: Ctx, x, p: str, b: spec.BlockSpec, res: int):
    """MBConvBlock.forward of efficientnet-pytorch 0.6.3 (Appendix A.3)."""
    h = x
    if b.expand != 1:
        h = _swish(_bn(c, _same_conv(c, h, p + "_expand_conv", 1, 1, res), p + "_bn0", spec.BN_EPS))
    h = _swish(_bn(c, _same_conv(c, h, p + "_depthwise_conv", b.kernel, b.stride, res, b.cexp),
                   p + "_bn1", spec.BN_EPS))
    h = c.aq(h)
    s = F.adaptive_avg_pool2d(h, 1)
    s = F.conv2d(_swish(F.conv2d(s, c.sd[p + "_se_reduce.weight"], c.sd[p + "_se_reduce.bias"])),
                 c.sd[p + "_se_expand.weight"], c.sd[p + "_se_expand.bias"])
    h = torch.sigmoid(s) * h
    h = _bn(c, _same_conv(c, h, p + "_project_conv", 1, 1, res), p + "_bn2", spec.BN_EPS)
    return h + c.aq(x) if b.skip else h


def effnet_block_prefixes(name: str, blocks_prefix: str) -> List[str]:
    """state_dict prefix of every MBConv block under the reference's 4-way split
    (ReSIDE/models/modules.py:185-196): blocks.0 = Sequential(stem, bn0, MBConv...)."""
    idx = spec.split_indices(name)
    out = []
    for g in range(4):
        for j, _ in enumerate(range(idx[g], idx[g + 1])):
            out.append(f"{blocks_prefix}{g}.{j + 2 if g == 0 else j}.")
    return out


def efficientnet_stages(c: Ctx, x, name: str, blocks_prefix: str) -> List[torch.Tensor]:
    """E_efficientnet.forward, ReSIDE/models/modules.py:153-161 (== lasinger2019.py:44-52).
    NB: no swish between the stem BN and block 0 (modules.py:185 slices children()[:2])."""
    stem, blocks, res = spec.block_specs(name)
    idx = spec.split_indices(name)
    prefixes = effnet_block_prefixes(name, blocks_prefix)
    x = _bn(c, _same_conv(c, x, blocks_prefix + "0.0", 3, 2, res), blocks_prefix + "0.1", spec.BN_EPS)
    outs = []
    for i, b in enumerate(blocks):
        x = mbconv(c, x, prefixes[i], b, res)
        if i + 1 in idx[1:]:
            outs.append(x)
            c.tap(f"enc{len(outs)}", x)
    return outs


# ----------------------------------------------------------------------------- Hu decoder
def up_projection(c: Ctx, x, p: str, size):
    """_UpProjection.forward, ReSIDE/models/modules.py:33-41."""
    x = _interp(c.aq(x), size)
    b1 = F.relu(_bn(c, _conv(c, x, p + "conv1", 1, 2), p + "bn1"))
    b1 = _bn(c, _conv(c, b1, p + "conv1_2", 1, 1), p + "bn1_2")
    b2 = _bn(c, _conv(c, x, p + "conv2", 1, 2), p + "bn2")
    return F.relu(b1 + c.aq(b2))


def hu_D(c: Ctx, b1, b2, b3, b4, p="D."):
    """D.forward, ReSIDE/models/modules.py:228-235."""
    x = F.relu(_bn(c, _conv(c, b4, p + "conv"), p + "bn"))
    x = up_projection(c, x, p + "up1.", b3.shape[2:])
    x = up_projection(c, x, p + "up2.", b2.shape[2:])
    x = up_projection(c, x, p + "up3.", b1.shape[2:])
    return up_projection(c, x, p + "up4.", (b1.shape[2] * 2, b1.shape[3] * 2))


def hu_MFF(c: Ctx, b1, b2, b3, b4, size, p="MFF."):
    """MFF.forward, ReSIDE/models/modules.py:259-268."""
    m = [up_projection(c, b, f"{p}up{i}.", size) for i, b in enumerate((b1, b2, b3, b4), start=1)]
    return F.relu(_bn(c, _conv(c, torch.cat(m, 1), p + "conv", 1, 2), p + "bn"))


def hu_R(c: Ctx, x, p="R."):
    """R.forward, ReSIDE/models/modules.py:287-298."""
    x = F.relu(_bn(c, _conv(c, x, p + "conv0", 1, 2), p + "bn0"))
    x = F.relu(_bn(c, _conv(c, x, p + "conv1", 1, 2), p + "bn1"))
    return _conv(c, x, p + "conv2", 1, 2)


def hu_encoder(c: Ctx, x, encoder: str):
    if encoder == "resnet50":
        return resnet50_stages(c, x, "E.", "E.layer{}.")
    return efficientnet_stages(c, x, encoder, "E.blocks.")


def hu_forward(sd: SD, x, encoder: str, calibrate=False, taps=None, ctx=None):
    """model.forward, ReSIDE/models/net.py:26-32."""
    c = ctx if ctx is not None else Ctx(sd, calibrate, taps)
    c.stage = "E"
    b1, b2, b3, b4 = hu_encoder(c, x, encoder)
    c.stage = "D"
    d = hu_D(c, b1, b2, b3, b4)
    c.tap("D", d)
    c.stage = "MFF"
    m = hu_MFF(c, b1, b2, b3, b4, d.shape[2:])
    c.tap("MFF", m)
    c.stage = "R"
    return hu_R(c, torch.cat((d, m), 1))


# ----------------------------------------------------------------------------- LR (MiDaS-style) model
def lr_residual_block(c: Ctx, x, p: str, has_ds: bool):
    """ResidualBlock.forward, ReSIDE/models/lasinger2019.py:151-163: pre-activation ReLU on
    the conv path, the skip uses the RAW x."""
    out = F.relu(x)
    out = F.relu(_bn(c, _conv(c, out, p + "conv1", 1, 1), p + "bn1"))
    out = _bn(c, _conv(c, out, p + "conv2", 1, 1), p + "bn2")
    skip = _bn(c, _conv(c, x, p + "down_sample.0"), p + "down_sample.1") if has_ds else x
    return out + skip


def lr_bottleneck_block(c: Ctx, x, p: str, has_ds: bool):
    """BottleneckBlock.forward, ReSIDE/models/lasinger2019.py:197-211."""
    out = F.relu(x)
    out = F.relu(_bn(c, _conv(c, out, p + "conv1"), p + "bn1"))
    out = F.relu(_bn(c, _conv(c, out, p + "conv2", 1, 1), p + "bn2"))
    out = _bn(c, _conv(c, out, p + "conv3"), p + "bn3")
    skip = _bn(c, _conv(c, x, p + "down_sample.0"), p + "down_sample.1") if has_ds else x
    return out + skip


def lr_decoder(c: Ctx, enc: Sequence[torch.Tensor], output_size, num_features: int,
               non_negative=False, p="decoder."):
    """Decoder.forward, ReSIDE/models/lasinger2019.py:279-295 + DecoderBlock :238-246."""
    sizes = [tuple(e.shape[-2:]) for e in reversed(enc[:-1])]
    sizes.append((sizes[-1][0] * 2, sizes[-1][1] * 2))
    out = None
    # Decoder(block_type=...) (lasinger2019.py:252-262): the checkpoint tells which block type was trained
    block = lr_bottleneck_block if f"{p}blocks.0.res_block1.conv3.weight" in c.sd else lr_residual_block
    for i, (e, size) in enumerate(zip(reversed(enc), sizes)):
        bp = f"{p}blocks.{i}."
        r = block(c, e, bp + "res_block1.", e.shape[1] != num_features)
        out = r if out is None else block(c, out + r, bp + "res_block2.", False)
        out = _interp(out, size)
    out = F.relu(_bn(c, _conv(c, out, p + "conv1", 1, 1), p + "bn1"))
    out = _interp(out, output_size)
    out = F.relu(_bn(c, _conv(c, out, p + "conv2", 1, 1), p + "bn2"))
    out = _conv(c, out, p + "conv3")
    return F.relu(out) if non_negative else out


def lr_encoder(c: Ctx, x, encoder: str):
    if encoder == "resnet50":
        # torchvision layout: blocks.0 = Sequential(conv1,bn1,relu,maxpool,layer1) -> index 4;
        # blocks.{1,2,3} = Sequential(layerk) (lasinger2019.py:107-111)
        return resnet50_stages_lr(c, x)
    return efficientnet_stages(c, x, encoder, "encoder.blocks.")


def resnet50_stages_lr(c: Ctx, x):
    x = F.relu(_bn(c, _conv(c, x, "encoder.blocks.0.0", 2, 3), "encoder.blocks.0.1"))
    x = F.max_pool2d(x, 3, 2, 1)
    outs = []
    for k, blocks in enumerate([3, 4, 6, 3], start=1):
        pre = "encoder.blocks.0.4." if k == 1 else f"encoder.blocks.{k - 1}.0."
        for i in range(blocks):
            x = bottleneck(c, x, f"{pre}{i}.", 2 if (i == 0 and k > 1) else 1, i == 0)
        outs.append(x)
        c.tap(f"enc{k}", x)
    return outs


def lr_forward(sd: SD, x, encoder: str, output_size_hw, calibrate=False, taps=None,
               non_negative=False):
    """MidasNet.forward, ReSIDE/models/lasinger2019.py:369-370 (`_output_size` is HW)."""
    c = Ctx(sd, calibrate, taps)
    enc = lr_encoder(c, x, encoder)
    nf = sd["decoder.conv1.weight"].shape[1]
    return lr_decoder(c, enc, output_size_hw, nf, non_negative)


# ----------------------------------------------------------------------------- Sobel / loss / metrics
SOBEL_X = ((1.0, 0.0, -1.0), (2.0, 0.0, -2.0), (1.0, 0.0, -1.0))     # ReSIDE/sobel.py:9
SOBEL_Y = ((1.0, 2.0, 1.0), (0.0, 0.0, 0.0), (-1.0, -2.0, -1.0))     # ReSIDE/sobel.py:10


def sobel(x):
    """Sobel.forward, ReSIDE/sobel.py:19-23: [N,1,H,W] -> [N,2,H,W] (dx, dy), zero pad."""
    w = torch.tensor([SOBEL_X, SOBEL_Y], dtype=x.dtype).view(2, 1, 3, 3)
    return F.conv2d(x, w, None, 1, 1)


def sobel_loss_terms(output, depth) -> Dict[str, torch.Tensor]:
    """Loss expression of ReSIDE/train.py:216-236 (cos = CosineSimilarity(dim=1, eps=0), :202)."""
    ones = torch.ones_like(depth)
    dg, og = sobel(depth), sobel(output)
    ddx, ddy, odx, ody = dg[:, 0:1], dg[:, 1:2], og[:, 0:1], og[:, 1:2]
    dn = torch.cat((-ddx, -ddy, ones), 1)
    on = torch.cat((-odx, -ody, ones), 1)
    cos = (on * dn).sum(1) / (on.norm(dim=1) * dn.norm(dim=1))
    t = {
        "depth": torch.log(torch.abs(output - depth) + 0.5).mean(),
        "dx": torch.log(torch.abs(odx - ddx) + 0.5).mean(),
        "dy": torch.log(torch.abs(ody - ddy) + 0.5).mean(),
        "normal": torch.abs(1 - cos).mean(),
    }
    t["loss"] = t["depth"] + t["normal"] + (t["dx"] + t["dy"])
    return t


def edge_magnitude(x):
    """edge_detection, ReSIDE/test.py:94-102."""
    g = sobel(x)
    return torch.sqrt(g[:, 0:1] ** 2 + g[:, 1:2] ** 2)


def metric_sums(outputs, labels) -> Dict[str, float]:
    """Raw sums behind MetricsTracker.update, ReSIDE/util.py:39-83 (fp64 accumulation)."""
    o, l = outputs.double(), labels.double()
    nan = torch.isnan(l)
    invalid = ~(l > 0)
    res = o - l
    rel = (res.abs() / l)
    rel[nan | invalid] = 0
    lg = (torch.log10(o) - torch.log10(l)).abs()
    lg[nan | invalid] = 0
    ratio = torch.max(o / l, l / o)
    return {
        "num_valid": float((~nan).sum()), "batch": float(l.shape[0]),
        "abs": float(res.abs().sum()), "sq": float((res ** 2).sum()), "rel": float(rel.sum()),
        "log10": float(lg.sum()),
        "d1": float((ratio <= 1.25).sum()), "d2": float((ratio <= 1.25 ** 2).sum()),
        "d3": float((ratio <= 1.25 ** 3).sum()),
    }


def metrics_from_sums(s: Dict[str, float]) -> Dict[str, float]:
    """One MetricsTracker.update + AverageMeter read-out (util.py:54-83,118-134) for a single
    batch, INCLUDING the reference's quirk that log10 lacks the batch_size factor (:68)."""
    nv, b = s["num_valid"], s["batch"]
    mse = b * s["sq"] / nv / b
    return {
        "mae": b * s["abs"] / nv / b, "mse": mse, "rmse": math.sqrt(mse),
        "abs_rel": b * s["rel"] / nv / b, "log10": s["log10"] / nv / b,
        "delta1": s["d1"] / nv * b / b, "delta2": s["d2"] / nv * b / b, "delta3": s["d3"] / nv * b / b,
    }


def relative_l1(y_new, y_ref) -> float:
    """Parity metric of SURVEY.md 8c: max over images of sum|new-ref| / sum|ref|."""
    d = (y_new.double() - y_ref.double()).abs().flatten(1).sum(1)
    return float((d / y_ref.double().abs().flatten(1).sum(1)).max())


# --------------------------------------------------------------------------- post-forward epilogue
def post_resize(output, size_hw):
    """ReSIDE/train.py:273-276, test.py:49-52, test_nyu.py:85: bilinear, align_corners=True."""
    return _interp(output, tuple(int(v) for v in size_hw))


def depth_to_mm_u16(resized):
    """ReSIDE/test_nyu.py:86-88: squeeze -> numpy; >10 m -> 0; x1000; astype(uint16).  [N,1,H,W] -> uint16 [N,H,W]."""
    import numpy as np
    a = resized[:, 0].detach().cpu().numpy().copy()
    a[a > 10.0] = 0.0
    with np.errstate(invalid="ignore"):
        # numpy's float32 -> uint16 cast is the C conversion through int32 on x86-64 for in-range values
        return (a * 1000.0).astype(np.int32).astype(np.uint16)


def edge_counts(output, depth, threshold=0.25):
    """ReSIDE/test.py:56-76: [#pred edges, #label edges, #both, #equal] summed over the batch."""
    e1, e2 = edge_magnitude(depth) > threshold, edge_magnitude(output) > threshold
    return [float(e2.sum()), float(e1.sum()), float((e1 & e2).sum()), float((e1 == e2).sum())]
