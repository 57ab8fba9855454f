"""Encoder wrappers and the Hu et al. decoder -- mirrors ``ReSIDE/models/modules.py`` of the reference
(_UpProjection :16-41, E_resnet :44-69, E_efficientnet :124-200, D :203-235, MFF :238-268, R :271-298).

Same constructors, attribute names and parameter shapes (state_dict contract); the
forward of every class is an operator emission into the B200 plan:
  * _UpProjection: bilinear(align_corners) resize -> ONE packed 5x5 conv with N = 2*Cout (conv1+bn1+relu
    and conv2+bn2 share their input, modules.py:35,37) -> 3x3 conv whose epilogue adds branch 2 and
    applies the final ReLU (modules.py:36,39).
  * MFF / model: `torch.cat` (modules.py:265, net.py:30) is producers writing channel slices.
"""
import contextlib
import os

from torch import nn

from ... import _cabi as A
from ...efficientnet_pytorch import EfficientNet
from ...engine import EdeModule, cat_views, emit


_LOWRES_UPCONV = os.environ.get("EDE_LOWRES_UPCONV", "1") != "0"
_LOWRES_MIN_CIN = int(os.environ.get("EDE_LOWRES_MIN_CIN", "192"))        # for moderate upsampling (4x in area)
_ADD_AS_TAP_MAX_LANES = 64     # [x1 | bran2] lanes up to which conv1_2 + add run as one GEMM (fits one 64-channel chunk)
# D.conv + that many up-projections at split-bf16 precision; -1 = by decoder width (measured, DESIGN.md section 5:
# ENB4-HU, 224 channels, needs 3 to hold 1e-2 on every image of a batch of 32; ENB0-HU, 160 channels, holds it with 0:
# max 4.6e-3 over a batch of 64, and each level costs it 0.03-0.4 ms per batch-64 step)
_ACCURATE_D = int(os.environ.get("EDE_ACCURATE_D", "-1"))
_ACCURATE_D_MAX_C = 256
_LOWRES_MIN_CIN_BIG = int(os.environ.get("EDE_LOWRES_MIN_CIN_BIG", "32"))  # for >= 8x in area: z is tiny


class _UpProjection(EdeModule, nn.Sequential):
    def __init__(self, num_input_features, num_output_features):
        super().__init__()
        cin, cout = num_input_features, num_output_features
        self.conv1 = nn.Conv2d(cin, cout, kernel_size=5, stride=1, padding=2, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.relu = nn.ReLU(inplace=True)
        self.conv1_2 = nn.Conv2d(cout, cout, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1_2 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cin, cout, kernel_size=5, stride=1, padding=2, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)

    def _emit(self, g, x, size, out=None, tag="up"):
        parts = [(self.conv1.weight, self.bn1, None, True), (self.conv2.weight, self.bn2, None, False)]
        grow = (int(size[0]) * int(size[1])) / float(x.h * x.w)
        if (_LOWRES_UPCONV and g.precision != "fp32" and grow >= 3.0 and 2 * self.conv1.out_channels <= 64
                and x.c >= (_LOWRES_MIN_CIN_BIG if grow >= 8.0 else _LOWRES_MIN_CIN)):
            # narrow projection of a strongly upsampled map (MFF): contract Cin on the low-resolution grid, then
            # gather the taps separably.  The intermediate z has 25*32 fp32 lanes per LOW-resolution pixel, so for
            # a mere 2x2 upsampling it only pays when Cin is large; from 8x in area on it is always smaller.
            pair = g.upconv(x, size, parts, (2, 2, 2, 2), A.ACT_RELU, tag=tag + ".conv1+conv2")
        else:
            x = g.resize(x, size, tag=tag + ".resize")
            pair = g.conv(x, parts, 1, (2, 2, 2, 2), A.ACT_RELU, tag=tag + ".conv1+conv2")
        half = pair.c_p // 2
        if g.precision != "fp32" and pair.c_p <= _ADD_AS_TAP_MAX_LANES:
            # narrow projection: the bran2 half rides along as extra input channels with an identity centre tap
            return g.conv_bn_act_add(pair, self.conv1_2, self.bn1_2, A.ACT_RELU, out=out, tag=tag + ".conv1_2")
        return g.conv_bn_act(pair.slice(0, half), self.conv1_2, self.bn1_2, A.ACT_RELU,
                             res=pair.slice(half, half), out=out, tag=tag + ".conv1_2")


class E_resnet(EdeModule, nn.Module):
    def __init__(self, original_model, num_features=2048):
        super().__init__()
        for name in ("conv1", "bn1", "relu", "maxpool", "layer1", "layer2", "layer3", "layer4"):
            setattr(self, name, getattr(original_model, name))

    def _emit(self, g, x):
        with g.encoder_scope():
            x = g.conv_bn_act(x, self.conv1, self.bn1, A.ACT_RELU, tag="E.conv1")
            x = g.maxpool3x3s2(x, tag="E.maxpool")
            feats = []
            for layer in (self.layer1, self.layer2, self.layer3, self.layer4):
                x = emit(layer, g, x)
                feats.append(x)
        return tuple(feats)


class _UnsupportedEncoder(nn.Module):
    """E_densenet / E_senet (modules.py:72-121 of the reference): importable -- Benchmark/reside_model.py:6 compares
    encoder TYPES against these names -- but not constructible: those encoders are outside the B200 depth path."""

    def __init__(self, original_model=None, num_features=None):
        super().__init__()
        raise RuntimeError(f"Unsupported encoder network '{type(self).__name__}': only E_resnet and E_efficientnet "
                           "are implemented on the B200 depth path")


class E_densenet(_UnsupportedEncoder):
    pass


class E_senet(_UnsupportedEncoder):
    pass


_EFFNET_SPLITS = {"efficientnet-b0": (3, 5, 8), "efficientnet-b1": (5, 8, 16), "efficient-b2": (5, 8, 16),
                  "efficientnet-b3": (5, 8, 18), "efficientnet-b4": (6, 10, 22), "efficientnet-b5": (8, 13, 27),
                  "efficientnet-b6": (9, 15, 31), "efficientnet-b7": (11, 18, 38)}


def split_efficientnet(encoder_name, pretrained):
    """The reference's 4-way split of an EfficientNet (modules.py:165-200 == lasinger2019.py:55-91):
    blocks[0] = Sequential(conv_stem, bn0, MBConv[0:i1]) -- NB no swish after the stem BN -- and
    blocks[k] = Sequential(MBConv[ik:ik+1]); out channels = `_bn2.num_features` of each last block."""
    if encoder_name not in _EFFNET_SPLITS:
        raise RuntimeError(f"Unsupported encoder network '{encoder_name}'.")
    net = EfficientNet.from_pretrained(encoder_name) if pretrained else EfficientNet.from_name(encoder_name)
    cuts = (0,) + _EFFNET_SPLITS[encoder_name] + (len(net._blocks),)
    stem = list(net.children())[:2]
    blocks = nn.ModuleList(
        nn.Sequential(*(stem if a == 0 else []), *net._blocks[a:b]) for a, b in zip(cuts[:-1], cuts[1:]))
    return blocks, [list(b.modules())[-2].num_features for b in blocks]


class E_efficientnet(EdeModule, nn.Module):
    """An encoder that uses a (pre-trained) EfficientNet."""

    def __init__(self, name="efficientnet-b0", pretrained=False, freeze_weights=False):
        super().__init__()
        self.name = name.lower()
        self.pretrained = pretrained
        self.freeze_weights = freeze_weights
        if not name.startswith("efficientnet"):
            raise RuntimeError(f"Unsupported encoder network '{self.name}'.")
        self.blocks, self._block_out_channels = self._make_efficientnet_encoder(self.name, pretrained)
        self.num_features = self.block_out_channels[-1]
        if freeze_weights:
            self.blocks.requires_grad_(False)

    @property
    def block_out_channels(self):
        return self._block_out_channels

    _make_efficientnet_encoder = staticmethod(split_efficientnet)

    def _emit(self, g, x):
        outs = []
        with g.encoder_scope():
            for block in self.blocks:
                x = emit(block, g, x)
                outs.append(x)
        return outs


class D(EdeModule, nn.Module):
    def __init__(self, num_features=2048):
        super().__init__()
        c = num_features // 2
        self.conv = nn.Conv2d(num_features, c, kernel_size=1, stride=1, bias=False)
        self.bn = nn.BatchNorm2d(c)
        for i in range(1, 5):
            setattr(self, f"up{i}", _UpProjection(num_input_features=c, num_output_features=c // 2))
            c //= 2
        self.out_channels = c

    def _emit(self, g, b1, b2, b3, b4, out=None):
        # Narrow decoders (EfficientNet: 160 / 224 channels after D.conv) run D.conv and the first `level`
        # up-projections at the encoder's split-bf16 precision: those low-resolution stages are a few per cent of the
        # FLOPs but the largest single source of bf16 error, which a random-init net amplifies.
        # ResNet-50's D (1024 channels, 30 % of the FLOPs) stays plain bf16.
        c = self.conv.out_channels
        level = 0 if c > _ACCURATE_D_MAX_C else (_ACCURATE_D if _ACCURATE_D >= 0 else (3 if c > 192 else 0))
        sizes = ([b3.h, b3.w], [b2.h, b2.w], [b1.h, b1.w], [b1.h * 2, b1.w * 2])
        ups = (self.up1, self.up2, self.up3, self.up4)
        with (g.encoder_scope() if level > 0 else contextlib.nullcontext()):
            x = g.conv_bn_act(b4, self.conv, self.bn, A.ACT_RELU, tag="D.conv")
        for i, (up, size) in enumerate(zip(ups, sizes)):
            with (g.encoder_scope() if i < min(level, 3) else contextlib.nullcontext()):
                x = up._emit(g, x, size, out=out if i == 3 else None, tag=f"D.up{i + 1}")
        return x


class MFF(EdeModule, nn.Module):
    def __init__(self, block_channel, num_features=64):
        super().__init__()
        for i in range(4):
            setattr(self, f"up{i + 1}", _UpProjection(num_input_features=block_channel[i], num_output_features=16))
        self.conv = nn.Conv2d(num_features, num_features, kernel_size=5, stride=1, padding=2, bias=False)
        self.bn = nn.BatchNorm2d(num_features)
        self.out_channels = num_features

    def _emit(self, g, b1, b2, b3, b4, size, out=None):
        slots = g.alloc_concat(b1.n, int(size[0]), int(size[1]), [16, 16, 16, 16])
        ms = [up._emit(g, b, size, out=slot, tag=f"MFF.up{i + 1}")
              for i, (up, b, slot) in enumerate(zip((self.up1, self.up2, self.up3, self.up4), (b1, b2, b3, b4), slots))]
        return g.conv_bn_act(cat_views(ms), self.conv, self.bn, A.ACT_RELU, out=out, tag="MFF.conv")


class R(EdeModule, nn.Module):
    def __init__(self, block_channel):
        super().__init__()
        c = 64 + block_channel[3] // 32
        self.conv0 = nn.Conv2d(c, c, kernel_size=5, stride=1, padding=2, bias=False)
        self.bn0 = nn.BatchNorm2d(c)
        self.conv1 = nn.Conv2d(c, c, kernel_size=5, stride=1, padding=2, bias=False)
        self.bn1 = nn.BatchNorm2d(c)
        self.conv2 = nn.Conv2d(c, 1, kernel_size=5, stride=1, padding=2, bias=True)

    def _emit(self, g, x):
        x = g.conv_bn_act(x, self.conv0, self.bn0, A.ACT_RELU, tag="R.conv0")
        x = g.conv_bn_act(x, self.conv1, self.bn1, A.ACT_RELU, tag="R.conv1")
        return g.conv_bn_act(x, self.conv2, None, A.ACT_NONE, out_dtype=A.F32, tag="R.conv2")
