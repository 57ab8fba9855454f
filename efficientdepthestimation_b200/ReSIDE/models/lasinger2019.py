""""LR" MiDaS-style model -- mirrors ``ReSIDE/models/lasinger2019.py`` of the reference
(Encoder :14-120, ResidualBlock :123-163, DecoderBlock :214-246, Decoder :249-295, MidasNet :298-422).

Same constructors / parameter names / checkpoint dict format; forwards are operator emissions:
  * ResidualBlock: pre-activation ReLU on the conv path only, the skip uses the RAW input (:153-161);
    the skip add is fused into conv2's epilogue.
  * DecoderBlock: `input_from_decoder + res_block1(enc)` (:241) is one vectorised NHWC add; the
    bilinear resize (:244) writes the next block's input.
"""
from typing import List, Optional, Union
from warnings import warn

import torch
from torch import nn

from ... import _cabi as A
from ...engine import EdeModule, emit
from .modules import split_efficientnet

__version__ = '0.2.0'


class Encoder(EdeModule, nn.Module):
    """An encoder that uses a (pre-trained) backbone network."""

    def __init__(self, name="efficientnet-b0", pretrained=False, freeze_weights=False):
        super().__init__()
        self.name = name.lower()
        self.pretrained = pretrained
        self.freeze_weights = freeze_weights
        if name.startswith("efficientnet"):
            self.blocks, self._block_out_channels = Encoder._make_efficientnet_encoder(self.name, pretrained)
        elif name.startswith("resnet"):
            self.blocks, self._block_out_channels = Encoder._make_resnet_encoder(self.name, pretrained)
        else:
            raise RuntimeError(f"Unsupported encoder network '{self.name}'.")
        if freeze_weights:
            self.blocks.requires_grad_(False)

    @property
    def block_out_channels(self) -> List[int]:
        return self._block_out_channels

    _make_efficientnet_encoder = staticmethod(split_efficientnet)

    @staticmethod
    def _make_resnet_encoder(encoder_name, pretrained):
        """torchvision layout (lasinger2019.py:93-120): blocks[0] = Sequential(conv1, bn1, relu, maxpool,
        layer1), blocks[k] = Sequential(layer{k+1}).  Offline, `pretrained` cannot download; callers
        overwrite the weights with load_state_dict anyway."""
        import torchvision
        ctor = {"resnet50": torchvision.models.resnet50, "resnet101": torchvision.models.resnet101,
                "resnet152": torchvision.models.resnet152}.get(encoder_name)
        if ctor is None:
            raise RuntimeError(f"Unsupported encoder network '{encoder_name}'.")
        layers = list(ctor(weights=None).children())[:-2]
        blocks = nn.ModuleList([nn.Sequential(*layers[:5])] + [nn.Sequential(layer) for layer in layers[5:]])
        return blocks, [list(block.modules())[-2].num_features for block in blocks]

    def _emit(self, g, x):
        outs = []
        with g.encoder_scope():
            for block in self.blocks:
                x = emit(block, g, x)
                outs.append(x)
        return outs


class ResidualBlock(EdeModule, nn.Module):
    """A simple convolutional block with a skip connection."""

    def __init__(self, in_channels, out_channels=None, stride=1):
        super().__init__()
        out_channels = in_channels if out_channels is None else out_channels
        self.conv1 = nn.Conv2d(in_channels, out_channels, kernel_size=3, padding=1, stride=stride, bias=False)
        self.bn1 = nn.BatchNorm2d(out_channels)
        self.conv2 = nn.Conv2d(out_channels, out_channels, kernel_size=3, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(out_channels)
        if in_channels != out_channels:
            self.down_sample = nn.Sequential(
                nn.Conv2d(in_channels, out_channels, kernel_size=1, stride=stride, bias=False),
                nn.BatchNorm2d(out_channels))
        else:
            self.down_sample = None

    def _emit(self, g, x, tag="res_block"):
        out = g.conv_bn_act(x, self.conv1, self.bn1, A.ACT_RELU, relu_in=True, tag=tag + ".conv1")
        skip = x
        if self.down_sample is not None:
            ds = list(self.down_sample.children())
            skip = g.conv_bn_act(x, ds[0], ds[1], A.ACT_NONE, tag=tag + ".down_sample")
        return g.conv_bn_act(out, self.conv2, self.bn2, A.ACT_NONE, res=skip, tag=tag + ".conv2")


class BottleneckBlock(EdeModule, nn.Module):
    """The bottleneck version of the residual block (lasinger2019.py:166-211 of the reference): pre-activation ReLU ->
    1x1 (to out/4) + BN + ReLU -> 3x3 + BN + ReLU -> 1x1 + BN, plus the RAW input (through a 1x1 + BN projection when
    the channel counts differ); no ReLU on the sum."""

    def __init__(self, in_channels, out_channels=None):
        super().__init__()
        out_channels = in_channels if out_channels is None else out_channels
        mid = max(1, out_channels // 4)           # the ratio of torchvision's Bottleneck (lasinger2019.py:180-182)
        self.conv1 = nn.Conv2d(in_channels, mid, kernel_size=1, bias=False)
        self.bn1 = nn.BatchNorm2d(mid)
        self.conv2 = nn.Conv2d(mid, mid, kernel_size=3, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(mid)
        self.conv3 = nn.Conv2d(mid, out_channels, kernel_size=1, bias=False)
        self.bn3 = nn.BatchNorm2d(out_channels)
        if in_channels != out_channels:
            self.down_sample = nn.Sequential(nn.Conv2d(in_channels, out_channels, kernel_size=1, bias=False),
                                             nn.BatchNorm2d(out_channels))
        else:
            self.down_sample = None

    def _emit(self, g, x, tag="bottleneck_block"):
        out = g.conv_bn_act(x, self.conv1, self.bn1, A.ACT_RELU, relu_in=True, tag=tag + ".conv1")
        out = g.conv_bn_act(out, self.conv2, self.bn2, A.ACT_RELU, tag=tag + ".conv2")
        skip = x
        if self.down_sample is not None:
            ds = list(self.down_sample.children())
            skip = g.conv_bn_act(x, ds[0], ds[1], A.ACT_NONE, tag=tag + ".down_sample")
        return g.conv_bn_act(out, self.conv3, self.bn3, A.ACT_NONE, res=skip, tag=tag + ".conv3")


class DecoderBlock(EdeModule, nn.Module):
    """Transforms the previous decoder block's output and the corresponding encoder block's output."""

    def __init__(self, in_channels, out_channels, block_type=ResidualBlock):
        super().__init__()
        self.res_block1 = block_type(in_channels, out_channels)
        self.res_block2 = block_type(out_channels)

    def _emit(self, g, input_from_encoder, input_from_decoder, output_size, tag="decoder_block"):
        out = self.res_block1._emit(g, input_from_encoder, tag=tag + ".res_block1")
        if input_from_decoder is not None:
            out = g.eltwise(input_from_decoder, out, tag=tag + ".add")
            out = self.res_block2._emit(g, out, tag=tag + ".res_block2")
        return g.resize(out, output_size, tag=tag + ".resize")


class Decoder(EdeModule, nn.Module):
    """The decoder sub-network of the encoder-decoder architecture."""

    def __init__(self, encoder_block_channels, num_features: Union[int, str] = 'auto', non_negative=False,
                 block_type=ResidualBlock):
        super().__init__()
        self.num_features = encoder_block_channels[0] if num_features == 'auto' else int(num_features)
        self.non_negative = non_negative
        self.blocks = nn.ModuleList(
            [DecoderBlock(c, self.num_features, block_type) for c in reversed(encoder_block_channels)])
        self.conv1 = nn.Conv2d(self.num_features, 128, kernel_size=3, padding=1, bias=False)
        self.bn1 = nn.BatchNorm2d(128)
        self.conv2 = nn.Conv2d(128, 32, kernel_size=3, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(32)
        self.conv3 = nn.Conv2d(32, 1, kernel_size=1, padding=0, bias=True)

    def _emit(self, g, encoder_outputs, output_size):
        sizes = [(e.h, e.w) for e in reversed(encoder_outputs[:-1])]
        sizes.append((sizes[-1][0] * 2, sizes[-1][1] * 2))
        out = None
        for i, (block, enc, size) in enumerate(zip(self.blocks, reversed(encoder_outputs), sizes)):
            out = block._emit(g, enc, out, size, tag=f"decoder.blocks.{i}")
        out = g.conv_bn_act(out, self.conv1, self.bn1, A.ACT_RELU, tag="decoder.conv1")
        out = g.resize(out, output_size, tag="decoder.resize")
        out = g.conv_bn_act(out, self.conv2, self.bn2, A.ACT_RELU, tag="decoder.conv2")
        return g.conv_bn_act(out, self.conv3, None, A.ACT_RELU if self.non_negative else A.ACT_NONE,
                             out_dtype=A.F32, tag="decoder.conv3")


class MidasNet(EdeModule, nn.Module):
    """Re-implementation of MiDaS (Lasinger/Ranftl et al. 2019) as the reference defines it."""

    def __init__(self, encoder: Encoder, decoder: Decoder, output_size: tuple, input_size: Optional[tuple] = None,
                 adversarial_training=False):
        super().__init__()
        self.encoder = encoder
        self.decoder = decoder
        self.input_size = input_size if input_size else output_size
        self.output_size = output_size
        self.adversarial_training = adversarial_training

    # sizes are exposed in WH order and stored in HW order, as the reference does (:343-367)
    @property
    def input_size(self):
        return self._input_size[1], self._input_size[0]

    @input_size.setter
    def input_size(self, size):
        self._input_size = (size[1], size[0])

    @property
    def output_size(self):
        return self._output_size[1], self._output_size[0]

    @output_size.setter
    def output_size(self, size):
        self._output_size = (size[1], size[0])
        if isinstance(self, EdeModule):
            self.invalidate_plans()

    def _emit(self, g, x):
        return self.decoder._emit(g, self.encoder._emit(g, x), self._output_size)

    def save(self, f):
        """Checkpoint dict format of the reference (lasinger2019.py:372-394)."""
        torch.save({
            'encoder': {'name': self.encoder.name, 'freeze_weights': self.encoder.freeze_weights},
            'decoder': {'num_features': self.decoder.num_features, 'non_negative': self.decoder.non_negative},
            'input_size': self.input_size, 'output_size': self.output_size,
            'adversarial_training': self.adversarial_training,
            'weights': self.state_dict(), 'version': __version__}, f)

    @staticmethod
    def load(f):
        state = torch.load(f)
        if state['version'] != __version__:
            warn(f"Version mismatch: the loaded weights are for version {state['version']}, however the "
                 f"current version is {__version__}. This may cause issues loading the model.")
        encoder = Encoder(**state['encoder'])
        decoder = Decoder(encoder.block_out_channels, **state['decoder'])
        net = MidasNet(encoder, decoder, state['output_size'], state['input_size'], state['adversarial_training'])
        net.load_state_dict(state['weights'])
        return net

    @property
    def num_parameters(self):
        return sum(p.nelement() for p in self.parameters())
