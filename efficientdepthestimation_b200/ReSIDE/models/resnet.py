"""ResNet-50 definition -- mirrors ``ReSIDE/models/resnet.py`` of the reference (Bottleneck :59-95,
ResNet :98-155, resnet50 :179-187): identical attribute names / shapes, stride on the 3x3.
Only what the depth path uses is provided (Bottleneck nets); the classifier head (avgpool, fc)
is kept as parameters for state_dict compatibility but never executed.
"""
import math

from torch import nn

from ...engine import EdeModule, emit, emit_bottleneck

__all__ = ["ResNet", "Bottleneck", "resnet50", "resnet101", "resnet152"]


class Bottleneck(EdeModule, nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, kernel_size=1, bias=False)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = nn.Conv2d(planes, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = nn.Conv2d(planes, planes * 4, kernel_size=1, bias=False)
        self.bn3 = nn.BatchNorm2d(planes * 4)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def _emit(self, g, x):
        return emit_bottleneck(self, g, x)


class ResNet(EdeModule, nn.Module):
    """Stem (7x7 s2 conv, BN, ReLU, 3x3 s2 max-pool) + four stages of `block`; registration order and
    names (conv1, bn1, relu, maxpool, layer1..4, avgpool, fc) are the state_dict contract."""
    _STAGES = ((64, 1), (128, 2), (256, 2), (512, 2))          # (planes, stride of the first block)

    def __init__(self, block, layers, num_classes=1000):
        super().__init__()
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        for k, ((planes, stride), count) in enumerate(zip(self._STAGES, layers), start=1):
            setattr(self, f"layer{k}", self._make_layer(block, planes, count, stride))
        self.avgpool = nn.AvgPool2d(7, stride=1)
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        self._he_init()

    def _he_init(self):
        """N(0, sqrt(2 / (k*k*cout))) conv weights, unit BN -- the reference's init (resnet.py:115-121)."""
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                fan_out = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
                nn.init.normal_(m.weight, 0.0, math.sqrt(2.0 / fan_out))
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)

    def _make_layer(self, block, planes, blocks, stride=1):
        width = planes * block.expansion
        shortcut = None
        if stride != 1 or self.inplanes != width:
            shortcut = nn.Sequential(nn.Conv2d(self.inplanes, width, kernel_size=1, stride=stride, bias=False),
                                     nn.BatchNorm2d(width))
        stage = [block(self.inplanes, planes, stride, shortcut)]
        self.inplanes = width
        stage.extend(block(width, planes) for _ in range(blocks - 1))
        return nn.Sequential(*stage)

    def _emit(self, g, x):
        raise RuntimeError("the ImageNet classifier forward of ResNet is outside the depth-estimation path; "
                           "wrap it in modules.E_resnet")


# `pretrained` is accepted for signature parity only: the reference downloads ImageNet weights there
# (resnet.py:185-186) and every caller then overwrites them with load_state_dict; offline the random-init
# network is returned.
def resnet50(pretrained=False, **kwargs):
    return ResNet(Bottleneck, [3, 4, 6, 3], **kwargs)


def resnet101(pretrained=False, **kwargs):
    return ResNet(Bottleneck, [3, 4, 23, 3], **kwargs)


def resnet152(pretrained=False, **kwargs):
    return ResNet(Bottleneck, [3, 8, 36, 3], **kwargs)
