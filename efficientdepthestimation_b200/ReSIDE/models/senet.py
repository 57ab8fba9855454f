"""Stand-in for ``ReSIDE/models/senet.py`` (see densenet.py in this directory): SENet-154 is not part of the
B200 depth path; constructing it raises."""


def _unsupported(*_a, **_kw):
    raise RuntimeError("Unsupported encoder network 'senet': SENet encoders are not part of the B200 depth path "
                       "(efficientdepthestimation_b200 implements ResNet-50 and EfficientNet-B0..B7)")


senet154 = se_resnet50 = se_resnet101 = se_resnet152 = se_resnext50_32x4d = se_resnext101_32x4d = _unsupported
