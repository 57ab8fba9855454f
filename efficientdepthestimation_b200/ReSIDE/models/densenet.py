"""Stand-in for ``ReSIDE/models/densenet.py`` so that the reference's import lines (`from ReSIDE.models import
modules, resnet, densenet, net, senet`, ReSIDE/test.py:7, train.py:15) work against the mirror package.
DenseNet-161 is not part of the B200 depth path (BASELINE.json north_star): constructing it
raises the reference's own error type."""


def _unsupported(*_a, **_kw):
    raise RuntimeError("Unsupported encoder network 'densenet': DenseNet encoders are not part of the B200 depth path "
                       "(efficientdepthestimation_b200 implements ResNet-50 and EfficientNet-B0..B7)")


densenet121 = densenet169 = densenet201 = densenet161 = _unsupported
