"""B200-native (sm_100a) implementation of the ReSIDE depth-estimation forward path.

Public surface: the reference's own module API re-provided under ``efficientdepthestimation_b200.ReSIDE``
(``models.net.model``, ``models.modules.*``, ``models.resnet.resnet50``, ``models.lasinger2019.*``,
``sobel.Sobel``, ``util.MetricsTracker``, ``train.define_model``), ``loss`` (fused Sobel loss + metrics),
``parallel`` (batch sharding + scalar all-reduce) and ``dropin.install()`` which aliases those modules as
top-level ``ReSIDE`` / ``efficientnet_pytorch`` so the reference scripts run unchanged.
Arithmetic lives in ``lib/libede_sm100.so`` (C ABI: include/ede.h); there is no CPU fallback.
"""
__version__ = "0.1.0"
