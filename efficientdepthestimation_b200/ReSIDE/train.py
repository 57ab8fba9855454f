"""Model factory -- mirrors ``define_model`` of ``ReSIDE/train.py:20-38`` (the single constructor every
reference script calls).  The training loop itself is outside the forward path."""
from .models import modules, net, resnet
from .models.lasinger2019 import Decoder, Encoder, MidasNet  # noqa: F401  (re-exported like the reference)


def define_model(is_resnet=False, is_densenet=False, is_senet=False, is_efficientnet=False,
                 efficientnet_variant="efficientnet-b0"):
    if is_resnet:
        encoder = modules.E_resnet(resnet.resnet50(pretrained=True))
        return net.model(encoder, num_features=2048, block_channel=[256, 512, 1024, 2048])
    if is_efficientnet:
        encoder = modules.E_efficientnet(name=efficientnet_variant, pretrained=True)
        return net.model(encoder, num_features=encoder.num_features, block_channel=encoder.block_out_channels)
    if is_densenet or is_senet:
        raise RuntimeError("Unsupported encoder network: DenseNet-161 / SENet-154 are not part of the B200 depth path")
    raise RuntimeError("define_model: no encoder selected")
