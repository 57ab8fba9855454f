"""Named constructors for the six model variants of the reference (README.md:49-56), built with the
reference's own constructor calls (ReSIDE/train.py:20-38,86-91) minus the downloads."""
from typing import Tuple

from .ReSIDE.models import modules, net, resnet
from .ReSIDE.models.lasinger2019 import BottleneckBlock, Decoder, Encoder, MidasNet, ResidualBlock

ARCHS = ("rn50-hu", "enb0-hu", "enb4-hu", "rn50-lr", "enb0-lr", "enb4-lr")
_ENC = {"rn50": "resnet50", "enb0": "efficientnet-b0", "enb4": "efficientnet-b4"}


def input_hw(arch: str) -> Tuple[int, int]:
    return (228, 304) if arch.endswith("-hu") else (384, 384)


def build_model(arch: str, lr_output_wh=None, lr_block: str = "residual"):
    """arch = '<rn50|enb0|enb4>-<hu|lr>'; random-init weights (load a state_dict afterwards)."""
    if arch not in ARCHS:
        raise RuntimeError(f"Unsupported model '{arch}' (expected one of {', '.join(ARCHS)})")
    enc, dec = arch.split("-")
    if dec == "hu":
        if enc == "rn50":
            return net.model(modules.E_resnet(resnet.resnet50(pretrained=False)), 2048, [256, 512, 1024, 2048])
        e = modules.E_efficientnet(_ENC[enc], pretrained=False)
        return net.model(e, e.num_features, e.block_out_channels)
    e = Encoder(_ENC[enc], pretrained=False)
    h, w = input_hw(arch)
    block = {"residual": ResidualBlock, "bottleneck": BottleneckBlock}[lr_block]      # Decoder(block_type=...), :252-262
    return MidasNet(e, Decoder(e.block_out_channels, num_features="auto", block_type=block),
                    output_size=lr_output_wh or (w // 2, h // 2), input_size=(w, h))
