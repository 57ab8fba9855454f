"""Top-level Hu et al. model -- mirrors ``ReSIDE/models/net.py`` of the reference.

``model(Encoder, num_features, block_channel).forward(x) -> [N,1,2*h1,2*w1]`` (net.py:18-32):
E -> (D, MFF) -> cat -> R.  Here the ``torch.cat`` of net.py:30 is free: D.up4 and MFF.conv write
adjacent channel slices of one NHWC buffer that R.conv0 reads.
"""
from torch import nn

from ...engine import EdeModule, cat_views, emit
from . import modules


class model(EdeModule, nn.Module):
    def __init__(self, Encoder, num_features, block_channel):
        super().__init__()
        self.E = Encoder
        self.D = modules.D(num_features)
        self.MFF = modules.MFF(block_channel)
        self.R = modules.R(block_channel)

    def _emit(self, g, x):
        b1, b2, b3, b4 = emit(self.E, g, x)
        oh, ow = b1.h * 2, b1.w * 2
        d_slot, m_slot = g.alloc_concat(x.n, oh, ow, [self.D.out_channels, self.MFF.out_channels])
        d = self.D._emit(g, b1, b2, b3, b4, out=d_slot)
        m = self.MFF._emit(g, b1, b2, b3, b4, [d.h, d.w], out=m_slot)
        return self.R._emit(g, cat_views([d, m]))

    @property
    def num_parameters(self):
        """Total number of parameters of the model and its sub-modules (net.py:34-39)."""
        return sum(p.nelement() for p in self.parameters())
