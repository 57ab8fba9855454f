"""Sobel gradients -- mirrors ``ReSIDE/sobel.py`` (Sobel :5-23): frozen 3x3 x/y taps stored as
``edge_conv.weight`` (state_dict key kept), forward = ede_sobel (fp32 NCHW [N,1,H,W] -> [N,2,H,W])."""
import torch
from torch import nn

from .. import _cabi as A


class Sobel(nn.Module):
    def __init__(self):
        super().__init__()
        self.edge_conv = nn.Conv2d(1, 2, kernel_size=3, stride=1, padding=1, bias=False)
        kx = torch.tensor([[1., 0., -1.], [2., 0., -2.], [1., 0., -1.]])
        ky = torch.tensor([[1., 2., 1.], [0., 0., 0.], [-1., -2., -1.]])
        self.edge_conv.weight = nn.Parameter(torch.stack((kx, ky)).view(2, 1, 3, 3))
        for p in self.parameters():
            p.requires_grad = False

    def forward(self, x):
        if x.device.type != "cuda":
            raise RuntimeError("Sobel.forward: efficientdepthestimation_b200 runs on CUDA (sm_100a) only; no CPU fallback")
        if x.dim() != 4 or x.size(1) != 1:
            raise RuntimeError(f"Sobel.forward expects [N,1,H,W], got {tuple(x.shape)}")
        x = x.detach().float().contiguous()
        n, _, h, w = x.shape
        out = torch.empty(n, 2, h, w, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            A.check(A.lib().ede_sobel(x.data_ptr(), out.data_ptr(), n, h, w,
                                      torch.cuda.current_stream(x.device).cuda_stream), "ede_sobel")
        return out
