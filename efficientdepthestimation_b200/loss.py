"""Fused Sobel depth/gradient/normal loss + depth metrics (host side of ede_sobel_loss_metrics).

Replaces the ~25 small kernels of the loss expression (ReSIDE/train.py:216-236) and the 9 host
syncs of MetricsTracker.update (ReSIDE/util.py:39-92) by ONE pass over (pred, label) producing a
16-double vector of partial sums on the device; every reported quantity is recovered from those
sums, so a multi-GPU run only has to all-reduce 128 bytes (parallel.py).
"""
import math
from typing import Dict

import torch

from . import _cabi as A

SUM_NAMES = ("log_depth", "normal", "log_dx", "log_dy", "abs", "sq", "rel", "log10", "d1", "d2", "d3",
             "num_valid", "num_pixels", "num_images", "_14", "_15")


def _check_pair(pred: torch.Tensor, label: torch.Tensor):
    if pred.device.type != "cuda" or label.device.type != "cuda":
        raise RuntimeError("efficientdepthestimation_b200.loss: CUDA tensors required (no CPU fallback)")
    if pred.dim() != 4 or pred.size(1) != 1 or pred.shape != label.shape:
        raise RuntimeError(f"expected pred/label of identical shape [N,1,H,W], got {tuple(pred.shape)} / {tuple(label.shape)}")
    return pred.detach().float().contiguous(), label.detach().float().contiguous()


def loss_metric_sums(pred: torch.Tensor, label: torch.Tensor, out: torch.Tensor = None) -> torch.Tensor:
    """16 fp64 partial sums (layout in include/ede.h) ACCUMULATED into `out` (device tensor); async.

    Accumulating several batches into one `out` means one NaN prediction poisons sums [0..7] of the whole run; the
    reference's AverageMeter drops only the offending batch (ReSIDE/util.py:131).  To keep that behaviour pass a fresh
    `out` per batch (the default) and feed it to ``MetricsTracker.update_from_sums`` -- the tracker checks per batch."""
    pred, label = _check_pair(pred, label)
    if out is None:
        out = torch.zeros(16, dtype=torch.float64, device=pred.device)
    n, _, h, w = pred.shape
    with torch.cuda.device(pred.device):
        A.check(A.lib().ede_sobel_loss_metrics(pred.data_ptr(), label.data_ptr(), n, h, w, out.data_ptr(),
                                               torch.cuda.current_stream(pred.device).cuda_stream),
                "ede_sobel_loss_metrics")
    return out


def edge_counts(pred: torch.Tensor, label: torch.Tensor, threshold: float = 0.25, out: torch.Tensor = None):
    """[#pred edges, #label edges, #both, #equal] of ReSIDE/test.py:56-76,94-102 as 4 fp64 on device."""
    pred, label = _check_pair(pred, label)
    if out is None:
        out = torch.zeros(4, dtype=torch.float64, device=pred.device)
    n, _, h, w = pred.shape
    with torch.cuda.device(pred.device):
        A.check(A.lib().ede_edge_counts(pred.data_ptr(), label.data_ptr(), n, h, w, float(threshold), out.data_ptr(),
                                        torch.cuda.current_stream(pred.device).cuda_stream), "ede_edge_counts")
    return out


def depth_epilogue(pred: torch.Tensor, size=None, label: torch.Tensor = None, *, want_resized: bool = False,
                   want_depth_mm: bool = False, want_sums: bool = None, want_edges: bool = False,
                   edge_threshold: float = 0.25, cap: float = 10.0, mm_scale: float = 1000.0,
                   sums: torch.Tensor = None, edges: torch.Tensor = None) -> Dict[str, torch.Tensor]:
    """ONE pass for everything the reference's eval callers do right after ``model(image)``:
    ``F.interpolate(output, size, 'bilinear', align_corners=True)`` (train.py:273-276, test.py:49-52,
    test_nyu.py:85) -> MetricsTracker sums + Sobel loss sums (util.py:39-83, train.py:216-236) and
    edge counts (test.py:56-76) against ``label`` -> cap / x1000 / uint16 (test_nyu.py:87-88).
    Returns the requested device tensors {"resized", "depth_mm", "sums", "edges"}; async, no host sync."""
    if pred.device.type != "cuda":
        raise RuntimeError("efficientdepthestimation_b200.loss: CUDA tensors required (no CPU fallback)")
    if pred.dim() != 4 or pred.size(1) != 1:
        raise RuntimeError(f"expected pred of shape [N,1,h,w], got {tuple(pred.shape)}")
    pred = pred.detach().float().contiguous()
    n, _, h, w = pred.shape
    if label is not None:
        if label.device != pred.device or label.dim() != 4 or label.size(0) != n or label.size(1) != 1:
            raise RuntimeError(f"expected label of shape [{n},1,H,W] on {pred.device}, got {tuple(label.shape)}")
        label = label.detach().float().contiguous()
        if size is None:
            size = tuple(label.shape[2:])
        elif tuple(size) != tuple(label.shape[2:]):
            raise RuntimeError(f"size {tuple(size)} does not match the label {tuple(label.shape[2:])}")
    if size is None:
        raise RuntimeError("depth_epilogue needs `size` or `label`")
    oh, ow = int(size[0]), int(size[1])
    if want_sums is None:
        want_sums = label is not None
    if (want_sums or want_edges) and label is None:
        raise RuntimeError("metric sums / edge counts need a label")
    out = {}
    dev = pred.device
    if want_resized:
        out["resized"] = torch.empty(n, 1, oh, ow, dtype=torch.float32, device=dev)
    if want_depth_mm:
        out["depth_mm"] = torch.empty(n, oh, ow, dtype=torch.uint16, device=dev)
    if want_sums:
        out["sums"] = sums if sums is not None else torch.zeros(16, dtype=torch.float64, device=dev)
    if want_edges:
        out["edges"] = edges if edges is not None else torch.zeros(4, dtype=torch.float64, device=dev)
    if not out:
        raise RuntimeError("depth_epilogue: no output requested")
    ptr = lambda k: out[k].data_ptr() if k in out else None  # noqa: E731
    a = A.EdeDepthEpilogue(pred=pred.data_ptr(), label=label.data_ptr() if label is not None else None,
                           resized=ptr("resized"), depth_mm=ptr("depth_mm"), sums16=ptr("sums"), edges4=ptr("edges"),
                           n=n, h=h, w=w, oh=oh, ow=ow, edge_thr=float(edge_threshold), cap=float(cap),
                           mm_scale=float(mm_scale))
    import ctypes
    with torch.cuda.device(dev):
        A.check(A.lib().ede_depth_epilogue(ctypes.byref(a), torch.cuda.current_stream(dev).cuda_stream),
                "ede_depth_epilogue")
    return out


def loss_from_sums(s) -> Dict[str, float]:
    """train.py:231-236 from global sums: every term is a mean over N*H*W pixels."""
    s = [float(v) for v in s]
    px = s[12]
    t = {"depth": s[0] / px, "normal": s[1] / px, "dx": s[2] / px, "dy": s[3] / px}
    t["loss"] = t["depth"] + t["normal"] + (t["dx"] + t["dy"])
    return t


def metrics_from_sums(s) -> Dict[str, float]:
    """One MetricsTracker.update + read-out (util.py:54-83) from sums, INCLUDING the reference's quirk
    that log10 lacks the batch_size factor (util.py:68)."""
    s = [float(v) for v in s]
    nv, b = s[11], s[13]
    mse = s[5] / nv
    return {"mae": s[4] / nv, "mse": mse, "rmse": math.sqrt(mse) if mse == mse and mse >= 0 else float("nan"),
            "abs_rel": s[6] / nv, "log10": s[7] / nv / b, "delta1": s[8] / nv, "delta2": s[9] / nv,
            "delta3": s[10] / nv}


def sobel_loss(output: torch.Tensor, depth: torch.Tensor) -> torch.Tensor:
    """The scalar loss of train.py:236 as a 0-dim device tensor (forward value only)."""
    s = loss_metric_sums(output, depth)
    return ((s[0] + s[1] + s[2] + s[3]) / s[12]).float()
