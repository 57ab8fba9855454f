"""TEST INFRASTRUCTURE ONLY (imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs).

CPU restatement (numpy, integer arithmetic) of the reference's video-frame preprocessing:

    transforms.Compose([Scale([320, 240]), CenterCrop([304, 228]), ToTensor(), Normalize(mean, std)])
        ReSIDE/inference_benchmark.py:46-52, ReSIDE/demo_transform.py:23-127

`Scale` is `PIL.Image.resize((ow, oh), Image.BILINEAR)` (demo_transform.py:31-34).  Pillow is a third-party
dependency of the reference (requirements.txt: Pillow; this container has Pillow 12.2.0); its published algorithm
(src/libImaging/Resample.c: precompute_coeffs, normalize_coeffs_8bpc, ImagingResampleHorizontal_8bpc,
ImagingResampleVertical_8bpc) is restated here and PINNED against the real library in
tests/test_oracle_golden.py (live when PIL imports, and through tests/golden/preprocess_frames.npz).
"""
import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2
IMAGENET_MEAN = (0.485, 0.456, 0.406)      # inference_benchmark.py:43-44
IMAGENET_STD = (0.229, 0.224, 0.225)


def bilinear_coeffs(in_size: int, out_size: int):
    """precompute_coeffs + normalize_coeffs_8bpc for the triangle filter (support 1.0)."""
    scale = in_size / out_size
    filterscale = max(scale, 1.0)
    support = 1.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    ss = 1.0 / filterscale
    bounds = np.zeros((out_size, 2), np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)
        xmax = min(int(center + support + 0.5), in_size) - xmin
        w = []
        for x in range(xmax):
            t = abs((x + xmin - center + 0.5) * ss)
            w.append(1.0 - t if t < 1.0 else 0.0)
        ww = sum(w)          # left-to-right double additions, as in the C loop
        for x in range(xmax):
            v = w[x] / ww if ww != 0.0 else w[x]
            kk[xx, x] = int(-0.5 + v * (1 << PRECISION_BITS)) if v < 0 else int(0.5 + v * (1 << PRECISION_BITS))
        bounds[xx] = (xmin, xmax)
    return bounds, kk


def _resample_axis0(img: np.ndarray, out_size: int) -> np.ndarray:
    """One 8bpc pass along axis 0 of a uint8 array [in, ...]."""
    bounds, kk = bilinear_coeffs(img.shape[0], out_size)
    out = np.empty((out_size,) + img.shape[1:], np.uint8)
    src = img.astype(np.int64)
    for i in range(out_size):
        x0, n = bounds[i]
        k = kk[i, :n].astype(np.int64).reshape((n,) + (1,) * (img.ndim - 1))
        s = (1 << (PRECISION_BITS - 1)) + (src[x0:x0 + n] * k).sum(0)
        out[i] = np.clip(s >> PRECISION_BITS, 0, 255).astype(np.uint8)
    return out


def pil_resize_bilinear(frame: np.ndarray, size_wh) -> np.ndarray:
    """Image.resize(size_wh, Image.BILINEAR) of a uint8 HWC frame: horizontal pass, then vertical (Resample.c)."""
    ow, oh = size_wh
    h, w = frame.shape[:2]
    x = frame
    if ow != w:
        x = _resample_axis0(np.ascontiguousarray(x.transpose(1, 0, 2)), ow).transpose(1, 0, 2)
    if oh != h:
        x = _resample_axis0(np.ascontiguousarray(x), oh)
    return np.ascontiguousarray(x)


def center_crop_box(size_wh, crop_wh):
    """CenterCrop.centerCrop, demo_transform.py:45-58: (x1, y1)."""
    (w1, h1), (tw, th) = size_wh, crop_wh
    if w1 == tw and h1 == th:
        return 0, 0
    return int(round((w1 - tw) / 2.0)), int(round((h1 - th) / 2.0))


def preprocess_frames(frames: np.ndarray, scale_wh=(320, 240), crop_wh=(304, 228), mean=IMAGENET_MEAN,
                      std=IMAGENET_STD) -> np.ndarray:
    """uint8 [N,H,W,3] -> float32 [N,3,crop_h,crop_w] through Scale, CenterCrop, ToTensor, Normalize."""
    out = []
    x1, y1 = center_crop_box(scale_wh, crop_wh)
    for f in frames:
        r = pil_resize_bilinear(f, scale_wh)[y1:y1 + crop_wh[1], x1:x1 + crop_wh[0]]
        t = r.transpose(2, 0, 1).astype(np.float32) / np.float32(255)            # ToTensor: float().div(255)
        for c in range(3):                                                          # Normalize: sub_(m).div_(s) in fp32
            t[c] = (t[c] - np.float32(mean[c])) / np.float32(std[c])
        out.append(t)
    return np.stack(out)
