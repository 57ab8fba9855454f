"""Golden fixtures for the rows either side of the forward path, made by the REAL reference code.

Run in the build container only (needs /root/reference and Pillow):
    python tests/golden/make_golden_pipeline.py

* preprocess_frames.npz -- uint8 frames pushed through the reference's own transform classes
  (ReSIDE/demo_transform.py: Scale, CenterCrop, ToTensor, Normalize, composed as in
  ReSIDE/inference_benchmark.py:46-52) on top of the real Pillow; same aspect ratios as the 640x480 -> 320x240 ->
  304x228 production case, plus an up-scaling and a 4.5x down-scaling case.
* depth_epilogue.npz -- the verbatim post-forward lines of ReSIDE/test_nyu.py:85-88 (interpolate, cap, x1000,
  uint16), ReSIDE/test.py:49-76,94-102 (metrics + edge counts) on seeded maps.
It asserts that the oracle restatements (oracle/preprocess_cpu.py, oracle/reside_cpu.py) reproduce them exactly.
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle", "_shim"), "/root/reference", "/root/reference/ReSIDE", ROOT]

from PIL import Image                                   # noqa: E402
from ReSIDE import demo_transform as T                  # noqa: E402
from ReSIDE import sobel as ref_sobel                   # noqa: E402
from ReSIDE.util import MetricsTracker                  # noqa: E402
from oracle import preprocess_cpu as P                  # noqa: E402
from oracle import reside_cpu as rc                     # noqa: E402
from oracle import weights as W                         # noqa: E402

MEAN, STD = [0.485, 0.456, 0.406], [0.229, 0.224, 0.225]


def reference_transform(frame, scale_wh, crop_wh):
    img = Image.fromarray(frame)
    for t in (T.Scale(list(scale_wh)), T.CenterCrop(list(crop_wh)), T.ToTensor(), T.Normalize(MEAN, STD)):
        img = t(img)
    return img.numpy()


def make_preprocess():
    rng = np.random.default_rng(7)
    cases = {}
    for name, (h, w), n, scale_wh, crop_wh in [("down2", (120, 160), 2, (80, 60), (76, 57)),
                                               ("up", (45, 70), 1, (80, 60), (76, 57)),
                                               ("down4p5", (135, 240), 1, (53, 30), (50, 28)),
                                               ("same_h", (60, 97), 1, (80, 60), (80, 60))]:
        frames = rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8)
        frames[0, : h // 3] = (np.linspace(0, 255, w)[None, :, None] * np.ones((h // 3, 1, 3))).astype(np.uint8)
        ref = np.stack([reference_transform(f, scale_wh, crop_wh) for f in frames])
        mine = P.preprocess_frames(frames, scale_wh, crop_wh, MEAN, STD)
        assert ref.dtype == np.float32 and np.array_equal(ref, mine), name
        cases[name + "_frames"] = frames
        cases[name + "_out"] = ref
        cases[name + "_cfg"] = np.array(list(scale_wh) + list(crop_wh), np.int32)
    # production geometry, checked here (too large to commit): 640x480 -> 320x240 -> 304x228
    f = rng.integers(0, 256, (1, 480, 640, 3), dtype=np.uint8)
    assert np.array_equal(reference_transform(f[0], (320, 240), (304, 228)), P.preprocess_frames(f)[0])
    np.savez_compressed(os.path.join(HERE, "preprocess_frames.npz"), **cases)
    print("preprocess_frames.npz:", {k: v.shape for k, v in cases.items()})


def edge_detection(depth):
    """ReSIDE/test.py:94-102 with `.cuda()` dropped."""
    get_edge = ref_sobel.Sobel()
    edge_xy = get_edge(depth)
    edge_sobel = torch.pow(edge_xy[:, 0, :, :], 2) + torch.pow(edge_xy[:, 1, :, :], 2)
    return torch.sqrt(edge_sobel)


def make_epilogue():
    g = torch.Generator().manual_seed(11)
    out = {}
    for name, n, (h, w), (oh, ow), holes in [("nyu", 2, (57, 76), (120, 160), False),
                                             ("holes", 3, (29, 38), (57, 77), True),
                                             ("same", 1, (16, 20), (16, 20), False)]:
        pred = 0.2 + 11.5 * torch.rand(n, 1, h, w, generator=g)          # some pixels above the 10 m cap
        pred[0, 0, 0, :4] = torch.tensor([10.0, 10.0005, 9.9995, 0.0004])
        depth = W.make_labels(n, oh, ow, seed=5, holes=holes)
        with torch.no_grad():
            output = F.interpolate(pred, size=[depth.size(2), depth.size(3)], mode="bilinear", align_corners=True)
            outputs = output.squeeze(1).cpu().numpy().copy()               # test_nyu.py:86 (squeeze of [N,1,H,W])
            outputs[outputs > 10.0] = 0.0
            outputs = (outputs * 1000.0).astype(np.uint16)
            m = MetricsTracker()
            m.update(outputs=output, labels=depth)
            thre = 0.25
            e1, e2 = edge_detection(depth) > thre, edge_detection(output) > thre
            edges = np.array([float(e2.sum()), float(e1.sum()), float(((e1.int() + e2.int()) == 2).sum()),
                              float(torch.eq(e1, e2).float().sum())])
        assert torch.equal(rc.post_resize(pred, (oh, ow)), output)
        assert np.array_equal(rc.depth_to_mm_u16(output), outputs), name
        assert rc.edge_counts(output, depth, thre) == list(edges), name
        out[name + "_pred"] = pred.numpy()
        out[name + "_label"] = depth.numpy()
        out[name + "_resized"] = output.numpy()
        out[name + "_mm"] = outputs
        out[name + "_edges"] = edges
        md = m.to_dict()
        out[name + "_metrics"] = np.array([md[k] for k in ("mae", "mse", "rmse", "abs_rel", "log10", "delta1", "delta2", "delta3")], np.float64)
    np.savez_compressed(os.path.join(HERE, "depth_epilogue.npz"), **out)
    print("depth_epilogue.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    make_preprocess()
    make_epilogue()
