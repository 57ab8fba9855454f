"""not gpu: the oracle restatements of the rows either side of the forward path against the
golden vectors made by the REAL reference code (tests/golden/make_golden_pipeline.py), Pillow itself when it
imports, and the library's HOST coefficient routine (no GPU needed) against the oracle's tables."""
import os

import numpy as np
import pytest
import torch

from oracle import preprocess_cpu as P
from oracle import reside_cpu as rc

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
PRE_CASES = ["down2", "up", "down4p5", "same_h"]
EPI_CASES = ["nyu", "holes", "same"]


@pytest.mark.parametrize("case", PRE_CASES)
def test_preprocess_oracle_matches_reference_transform(case):
    g = np.load(os.path.join(GOLDEN, "preprocess_frames.npz"))
    sw, sh, cw, ch = (int(v) for v in g[case + "_cfg"])
    out = P.preprocess_frames(g[case + "_frames"], (sw, sh), (cw, ch))
    assert out.dtype == np.float32 and np.array_equal(out, g[case + "_out"])       # bit-exact


def test_resample_oracle_matches_pillow_live():
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(3)
    for (h, w), (ow, oh) in [((96, 128), (64, 48)), ((50, 31), (64, 48)), ((200, 333), (37, 21)), ((48, 100), (64, 48))]:
        f = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        ref = np.asarray(Image.fromarray(f).resize((ow, oh), Image.BILINEAR))
        assert np.array_equal(P.pil_resize_bilinear(f, (ow, oh)), ref), ((h, w), (ow, oh))


def test_center_crop_box():
    assert P.center_crop_box((320, 240), (304, 228)) == (8, 6)          # inference_benchmark.py:47-48
    assert P.center_crop_box((304, 228), (304, 228)) == (0, 0)
    assert P.center_crop_box((53, 30), (50, 28)) == (2, 1)               # round-half-even of 1.5 -> 2


@pytest.mark.parametrize("in_size,out_size", [(640, 320), (480, 240), (70, 80), (240, 53), (1920, 320), (97, 80),
                                              (5, 5), (1, 7), (7, 1)])
def test_library_host_coefficients_equal_oracle(lib, in_size, out_size):
    from efficientdepthestimation_b200 import preprocess as PP
    b, k = PP.resample_tables(in_size, out_size, 0, out_size)
    b2, k2 = P.bilinear_coeffs(in_size, out_size)
    assert np.array_equal(b, b2) and np.array_equal(k, k2)
    if out_size > 4:                                                     # a cropped window is a slice of the table
        b3, k3 = PP.resample_tables(in_size, out_size, 2, out_size - 3)
        assert np.array_equal(b3, b2[2:out_size - 1]) and np.array_equal(k3, k2[2:out_size - 1])
    assert lib.ede_resample_coeffs(0, 4, 0, 4, None, None) < 0 and b"ede_resample_coeffs" in lib.ede_last_error()
    assert lib.ede_resample_coeffs(8, 4, 2, 3, None, None) < 0          # window past the end


@pytest.mark.parametrize("case", EPI_CASES)
def test_epilogue_oracle_matches_reference_lines(case):
    g = np.load(os.path.join(GOLDEN, "depth_epilogue.npz"))
    pred, label = torch.from_numpy(g[case + "_pred"]), torch.from_numpy(g[case + "_label"])
    r = rc.post_resize(pred, label.shape[2:])
    assert np.array_equal(r.numpy(), g[case + "_resized"])
    assert np.array_equal(rc.depth_to_mm_u16(r), g[case + "_mm"])
    assert rc.edge_counts(r, label) == list(g[case + "_edges"])
    m = rc.metrics_from_sums(rc.metric_sums(r, label))
    for k, v in zip(("mae", "mse", "rmse", "abs_rel", "log10", "delta1", "delta2", "delta3"), g[case + "_metrics"]):
        assert (m[k] != m[k] and v != v) or abs(m[k] - v) <= 1e-5 * max(1.0, abs(v)), (k, m[k], v)


def test_product_hosts_refuse_cpu_tensors(lib):
    from efficientdepthestimation_b200 import loss as L
    from efficientdepthestimation_b200 import preprocess as PP
    with pytest.raises(RuntimeError):
        PP.FramePreprocessor()(torch.zeros(1, 8, 8, 3, dtype=torch.uint8))
    with pytest.raises(RuntimeError):
        L.depth_epilogue(torch.zeros(1, 1, 4, 4), size=(8, 8), want_resized=True)
    with pytest.raises(RuntimeError):
        PP.FramePreprocessor(scale_wh=(10, 10), crop_wh=(12, 8))
