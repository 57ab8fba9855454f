"""-m gpu: the fused kernels either side of the forward path, through the C ABI, against the oracle and the golden
vectors of the real reference code (SURVEY 8f-1 ede_depth_epilogue, 8f-2 ede_preprocess_frames)."""
import os

import numpy as np
import pytest
import torch

from efficientdepthestimation_b200 import loss as L
from efficientdepthestimation_b200 import preprocess as PP
from oracle import preprocess_cpu as P
from oracle import reside_cpu as rc
from oracle import weights as W

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


# ------------------------------------------------------------------------------------------ preprocessing (bit-exact)
@pytest.mark.parametrize("case", ["down2", "up", "down4p5", "same_h"])
def test_preprocess_matches_reference_golden_bit_exact(cuda, lib, case):
    g = np.load(os.path.join(GOLDEN, "preprocess_frames.npz"))
    sw, sh, cw, ch = (int(v) for v in g[case + "_cfg"])
    pre = PP.FramePreprocessor((sw, sh), (cw, ch))
    y = pre(torch.from_numpy(g[case + "_frames"]).to(cuda))
    torch.cuda.synchronize()
    assert y.dtype == torch.float32 and y.is_contiguous()
    assert np.array_equal(y.cpu().numpy(), g[case + "_out"])


@pytest.mark.parametrize("n,hw,band", [(3, (480, 640), 4), (1, (1080, 1920), 4), (2, (240, 320), 7), (1, (228, 304), 1),
                                       (2, (333, 517), 5)])
def test_preprocess_production_geometry_vs_oracle(cuda, lib, n, hw, band):
    """640x480 -> 320x240 -> 304x228 (inference_benchmark.py:46-52) and other source sizes; ragged last band."""
    rng = np.random.default_rng(hw[0] + n)
    frames = rng.integers(0, 256, (n,) + hw + (3,), dtype=np.uint8)
    frames[0, :7] = 255
    frames[-1, -5:] = 0
    y = PP.FramePreprocessor(band=band)(torch.from_numpy(frames).to(cuda))
    torch.cuda.synchronize()
    assert tuple(y.shape) == (n, 3, 228, 304)
    assert np.array_equal(y.cpu().numpy(), P.preprocess_frames(frames))


def test_preprocess_properties_full_batch(cuda, lib):
    """Batch 64 of 640x480 frames: constant frames stay constant ((c/255-mean)/std exactly), frames are independent
    (batch == per-frame), and a vertical flip of the input flips the output (the filter is symmetric)."""
    n, h, w = 64, 480, 640
    g = torch.Generator().manual_seed(0)
    frames = torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8, generator=g)
    frames[1] = 200
    pre = PP.FramePreprocessor()
    d = frames.to(cuda)
    y = pre(d)
    for c in range(3):
        want = (np.float32(200) / np.float32(255) - np.float32(P.IMAGENET_MEAN[c])) / np.float32(P.IMAGENET_STD[c])
        assert bool((y[1, c] == float(want)).all())
    assert torch.equal(pre(d[5:6]), y[5:6])
    assert torch.equal(pre(d[7:8].flip(1)), y[7:8].flip(2))
    assert np.array_equal(y[63].cpu().numpy(), P.preprocess_frames(frames[63:64].numpy())[0])


def test_preprocess_feeds_the_forward(cuda, lib):
    from efficientdepthestimation_b200 import factory
    m = factory.build_model("enb0-hu").cuda().eval()
    frames = torch.randint(0, 256, (2, 480, 640, 3), dtype=torch.uint8, generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        out = m(PP.FramePreprocessor()(frames.to(cuda)))
    assert tuple(out.shape) == (2, 1, 114, 152) and bool(torch.isfinite(out).all())


# ------------------------------------------------------------------------------------------ post-forward epilogue
@pytest.mark.parametrize("case", ["nyu", "holes", "same"])
def test_epilogue_matches_reference_golden(cuda, lib, case):
    g = np.load(os.path.join(GOLDEN, "depth_epilogue.npz"))
    pred, label = torch.from_numpy(g[case + "_pred"]).to(cuda), torch.from_numpy(g[case + "_label"]).to(cuda)
    r = L.depth_epilogue(pred, label=label, want_resized=True, want_depth_mm=True, want_edges=True)
    torch.cuda.synchronize()
    resized, mm = r["resized"].cpu(), r["depth_mm"].cpu().numpy()
    ref = torch.from_numpy(g[case + "_resized"])
    assert torch.allclose(resized, ref, rtol=1e-6, atol=1e-6)
    # the conversion itself is exact on the map the kernel produced ...
    assert np.array_equal(mm, rc.depth_to_mm_u16(resized))
    # ... and against the reference's map it can only differ where the last ulp of the resize crosses an integer mm
    # or the 10 m cap (ATen's CPU kernel orders the four products differently)
    diff = np.abs(mm.astype(np.int64) - g[case + "_mm"].astype(np.int64))
    near_cap = np.abs(g[case + "_resized"][:, 0] - 10.0) < 1e-5
    assert bool(((diff <= 1) | near_cap).all()) and float((diff > 0).mean()) < 1e-3
    assert r["edges"].cpu().tolist() == list(g[case + "_edges"])
    sums = r["sums"].cpu().tolist()
    assert sums[12] == label.numel() and sums[13] == label.shape[0]
    m = L.metrics_from_sums(sums)
    for k, v in zip(("mae", "mse", "rmse", "abs_rel", "log10", "delta1", "delta2", "delta3"), g[case + "_metrics"]):
        assert (m[k] != m[k] and v != v) or abs(m[k] - v) <= 1e-5 * max(1.0, abs(v)), (k, m[k], v)


def test_epilogue_equals_unfused_kernels(cuda, lib):
    """Fused pass == ede_resize_planes_f32 -> ede_sobel_loss_metrics / ede_edge_counts (same expression trees):
    resized map bit-identical, counters exact, fp64 sums equal up to the order of the atomics."""
    from efficientdepthestimation_b200 import _cabi as A
    n, (h, w), (oh, ow) = 32, (114, 152), (228, 304)                     # BASELINE config-4 size
    pred = (0.3 + 10.5 * torch.rand(n, 1, h, w, generator=torch.Generator().manual_seed(2))).to(cuda)
    label = W.make_labels(n, oh, ow, seed=1).to(cuda)
    r = L.depth_epilogue(pred, label=label, want_resized=True, want_depth_mm=True, want_edges=True)
    up = torch.empty(n, 1, oh, ow, device=cuda)
    A.check(lib.ede_resize_planes_f32(pred.data_ptr(), up.data_ptr(), n, h, w, oh, ow, torch.cuda.current_stream().cuda_stream))
    s2, e2 = L.loss_metric_sums(up, label), L.edge_counts(up, label, 0.25)
    torch.cuda.synchronize()
    assert torch.equal(r["resized"], up)
    assert torch.equal(r["edges"], e2) and torch.equal(r["sums"][8:14], s2[8:14])
    assert torch.allclose(r["sums"], s2, rtol=1e-12, atol=0)
    assert np.array_equal(r["depth_mm"].cpu().numpy(), rc.depth_to_mm_u16(up.cpu()))
    # oracle on the same inputs (train.py:216-236 loss terms at the resized size, util.py metrics)
    ref_up = rc.post_resize(pred.cpu(), (oh, ow))
    t = rc.sobel_loss_terms(ref_up, label.cpu())
    mine = L.loss_from_sums(r["sums"].cpu().tolist())
    for k in ("depth", "normal", "dx", "dy", "loss"):
        assert abs(mine[k] - float(t[k])) <= 2e-6 * max(1.0, abs(float(t[k]))), k
    # outputs are optional and independent: 16-bit map only (test_nyu.py path, no label), accumulation into `sums`
    only = L.depth_epilogue(pred, size=(480, 640), want_depth_mm=True)
    assert set(only) == {"depth_mm"} and tuple(only["depth_mm"].shape) == (n, 480, 640)
    want = rc.depth_to_mm_u16(rc.post_resize(pred[:2].cpu(), (480, 640))).astype(np.int64)
    d = np.abs(only["depth_mm"][:2].cpu().numpy().astype(np.int64) - want)
    assert float((d > 1).mean()) < 1e-4 and float((d > 0).mean()) < 1e-2      # last-ulp crossings of an integer mm / the cap
    acc = L.depth_epilogue(pred[:16], label=label[:16])["sums"]
    acc = L.depth_epilogue(pred[16:], label=label[16:], sums=acc)["sums"]
    assert torch.equal(acc[8:14], s2[8:14]) and torch.allclose(acc, s2, rtol=1e-12)


def test_epilogue_edge_cases_and_errors(cuda, lib):
    pred = torch.tensor([[[[1.5, -0.0007], [10.0, 65.0]]]], device=cuda)
    r = L.depth_epilogue(pred, size=(2, 2), want_depth_mm=True, want_resized=True)
    # -0.0007 * 1000 = -0.7 -> int 0; exactly 10 m is kept (the cap is a strict >); 65 m > cap -> 0
    assert torch.equal(r["resized"], pred) and r["depth_mm"].cpu().tolist() == [[[1500, 0], [10000, 0]]]
    # a NaN poisons its bilinear neighbours even at weight 0 (0 * NaN), exactly as ATen does; NaN -> 0 in the cast
    pred_nan = torch.tensor([[[[1.5, 2.5], [float("nan"), 3.5]]]], device=cuda)
    rn = L.depth_epilogue(pred_nan, size=(2, 2), want_depth_mm=True, want_resized=True)
    assert bool(torch.isnan(rn["resized"][0, 0, :, 0]).all()) and rn["depth_mm"][0, :, 0].cpu().tolist() == [0, 0]
    assert rn["depth_mm"][0, 1, 1].item() == 3500
    one = L.depth_epilogue(torch.full((1, 1, 1, 1), 2.0, device=cuda), size=(5, 3), want_resized=True)["resized"]
    assert bool((one == 2.0).all())
    with pytest.raises(RuntimeError):
        L.depth_epilogue(pred, size=(4, 4))                                         # nothing requested
    with pytest.raises(RuntimeError):
        L.depth_epilogue(pred, size=(4, 4), want_edges=True)                        # edges need a label
    with pytest.raises(RuntimeError):
        L.depth_epilogue(pred, size=(4, 4), label=torch.zeros(1, 1, 3, 3, device=cuda))


@pytest.mark.parametrize("n,size", [(4, (480, 640)), (32, (228, 304)), (3, (231, 299))])
def test_depth_mm_bit_exact_against_the_reference_lines_on_the_gpu(cuda, lib, n, size):
    """ReSIDE/test_nyu.py:84-88 runs `F.interpolate(outputs, size, 'bilinear', align_corners=True)` ON THE GPU, then caps
    and converts on the host.  Integer outputs are held to bit-exactness: the fused epilogue's resized map must equal
    ATen's CUDA upsample_bilinear2d bit for bit (same expression order: src = scale * dst in fp32, lambdas in fp32,
    `h0 * (w0 * a + w1 * b) + h1 * (w0 * c + w1 * d)` with nvcc's FMA contraction), so the uint16 millimetre maps are
    equal.  torch on the GPU is the TEST ORACLE here; the product path never calls it."""
    import torch.nn.functional as F
    pred = (0.3 + 10.5 * torch.rand(n, 1, 114, 152, generator=torch.Generator().manual_seed(n))).to(cuda)
    r = L.depth_epilogue(pred, size=size, want_resized=True, want_depth_mm=True)
    outputs = F.interpolate(pred, size=size, mode="bilinear", align_corners=True)          # test_nyu.py:85
    torch.cuda.synchronize()
    assert torch.equal(r["resized"], outputs), float((r["resized"] - outputs).abs().max())
    out = outputs.squeeze().cpu().numpy()                                                    # :86
    out[out > 10.0] = 0.0                                                                    # :87  cap at 10 m
    out = (out * 1000.0).astype(np.uint16)                                                   # :88
    assert np.array_equal(r["depth_mm"].cpu().numpy(), out)
    only = L.depth_epilogue(pred, size=size, want_depth_mm=True)["depth_mm"]                 # the label-free kernel
    assert np.array_equal(only.cpu().numpy(), out)
