"""-m gpu: every C-ABI operator against the same op in plain PyTorch fp32 on seeded inputs
(called through ctypes exactly as the engine does), plus Sobel/loss/metrics against the golden vectors."""
import ctypes as C
import os

import pytest
import torch
import torch.nn.functional as F

from efficientdepthestimation_b200 import _cabi as A
from efficientdepthestimation_b200 import loss as L
from efficientdepthestimation_b200 import prepack as P
from oracle import reside_cpu as rc
from oracle import weights as W

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
TDT = {A.F32: torch.float32, A.BF16: torch.bfloat16}


def stream():
    return torch.cuda.current_stream().cuda_stream


def nhwc(x_nchw, c_p, dtype):
    n, c, h, w = x_nchw.shape
    out = torch.zeros(n, h, w, c_p, dtype=torch.float32, device=x_nchw.device)
    out[..., :c] = x_nchw.permute(0, 2, 3, 1)
    return out.to(TDT[dtype]).contiguous()


def split_store(x_nhwc_f32):
    """fp32 NHWC -> EDE_BF16X2 bytes: per 8 lanes [8 hi][8 lo] (as a bf16 tensor with 2x lanes)."""
    n, h, w, c = x_nhwc_f32.shape
    hi = x_nhwc_f32.to(torch.bfloat16)
    lo = (x_nhwc_f32 - hi.float()).to(torch.bfloat16)
    return torch.stack((hi.view(n, h, w, c // 8, 8), lo.view(n, h, w, c // 8, 8)), dim=4).reshape(n, h, w, 2 * c).contiguous()


def split_load(t_bf16, c):
    n, h, w, _ = t_bf16.shape
    v = t_bf16.float().view(n, h, w, c // 8, 2, 8)
    return (v[..., 0, :] + v[..., 1, :]).reshape(n, h, w, c)


CONV_CASES = [
    # name, n, h, w, cin, cout, k, s, (pt, pl, pb, pr), res, act
    ("1x1", 3, 15, 19, 256, 136, 1, 1, (0, 0, 0, 0), True, A.ACT_RELU),
    ("1x1_swish_ragged", 2, 57, 76, 24, 144, 1, 1, (0, 0, 0, 0), False, A.ACT_SWISH),
    ("3x3", 2, 29, 38, 64, 64, 3, 1, (1, 1, 1, 1), False, A.ACT_RELU),
    ("5x5_ragged_c", 1, 114, 152, 74, 74, 5, 1, (2, 2, 2, 2), False, A.ACT_RELU),
    ("5x5_wide_k", 1, 33, 41, 320, 32, 5, 1, (2, 2, 2, 2), False, A.ACT_RELU),
    ("5x5_head_bias", 2, 60, 80, 128, 1, 5, 1, (2, 2, 2, 2), False, A.ACT_NONE),
    ("3x3s2", 2, 57, 76, 128, 128, 3, 2, (1, 1, 1, 1), False, A.ACT_RELU),
    ("1x1s2", 2, 57, 76, 256, 512, 1, 2, (0, 0, 0, 0), False, A.ACT_NONE),
    ("7x7s2_stem", 2, 228, 304, 3, 64, 7, 2, (3, 3, 3, 3), False, A.ACT_RELU),
    ("3x3s2_asym_stem", 2, 228, 304, 3, 32, 3, 2, (0, 0, 1, 1), False, A.ACT_NONE),
    ("5x5s2_asym", 1, 57, 76, 144, 40, 5, 2, (1, 1, 2, 2), False, A.ACT_NONE),
    # kernel-row taps packed on the 128 M rows of the operand-swapped form (EdeConv.w_pack_g = 4 / 2)
    ("5x5_24_to_32_pack4", 2, 57, 76, 24, 32, 5, 1, (2, 2, 2, 2), False, A.ACT_RELU),
    ("5x5_64_to_64_pack2", 2, 40, 50, 64, 64, 5, 1, (2, 2, 2, 2), False, A.ACT_RELU),
    ("3x3_40_to_40_pack2_res", 2, 28, 38, 40, 40, 3, 1, (1, 1, 1, 1), True, A.ACT_RELU),
    ("3x3_16_to_8_pack4_wide", 1, 114, 152, 16, 8, 3, 1, (1, 1, 1, 1), True, A.ACT_SWISH),
    ("5x5_74_to_1_pack4_bias", 1, 114, 152, 74, 1, 5, 1, (2, 2, 2, 2), False, A.ACT_NONE),
    ("5x5_74_to_74_k64_plus_16", 2, 57, 76, 74, 74, 5, 1, (2, 2, 2, 2), False, A.ACT_RELU),     # K = 64 + 16
    ("3x3_90_to_40_k64_plus_32_pack2", 2, 40, 50, 90, 40, 3, 1, (1, 1, 1, 1), True, A.ACT_RELU),  # K = 64 + 32
]


@pytest.mark.parametrize("mode", ["bf16", "split", "fp32"])
@pytest.mark.parametrize("case", CONV_CASES, ids=[c[0] for c in CONV_CASES])
def test_conv2d(cuda, lib, case, mode):
    _run_conv_case(cuda, lib, case, mode)


# narrow stride-1 k x k convs in the N-packed pixels-on-M form (EDE_PACK_N | G: the taps of a kernel row on the UMMA N
# axis, lane-shifted sums in the epilogue, weights resident in shared memory), forced through EDE_UMMA_NPACK=1 so that
# the form is exercised whatever the tile picker's time model prefers
NPACK_CASES = [
    ("np_3x3_32_to_16", 2, 57, 76, 32, 16, 3, 1, (1, 1, 1, 1), False, A.ACT_RELU),          # conv1_2: G = 3, N = 48
    ("np_5x5_24_to_32", 2, 57, 76, 24, 32, 5, 1, (2, 2, 2, 2), False, A.ACT_RELU),          # D.up4: G = 5, N = 160
    ("np_5x5_24_to_48_res", 1, 41, 67, 24, 48, 5, 1, (2, 2, 2, 2), True, A.ACT_RELU),       # G = 5, N = 240, residual
    ("np_3x3_40_to_40_swish", 2, 28, 38, 40, 40, 3, 1, (1, 1, 1, 1), True, A.ACT_SWISH),    # G = 3, N = 144
    ("np_3x3_16_to_90_g2", 1, 30, 61, 16, 90, 3, 1, (1, 1, 1, 1), False, A.ACT_RELU),       # G = 2 < kw: 2 steps per kernel row
    ("np_3x3_90_to_24_two_chunks", 2, 20, 33, 90, 24, 3, 1, (1, 1, 1, 1), False, A.ACT_NONE),   # K = 64 + 32: two stages
    ("np_4x4_asym_16_to_8", 1, 37, 45, 16, 8, 4, 1, (1, 1, 2, 2), False, A.ACT_RELU),       # even kernel, asymmetric pad
]


@pytest.mark.parametrize("mode", ["bf16", "split"])
@pytest.mark.parametrize("case", NPACK_CASES, ids=[c[0] for c in NPACK_CASES])
def test_conv2d_taps_on_n(cuda, lib, case, mode, monkeypatch):
    monkeypatch.setenv("EDE_UMMA_NPACK", "1")
    cfg = _run_conv_case(cuda, lib, case, mode)
    if mode == "split" and not cfg.pack_g & P.PACK_N:
        pytest.skip("hi+lo weights over doubled lanes do not fit the resident-weight budget of this form")
    assert cfg.pack_g & P.PACK_N and cfg.pitch == 32 and cfg.swap == 0, (cfg.pack_g, cfg.pitch, cfg.swap)


def _run_conv_case(cuda, lib, case, mode):
    name, n, h, w, cin, cout, k, s, (pt, pl, pb, pr), use_res, act = case
    g = torch.Generator().manual_seed(len(name) * 131 + k)
    oh, ow = (h + pt + pb - k) // s + 1, (w + pl + pr - k) // s + 1
    cin_p, cout_p = P.ceil8(cin), P.ceil8(cout)
    x = torch.randn(n, cin, h, w, generator=g).to(cuda)
    wt = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).to(cuda)
    scale = (0.5 + torch.rand(cout, generator=g)).to(cuda)
    shift = (0.2 * torch.randn(cout, generator=g)).to(cuda)
    res = torch.randn(n, cout, oh, ow, generator=g).to(cuda) if use_res else None
    io_dt = {"bf16": A.BF16, "split": A.BF16X2, "fp32": A.F32}[mode]
    if mode == "split":
        xin = split_store(nhwc(x, cin_p, A.F32))
        in_lanes = [v for g8 in range(0, cin_p, 8) for v in P.identity_map(cin, cin_p)[g8:g8 + 8] * 2]
        mul, x_dt, passes = 2, A.BF16, 2
        x_eff = split_load(xin, cin_p)[..., :cin].permute(0, 3, 1, 2)
        w_eff = wt
        resin = split_store(nhwc(res, cout_p, A.F32)) if use_res else None
        res_eff = split_load(resin, cout_p)[..., :cout].permute(0, 3, 1, 2) if use_res else None
        y = torch.full((n, oh, ow, 2 * cout_p), float("nan"), device=cuda, dtype=torch.bfloat16)
    else:
        xin = nhwc(x, cin_p, io_dt)
        in_lanes, mul, x_dt, passes = P.identity_map(cin, cin_p), 1, io_dt, 1
        x_eff = xin.float()[..., :cin].permute(0, 3, 1, 2)
        w_eff = wt.to(torch.bfloat16).float() if mode == "bf16" else wt
        resin = nhwc(res, cout_p, io_dt) if use_res else None
        res_eff = resin.float()[..., :cout].permute(0, 3, 1, 2) if use_res else None
        y = torch.full((n, oh, ow, cout_p), float("nan"), device=cuda, dtype=TDT[io_dt])
    full = P.scatter_weight(wt, in_lanes, P.identity_map(cout, cout_p))
    w_direct = P.pack_direct(full)
    w_umma, co16, ci64 = P.pack_umma(full, passes)
    pack_g = 1
    sc = P.place_lanes(scale, P.identity_map(cout, cout_p))
    sf = P.place_lanes(shift, P.identity_map(cout, cout_p))
    a = A.EdeConv(x=xin.data_ptr(), y=y.data_ptr(), res=resin.data_ptr() if use_res else None, scale=sc.data_ptr(),
                  shift=sf.data_ptr(), w_umma=w_umma.data_ptr(), w_direct=w_direct.data_ptr(), n=n, h=h, w=w,
                  cin_p=cin_p * mul, x_cs=cin_p * mul, x_co=0, oh=oh, ow=ow, cout_p=cout_p, y_cs=cout_p, y_co=0,
                  res_cs=cout_p, res_co=0, kh=k, kw=k, sh=s, sw=s, pt=pt, pl=pl, act=act, act_c_end=cout_p,
                  x_dtype=x_dt, y_dtype=io_dt, x_layout=A.NHWC, relu_in=0, w_umma_cout_p=co16, w_umma_cin_p=ci64,
                  force_direct=1 if mode == "fp32" else 0, w_passes=passes, w_pack_g=-1)
    # as engine.Graph.conv does: ask the tile picker which tap packing it wants, prepack for exactly that
    q = A.umma_config(a)
    pack_g = int(q.pack_g) if q.eligible and q.pack_g > 1 else 1
    if pack_g > 1:
        w_umma, co16, ci64 = (P.pack_umma_ntaps(full, passes, pack_g & 15) if pack_g & P.PACK_N else
                              P.pack_umma_taps(full, passes, pack_g))
        a.w_umma, a.w_umma_cout_p, a.w_umma_cin_p = w_umma.data_ptr(), co16, ci64
    a.w_pack_g = pack_g
    cfg = A.umma_config(a)
    assert bool(cfg.eligible) == (mode != "fp32") and (not cfg.eligible or cfg.pack_g == pack_g)
    A.check(lib.ede_conv2d(C.byref(a), stream()), "ede_conv2d")
    torch.cuda.synchronize()
    ref = F.conv2d(F.pad(x_eff.double(), (pl, pr, pt, pb)), w_eff.double(), None, s)
    ref = ref * scale.double().view(1, -1, 1, 1) + shift.double().view(1, -1, 1, 1)
    if use_res:
        ref = ref + res_eff.double()
    ref = {A.ACT_RELU: F.relu, A.ACT_SWISH: lambda t: t * torch.sigmoid(t), A.ACT_NONE: lambda t: t}[act](ref)
    ref = ref.permute(0, 2, 3, 1).float()
    got = split_load(y, cout_p) if mode == "split" else y.float()
    assert not torch.isnan(got).any()
    assert float(got[..., cout:].abs().sum()) == 0.0                      # pad lanes stay zero
    err = (got[..., :cout] - ref).abs()
    # bf16 output rounding is 2^-9 relative; split keeps 16 bits; fp32 path is exact up to summation order.
    # The tensor core adds into its fp32 accumulator with truncation: a bias of ~2^-24 per K=16 step that
    # grows linearly with the K extent (measured: 2e-4 at K = 2 x 8000) -- budgeted explicitly here.
    rel = {"bf16": 4e-3, "split": 6e-5 + 2.0e-8 * cin * k * k, "fp32": 2e-5}[mode]
    tol = rel * ref.abs() + rel
    assert bool((err <= tol).all()), (name, mode, float(err.max()), float((err / (ref.abs() + 1)).max()))
    return cfg


@pytest.mark.parametrize("c", [44, 90, 150])      # 48 lanes (one 64-chunk), 96 (three 32-lane chunks), 152 (64+64+24)
@pytest.mark.parametrize("dt", [A.F32, A.BF16, A.BF16X2, A.F32_TO_BF16X2])
@pytest.mark.parametrize("k,s,pad", [(3, 1, (1, 1, 1, 1)), (3, 2, (0, 0, 1, 1)), (5, 2, (1, 1, 2, 2)), (5, 1, (2, 2, 2, 2))])
def test_dwconv_se(cuda, lib, dt, k, s, pad, c):
    """Depthwise + BN + swish + SE squeeze, SE FCs and the gate multiply (App. A.3 of SURVEY.md)."""
    g = torch.Generator().manual_seed(k * 10 + s)
    n, h, w, csq = 3, 29, 38, 6
    c_p = P.ceil8(c)
    pt, pl, pb, pr = pad
    oh, ow = (h + pt + pb - k) // s + 1, (w + pl + pr - k) // s + 1
    x = torch.randn(n, c, h, w, generator=g).to(cuda)
    wt = (torch.randn(c, 1, k, k, generator=g) / k).to(cuda)
    scale, shift = (0.5 + torch.rand(c, generator=g)).to(cuda), (0.2 * torch.randn(c, generator=g)).to(cuda)
    w1, b1 = torch.randn(csq, c, generator=g).to(cuda) / c ** 0.5, 0.1 * torch.randn(csq, generator=g).to(cuda)
    w2, b2 = torch.randn(c, csq, generator=g).to(cuda) / csq ** 0.5, 0.1 * torch.randn(c, generator=g).to(cuda)
    dw_dt = dt
    if dt == A.F32_TO_BF16X2:                        # fp32 lanes in (as an MBConv expand conv writes them), split bf16 out
        xin = nhwc(x, c_p, A.F32)
        x_eff = x
        y = torch.zeros(n, oh, ow, 2 * c_p, device=cuda, dtype=torch.bfloat16)
        z = torch.zeros_like(y)
        dt = A.BF16X2                                # everything downstream of the depthwise conv is split bf16
    elif dt == A.BF16X2:
        xin = split_store(nhwc(x, c_p, A.F32))
        x_eff = split_load(xin, c_p)[..., :c].permute(0, 3, 1, 2)
        y = torch.zeros(n, oh, ow, 2 * c_p, device=cuda, dtype=torch.bfloat16)
        z = torch.zeros_like(y)
    else:
        xin = nhwc(x, c_p, dt)
        x_eff = xin.float()[..., :c].permute(0, 3, 1, 2)
        y = torch.zeros(n, oh, ow, c_p, device=cuda, dtype=TDT[dt])
        z = torch.zeros_like(y)
    lanes = P.identity_map(c, c_p)
    wk = torch.zeros(k * k, c_p, device=cuda)
    wk[:, :c] = wt.reshape(c, k * k).t()
    sc, sf = P.place_lanes(scale, lanes), P.place_lanes(shift, lanes)
    pool = torch.zeros(n, c_p, device=cuda)
    a = A.EdeDwConv(x=xin.data_ptr(), y=y.data_ptr(), w=wk.data_ptr(), scale=sc.data_ptr(), shift=sf.data_ptr(),
                    pool_sum=pool.data_ptr(), n=n, h=h, w_=w, c_p=c_p, x_cs=c_p, x_co=0, oh=oh, ow=ow, y_cs=c_p,
                    y_co=0, k=k, s=s, pt=pt, pl=pl, act=A.ACT_SWISH, dtype=dw_dt)
    A.check(lib.ede_dwconv2d(C.byref(a), stream()), "ede_dwconv2d")
    w1p = torch.zeros(csq, c_p, device=cuda); w1p[:, :c] = w1
    w2p = torch.zeros(c_p, csq, device=cuda); w2p[:c] = w2
    b2p = torch.zeros(c_p, device=cuda); b2p[:c] = b2
    gate = torch.zeros(n, c_p, device=cuda)
    f = A.EdeSeFc(pool_sum=pool.data_ptr(), w1=w1p.data_ptr(), b1=b1.data_ptr(), w2=w2p.data_ptr(), b2=b2p.data_ptr(),
                  gate=gate.data_ptr(), n=n, c=c, c_p=c_p, csq=csq, inv_hw=1.0 / (oh * ow))
    A.check(lib.ede_se_fc(C.byref(f), stream()), "ede_se_fc")
    sa = A.EdeSeScale(x=y.data_ptr(), y=z.data_ptr(), gate=gate.data_ptr(), n=n, hw=oh * ow, c_p=c_p, x_cs=c_p, x_co=0,
                      y_cs=c_p, y_co=0, dtype=dt)
    A.check(lib.ede_se_scale(C.byref(sa), stream()), "ede_se_scale")
    torch.cuda.synchronize()
    ref = F.conv2d(F.pad(x_eff, (pl, pr, pt, pb)), wt, None, s, 0, 1, c)
    ref = ref * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)
    ref = ref * torch.sigmoid(ref)
    rel = {A.F32: 1e-5, A.BF16: 4e-3, A.BF16X2: 4e-5}[dt]
    got = (split_load(y, c_p) if dt == A.BF16X2 else y.float())[..., :c].permute(0, 3, 1, 2)
    assert bool(((got - ref).abs() <= rel * ref.abs() + rel).all())
    ref_pool = got.sum(dim=(2, 3))                       # the squeeze sees the stored (rounded) activations
    assert torch.allclose(pool[:, :c], ref_pool, rtol=1e-4, atol=1e-3)
    sref = ref_pool / (oh * ow)
    hid = sref @ w1.t() + b1
    hid = hid * torch.sigmoid(hid)
    gref = torch.sigmoid(hid @ w2.t() + b2)
    assert torch.allclose(gate[:, :c], gref, rtol=1e-4, atol=1e-5) and float(gate[:, c:].abs().sum()) == 0
    zgot = (split_load(z, c_p) if dt == A.BF16X2 else z.float())[..., :c].permute(0, 3, 1, 2)
    zref = got * gref.view(n, c, 1, 1)
    assert bool(((zgot - zref).abs() <= rel * zref.abs() + rel).all())
    # the form the engine emits: expand weights transposed ([csq][c_p]), squeeze sums cleared for the next forward
    gate2 = torch.zeros_like(gate)
    w2t = w2p.t().contiguous()
    f2 = A.EdeSeFc(pool_sum=pool.data_ptr(), w1=w1p.data_ptr(), b1=b1.data_ptr(), w2=w2t.data_ptr(), b2=b2p.data_ptr(),
                   gate=gate2.data_ptr(), n=n, c=c, c_p=c_p, csq=csq, inv_hw=1.0 / (oh * ow), clear_pool=1,
                   w2_transposed=1)
    A.check(lib.ede_se_fc(C.byref(f2), stream()), "ede_se_fc")
    torch.cuda.synchronize()
    assert torch.allclose(gate2, gate, rtol=1e-5, atol=1e-6) and float(pool.abs().sum()) == 0


@pytest.mark.parametrize("dt", [A.F32, A.BF16, A.BF16X2])
@pytest.mark.parametrize("k,s,cout,pad,act,hw", [(3, 2, 32, (0, 0, 1, 1), A.ACT_SWISH, (57, 76)),     # EfficientNet-B0 stem
                                                 (3, 2, 48, (0, 0, 1, 1), A.ACT_SWISH, (228, 304)),   # EfficientNet-B4 stem
                                                 (7, 2, 64, (3, 3, 3, 3), A.ACT_RELU, (228, 304)),    # ResNet conv1
                                                 (7, 2, 64, (3, 3, 3, 3), A.ACT_RELU, (37, 131))])
def test_stem_conv_from_nchw(cuda, lib, dt, k, s, cout, pad, act, hw):
    """First conv read straight from the fp32 NCHW image (resnet.py:103-106 / EfficientNet stem)."""
    g = torch.Generator().manual_seed(k * 100 + cout)
    n, cin, (h, w) = 3, 3, hw
    pt, pl, pb, pr = pad
    oh, ow = (h + pt + pb - k) // s + 1, (w + pl + pr - k) // s + 1
    assert lib.ede_stem_supported(k, s, cin, cout) == 1 and lib.ede_stem_supported(5, 1, cin, cout) == 0
    x = torch.randn(n, cin, h, w, generator=g).to(cuda)
    wt = (torch.randn(cout, cin, k, k, generator=g) / (k * cin ** 0.5)).to(cuda)
    scale, shift = (0.5 + torch.rand(cout, generator=g)).to(cuda), (0.2 * torch.randn(cout, generator=g)).to(cuda)
    wk = wt.permute(2, 3, 1, 0).reshape(k * k * cin, cout).contiguous()
    mul = 2 if dt == A.BF16X2 else 1
    cs = cout + 16                                                   # write into a slice of a wider buffer
    y = torch.full((n, oh, ow, cs * mul), 7.0, device=cuda, dtype=torch.bfloat16 if dt != A.F32 else torch.float32)
    a = A.EdeStem(x=x.data_ptr(), y=y.data_ptr(), w=wk.data_ptr(), scale=scale.data_ptr(), shift=shift.data_ptr(),
                  n=n, cin=cin, h=h, w_=w, oh=oh, ow=ow, cout_p=cout, y_cs=cs, y_co=8, k=k, s=s, pt=pt, pl=pl,
                  act=act, y_dtype=dt)
    A.check(lib.ede_stem_conv(C.byref(a), stream()), "ede_stem_conv")
    torch.cuda.synchronize()
    ref = F.conv2d(F.pad(x.double(), (pl, pr, pt, pb)), wt.double(), None, s).float()
    ref = ref * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1)
    ref = ref * torch.sigmoid(ref) if act == A.ACT_SWISH else torch.relu(ref)
    full = split_load(y, cs) if dt == A.BF16X2 else y.float()
    got = full[..., 8:8 + cout].permute(0, 3, 1, 2)
    rel = {A.F32: 2e-5, A.BF16: 4e-3, A.BF16X2: 4e-5}[dt]
    assert bool(((got - ref).abs() <= rel * ref.abs() + rel).all())
    untouched = torch.cat((full[..., :8], full[..., 8 + cout:]), dim=-1)      # neighbours of the slice stay as they were
    assert bool((untouched == (14.0 if dt == A.BF16X2 else 7.0)).all())


@pytest.mark.parametrize("dts", [(A.F32, A.F32), (A.BF16, A.BF16), (A.BF16X2, A.BF16)])
@pytest.mark.parametrize("hw,ohw", [((8, 10), (114, 152)), ((57, 76), (114, 152)), ((15, 19), (29, 38)),
                                    ((114, 152), (228, 304)), ((7, 9), (7, 9)), ((1, 1), (5, 3))])
def test_resize_bilinear_align_corners(cuda, lib, dts, hw, ohw):
    """Non-integer ratios (8->114, 10->152) as in modules.py:34; written into a channel SLICE of a wider buffer."""
    xd, yd = dts
    g = torch.Generator().manual_seed(hw[0] * 7 + ohw[0])
    n, c, (h, w), (oh, ow) = 2, 20, hw, ohw
    c_p, cs_out, co_out = 24, 48, 16
    x = torch.randn(n, c, h, w, generator=g).to(cuda)
    if xd == A.BF16X2:
        xin = split_store(nhwc(x, c_p, A.F32))
        x_eff = split_load(xin, c_p)[..., :c].permute(0, 3, 1, 2)
    else:
        xin = nhwc(x, c_p, xd)
        x_eff = xin.float()[..., :c].permute(0, 3, 1, 2)
    y = torch.full((n, oh, ow, cs_out), 7.0, device=cuda, dtype=TDT[yd])
    a = A.EdeResize(x=xin.data_ptr(), y=y.data_ptr(), n=n, h=h, w=w, c_p=c_p, x_cs=c_p, x_co=0, oh=oh, ow=ow,
                    y_cs=cs_out, y_co=co_out, x_dtype=xd, y_dtype=yd)
    A.check(lib.ede_resize_bilinear_ac(C.byref(a), stream()), "ede_resize_bilinear_ac")
    torch.cuda.synchronize()
    ref = F.interpolate(x_eff, size=(oh, ow), mode="bilinear", align_corners=True).permute(0, 2, 3, 1)
    got = y.float()
    assert float((got[..., :co_out] - 7.0).abs().max()) == 0 and float((got[..., co_out + c_p:] - 7.0).abs().max()) == 0
    tol = 1e-5 if yd == A.F32 else 4e-3
    assert bool(((got[..., co_out:co_out + c] - ref).abs() <= tol * ref.abs() + tol).all())
    # fp32 planes variant (post-forward depth resize, train.py:273-276)
    if xd == A.F32:
        yp = torch.empty(n, c, oh, ow, device=cuda)
        A.check(lib.ede_resize_planes_f32(x.contiguous().data_ptr(), yp.data_ptr(), n * c, h, w, oh, ow, stream()))
        torch.cuda.synchronize()
        assert torch.allclose(yp, F.interpolate(x, size=(oh, ow), mode="bilinear", align_corners=True), atol=1e-5)


@pytest.mark.parametrize("dt", [A.F32, A.BF16, A.BF16X2])
@pytest.mark.parametrize("hw", [(114, 152), (113, 151), (5, 4)])
def test_maxpool_and_eltwise_and_layout(cuda, lib, dt, hw):
    g = torch.Generator().manual_seed(hw[0])
    n, c, (h, w) = 2, 12, hw
    c_p = 16
    x = torch.randn(n, c, h, w, generator=g).to(cuda)
    # nchw -> nhwc (network input), checked by reading back through nhwc -> nchw
    elems = 2 * c_p if dt == A.BF16X2 else c_p
    store = torch.bfloat16 if dt != A.F32 else torch.float32
    xin = torch.full((n, h, w, elems), float("nan"), device=cuda, dtype=store)
    la = A.EdeLayout(src=x.data_ptr(), dst=xin.data_ptr(), n=n, c=c, h=h, w=w, c_p=c_p, cs=c_p, co=0, dtype=dt)
    A.check(lib.ede_nchw_to_nhwc(C.byref(la), stream()), "ede_nchw_to_nhwc")
    back = torch.empty(n, c, h, w, device=cuda)
    lb = A.EdeLayout(src=xin.data_ptr(), dst=back.data_ptr(), n=n, c=c, h=h, w=w, c_p=c_p, cs=c_p, co=0, dtype=dt)
    A.check(lib.ede_nhwc_to_nchw(C.byref(lb), stream()), "ede_nhwc_to_nchw")
    torch.cuda.synchronize()
    assert not torch.isnan(xin.float()).any()
    tol = {A.F32: 0.0, A.BF16: 4e-3, A.BF16X2: 2e-5}[dt]
    assert bool(((back - x).abs() <= tol * x.abs() + 1e-30).all())
    x_eff = back
    # 3x3 s2 p1 max pooling (resnet.py:107)
    oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    y = torch.zeros(n, oh, ow, elems, device=cuda, dtype=store)
    m = A.EdeMaxPool(x=xin.data_ptr(), y=y.data_ptr(), n=n, h=h, w=w, c_p=c_p, x_cs=c_p, x_co=0, oh=oh, ow=ow,
                     y_cs=c_p, y_co=0, dtype=dt)
    A.check(lib.ede_maxpool3x3s2(C.byref(m), stream()), "ede_maxpool3x3s2")
    # eltwise: relu(x) + x
    z = torch.zeros(n, h, w, elems, device=cuda, dtype=store)
    e = A.EdeEltwise(x=xin.data_ptr(), x2=xin.data_ptr(), y=z.data_ptr(), n_pix=n * h * w, c_p=c_p, x_cs=c_p, x_co=0,
                     x2_cs=c_p, x2_co=0, y_cs=c_p, y_co=0, act=A.ACT_RELU, dtype=dt)
    A.check(lib.ede_eltwise(C.byref(e), stream()), "ede_eltwise")
    torch.cuda.synchronize()
    dec = (lambda t: split_load(t, c_p)) if dt == A.BF16X2 else (lambda t: t.float())
    got = dec(y)[..., :c].permute(0, 3, 1, 2)
    assert torch.equal(got, F.max_pool2d(x_eff, 3, 2, 1))                 # max of stored values is exact
    gz = dec(z)[..., :c].permute(0, 3, 1, 2)
    rz = F.relu(x_eff) + x_eff
    assert bool(((gz - rz).abs() <= max(tol, 1e-7) * rz.abs() + 1e-30).all())


@pytest.mark.parametrize("holes", [False, True])
def test_sobel_loss_metrics_against_reference_golden(cuda, lib, holes):
    """ede_sobel / ede_sobel_loss_metrics vs the fixtures produced by ReSIDE/sobel.py, train.py:216-236 and
    util.MetricsTracker of the REAL reference (tests/golden/make_golden.py)."""
    g = torch.load(os.path.join(GOLDEN, f"sobel_loss_metrics{'_holes' if holes else ''}.pt"))
    out, lab = g["output"].to(cuda), g["labels"].to(cuda)
    from efficientdepthestimation_b200.ReSIDE.sobel import Sobel
    from efficientdepthestimation_b200.ReSIDE.util import MetricsTracker
    s = Sobel().to(cuda)(out)
    assert list(Sobel().state_dict().keys()) == ["edge_conv.weight"]
    assert torch.allclose(s.cpu(), g["sobel"], atol=1e-5)
    sums = L.loss_metric_sums(out, lab).cpu().tolist()
    assert sums[12] == out.numel() and sums[13] == out.shape[0]
    if not holes:
        mine = L.loss_from_sums(sums)
        for k, v in g["loss"].items():
            assert abs(mine[k] - v) <= 2e-6 * max(1.0, abs(v)), (k, mine[k], v)
        assert abs(float(L.sobel_loss(out, lab)) - g["loss"]["loss"]) <= 1e-5
    mt = MetricsTracker()
    mt.update(outputs=out, labels=lab)
    for k, v in g["metrics"].items():
        mine = mt.to_dict()[k]
        assert (mine != mine and v != v) or abs(mine - v) <= 1e-5 * max(1.0, abs(v)), (k, mine, v)


def test_loss_full_size_properties(cuda, lib):
    """BASELINE config-4 size (32 x 228x304): linearity of the sums over the batch axis (two half batches
    accumulate to the whole), exactness of the counters, and agreement with the oracle on one image."""
    n, h, w = 32, 228, 304
    out = (0.5 + 9.5 * torch.rand(n, 1, h, w, generator=torch.Generator().manual_seed(3))).to(cuda)
    lab = W.make_labels(n, h, w, seed=1, holes=True).to(cuda)
    whole = L.loss_metric_sums(out, lab)
    acc = L.loss_metric_sums(out[:16], lab[:16])
    acc = L.loss_metric_sums(out[16:], lab[16:], out=acc)
    torch.cuda.synchronize()
    a, b = whole.cpu(), acc.cpu()
    assert torch.equal(a[8:14], b[8:14]) and float(a[12]) == n * h * w and float(a[13]) == n
    assert float(a[11]) == float((~torch.isnan(lab)).sum())
    assert torch.allclose(a[[6, 7]], b[[6, 7]], rtol=1e-12)          # masked sums stay finite
    assert bool(torch.isnan(a[:6]).all()) and bool(torch.isnan(b[:6]).all())   # NaN labels poison the unmasked sums (util.py:54-57)
    lab0 = W.make_labels(n, h, w, seed=1).to(cuda)
    c0 = L.loss_metric_sums(out, lab0).cpu()
    c1 = L.loss_metric_sums(out[:16], lab0[:16])
    c1 = L.loss_metric_sums(out[16:], lab0[16:], out=c1).cpu()
    assert torch.allclose(c0[:8], c1[:8], rtol=1e-12) and torch.equal(c0[8:14], c1[8:14])
    one = L.loss_metric_sums(out[:1], lab[:1]).cpu().tolist()
    ref = rc.metric_sums(out[:1].cpu(), lab[:1].cpu())
    assert abs(one[6] - ref["rel"]) <= 1e-5 * ref["rel"] and one[8] == ref["d1"] and one[10] == ref["d3"]
    ec = L.edge_counts(out[:2], lab[:2], 0.25).cpu().tolist()
    po, pl_ = rc.edge_magnitude(out[:2].cpu()) > 0.25, rc.edge_magnitude(torch.nan_to_num(lab[:2].cpu(), nan=float("nan"))) > 0.25
    assert ec[0] == float(po.sum()) and ec[1] == float(pl_.sum()) and ec[2] == float((po & pl_).sum())


@pytest.mark.parametrize("hw,ohw,cin,couts,split", [((8, 10), (114, 152), 72, (16, 16), True),
                                                    ((57, 76), (114, 152), 24, (16, 16), False),
                                                    ((14, 19), (57, 76), 40, (8, 8), True),
                                                    ((3, 4), (9, 7), 16, (8,), False)])
@pytest.mark.parametrize("z_bf16", [0, 1])
def test_upconv_lowres_equals_conv_of_upsampled(cuda, lib, hw, ohw, cin, couts, split, z_bf16, monkeypatch):
    """engine.Graph.upconv (1x1 GEMM on the low-res grid + ede_upconv_taps) == relu/bn(conv5x5(bilinear_ac(x)))
    as _UpProjection computes it (modules.py:34-37): every tap, border and non-integer ratio (8->114).
    z_bf16 = 0: the per-tap products z stay fp32 (error = bf16 weights + bf16 output only); 1 (the plans' default): z is
    stored as bf16 -- every one of the 25 tap terms carries a 2^-9 rounding and the terms cancel in the sum, so the
    bound is set against the magnitude of the terms, not of the result (model-level effect: 3.29e-3 -> 3.34e-3)."""
    from efficientdepthestimation_b200 import engine
    monkeypatch.setenv("EDE_UPCONV_Z_BF16", str(z_bf16))
    g = torch.Generator().manual_seed(hw[0] * 31 + cin)
    n, (h, w), (oh, ow) = 2, hw, ohw
    x = torch.randn(n, cin, h, w, generator=g)
    parts, refs = [], []
    for i, co in enumerate(couts):
        wt = torch.randn(co, cin, 5, 5, generator=g) / (cin * 25) ** 0.5
        bn = torch.nn.BatchNorm2d(co)
        bn.weight.data = 0.5 + torch.rand(co, generator=g)
        bn.bias.data = 0.2 * torch.randn(co, generator=g)
        bn.running_mean = 0.1 * torch.randn(co, generator=g)
        bn.running_var = 0.5 + torch.rand(co, generator=g)
        import copy
        refs.append((wt, copy.deepcopy(bn).double().eval(), i == 0))
        parts.append((wt.to(cuda), bn.to(cuda).eval(), None, i == 0))
    gr = engine.Graph(cuda, "bf16")
    if not split:
        gr.encoder_dtype = A.BF16
    xin = gr.input_nchw(x.shape)
    out = gr.upconv(xin, (oh, ow), parts, (2, 2, 2, 2), A.ACT_RELU)
    gr.output_nchw(out)
    plan = gr.finalize()
    (y,) = plan.run([x.to(cuda)])
    torch.cuda.synchronize()
    x_eff = x if split else x.to(torch.bfloat16).float()
    up = F.interpolate(x_eff.double(), size=(oh, ow), mode="bilinear", align_corners=True)
    outs = []
    with torch.no_grad():
        for wt, bn, activated in refs:
            r = bn(F.conv2d(up, wt.to(torch.bfloat16).double(), None, 1, 2))
            outs.append(F.relu(r) if activated else r)
    ref = torch.cat(outs, 1).float()
    got = y.cpu()
    assert got.shape == ref.shape
    # output stored as bf16 (2^-9) on top of bf16 weights (identical on both sides)
    atol = 5e-3 if not z_bf16 else 5e-3 + 2.0 ** -9 * 5.0 * float(ref.abs().max())     # 25 terms, each <= ~max|ref|
    assert bool(((got - ref).abs() <= 5e-3 * ref.abs() + atol).all()), float((got - ref).abs().max())


@pytest.mark.parametrize("cout,cin,k,passes,pack_g,use_maps,use_bn,use_bias", [
    (74, 74, 5, 1, 1, False, True, False),        # R.conv0: identity lanes, BN
    (136, 250, 1, 2, 1, True, True, False),       # split-precision encoder 1x1 with a scattered lane order
    (16, 48, 3, 1, 4, True, False, True),         # 4 kernel-row taps on the 128 M rows, conv bias, no BN
    (40, 40, 5, 2, 2, False, True, True),         # 2 taps, hi+lo passes, BN + bias
    (1, 74, 5, 1, 1, False, False, True),         # R.conv2: one output channel + bias
    (16, 32, 3, 1, 16 | 3, False, True, False),   # conv1_2: the 3 taps of a kernel row on the UMMA N axis (EDE_PACK_N | 3)
    (30, 24, 5, 2, 16 | 5, False, True, True),    # 5 taps x 32 columns, hi+lo passes
    (74, 74, 5, 1, 16 | 3, False, False, False),  # R.conv0: 3 taps x 80 columns, 2 kernel-row steps
])
def test_prepack_conv_behind_the_c_abi(cuda, lib, cout, cin, k, passes, pack_g, use_maps, use_bn, use_bias):
    """ede_prepack_conv_to / ede_prepack_conv / ede_free: OIHW weight + BatchNorm vectors -> w_umma / w_direct / scale /
    shift, bit-exact against the host packing functions (prepack.py) that the round-1 engine used."""
    g = torch.Generator().manual_seed(cout * 131 + cin)
    w = torch.randn(cout, cin, k, k, generator=g)
    cin_p, cout_p = P.ceil8(cin) + (8 if use_maps else 0), P.ceil8(cout) + (8 if use_maps else 0)
    if pack_g > 1:
        cout_p = P.ceil8(cout)
    if use_maps:
        in_map = torch.full((cin_p,), -1, dtype=torch.int32)
        in_map[torch.randperm(cin_p, generator=g)[:cin]] = torch.randperm(cin, generator=g).int()
        out_map = torch.full((cout_p,), -1, dtype=torch.int32)
        out_map[torch.randperm(cout_p, generator=g)[:cout]] = torch.randperm(cout, generator=g).int()
    else:
        in_map = torch.tensor(P.identity_map(cin, cin_p), dtype=torch.int32)
        out_map = torch.tensor(P.identity_map(cout, cout_p), dtype=torch.int32)
    bn = dict(weight=0.5 + torch.rand(cout, generator=g), bias=torch.randn(cout, generator=g),
              mean=torch.randn(cout, generator=g), var=0.1 + torch.rand(cout, generator=g), eps=1e-3) if use_bn else None
    bias = torch.randn(cout, generator=g) if use_bias else None
    # expected: the host functions
    full = P.scatter_weight(w, in_map.tolist(), out_map.tolist())
    exp_umma, co16, ci64 = (P.pack_umma_ntaps(full, passes, pack_g & 15) if pack_g & P.PACK_N else
                            P.pack_umma_taps(full, passes, pack_g) if pack_g > 1 else P.pack_umma(full, passes))
    exp_direct = P.pack_direct(full)
    sc, sf = P.fold_bn(cout, bn, bias, "cpu")
    exp_scale, exp_shift = P.place_lanes(sc, out_map.tolist()), P.place_lanes(sf, out_map.tolist())
    d = lambda t: None if t is None else t.cuda()                      # noqa: E731
    wd, imd, omd, bd = d(w), d(in_map), d(out_map), d(bias)
    bnd = {kk: d(v) for kk, v in bn.items() if kk != "eps"} if bn else {}
    a = A.EdePrepackConv(w_oihw=wd.data_ptr(), in_map=imd.data_ptr() if use_maps else None,
                         out_map=omd.data_ptr() if use_maps else None,
                         bn_weight=bnd["weight"].data_ptr() if bn else None, bn_bias=bnd["bias"].data_ptr() if bn else None,
                         bn_mean=bnd["mean"].data_ptr() if bn else None, bn_var=bnd["var"].data_ptr() if bn else None,
                         bias=bd.data_ptr() if use_bias else None, bn_eps=1e-3, cout=cout, cin=cin, kh=k, kw=k,
                         cin_p=cin_p, cout_p=cout_p, passes=passes, pack_g=pack_g)
    b4, d2 = (C.c_int64 * 4)(), (C.c_int32 * 2)()
    A.check(lib.ede_prepack_conv_sizes(C.byref(a), b4, d2))
    assert list(d2) == [co16, ci64] and b4[0] == exp_umma.numel() * 2 and b4[1] == exp_direct.numel() * 4
    # (1) caller-provided buffers
    o_umma = torch.full((b4[0] // 2,), 7.0, dtype=torch.bfloat16, device="cuda")
    o_direct = torch.full((b4[1] // 4,), 7.0, device="cuda")
    o_scale, o_shift = torch.full((cout_p,), 7.0, device="cuda"), torch.full((cout_p,), 7.0, device="cuda")
    a.w_umma, a.w_direct, a.scale, a.shift = o_umma.data_ptr(), o_direct.data_ptr(), o_scale.data_ptr(), o_shift.data_ptr()
    A.check(lib.ede_prepack_conv_to(C.byref(a), stream()))
    torch.cuda.synchronize()
    assert torch.equal(o_umma.cpu().view(torch.int16), exp_umma.reshape(-1).view(torch.int16))
    assert torch.equal(o_direct.cpu(), exp_direct.reshape(-1))
    assert torch.allclose(o_scale.cpu(), exp_scale, rtol=2e-7, atol=0) and torch.allclose(o_shift.cpu(), exp_shift, rtol=4e-7, atol=1e-7)
    # (2) the allocating form a C-only caller would use
    a.w_umma = a.w_direct = a.scale = a.shift = None
    h = C.POINTER(A.EdePackedConv)()
    A.check(lib.ede_prepack_conv(C.byref(a), C.byref(h), stream()))
    torch.cuda.synchronize()
    hc = h.contents
    assert (hc.w_umma_cout_p, hc.w_umma_cin_p, hc.w_passes, hc.w_pack_g) == (co16, ci64, passes, max(1, pack_g))
    back = torch.empty(b4[0] // 2, dtype=torch.bfloat16, device="cuda")
    try:
        rt = C.CDLL("libcudart.so.12")
    except OSError:
        rt = C.CDLL("/usr/local/cuda/lib64/libcudart.so.12")
    assert rt.cudaMemcpy(C.c_void_p(back.data_ptr()), C.c_void_p(hc.w_umma), C.c_size_t(b4[0]), 3) == 0   # D2D
    assert torch.equal(back.cpu().view(torch.int16), exp_umma.reshape(-1).view(torch.int16))
    assert lib.ede_free(h) == 0


@pytest.mark.parametrize("shape", [(2, 14, 19, 57, 76, 32, 5), (1, 7, 9, 114, 152, 32, 5), (2, 8, 10, 33, 41, 16, 3),
                                   (1, 57, 76, 114, 152, 32, 5)])
def test_upconv_streaming_form_matches_the_separable_form(cuda, lib, shape):
    """ede_upconv_taps without a workspace (one-pass streaming kernel: x-lerped rows kept in registers while a thread
    walks down a band of output rows) == the two-pass form through the `rows` workspace; fp32 outputs, same
    arithmetic up to the order of the tap-row sum."""
    n, h, w, oh, ow, cgrp, k = shape
    g = torch.Generator().manual_seed(h * 131 + ow)
    z_cs = k * k * cgrp
    z = torch.randn(n, h, w, z_cs, generator=g).to(cuda)
    scale = (0.5 + torch.rand(cgrp, generator=g)).to(cuda)
    shift = (0.2 * torch.randn(cgrp, generator=g)).to(cuda)
    ys = [torch.full((n, oh, ow, cgrp), float("nan"), device=cuda) for _ in range(2)]
    rows = torch.empty(n * h * k * ow * cgrp, device=cuda)
    st = torch.cuda.current_stream().cuda_stream
    for y, r in ((ys[0], rows.data_ptr()), (ys[1], None)):
        a = A.EdeUpConv(z=z.data_ptr(), y=y.data_ptr(), scale=scale.data_ptr(), shift=shift.data_ptr(), rows=r,
                        n=n, h=h, w=w, z_cs=z_cs, z_co=0, oh=oh, ow=ow, cgrp=cgrp, y_cs=cgrp, y_co=0, y_dtype=A.F32,
                        kh=k, kw=k, pt=k // 2, pl=k // 2, act=A.ACT_RELU, act_c_end=cgrp // 2)
        assert lib.ede_upconv_taps(C.byref(a), st) == 0, lib.ede_last_error()
    torch.cuda.synchronize()
    assert not torch.isnan(ys[1]).any()
    assert torch.allclose(ys[0], ys[1], rtol=1e-5, atol=2e-5), float((ys[0] - ys[1]).abs().max())


@pytest.mark.parametrize("bands", [2, 3])
@pytest.mark.parametrize("dt", [A.BF16X2, A.F32_TO_BF16X2, A.BF16])
@pytest.mark.parametrize("k,s,pad", [(3, 1, (1, 1, 1, 1)), (5, 2, (1, 1, 2, 2)), (5, 1, (2, 2, 2, 2))])
def test_dwconv_row_bands(cuda, lib, dt, k, s, pad, bands, monkeypatch):
    """The row-streaming depthwise kernel cut into row bands per image (small batches of large maps: ENB4-LR 192x192 at
    batch 8 ran 24 blocks on 148 SMs): every band re-reads its K - S halo rows, the squeeze sums land in one slab per
    (column strip, band) and add up to the unsplit result."""
    g = torch.Generator().manual_seed(k + s + bands)
    n, c, h, w = 2, 40, 61, 47
    c_p = P.ceil8(c)
    pt, pl, pb, pr = pad
    oh, ow = (h + pt + pb - k) // s + 1, (w + pl + pr - k) // s + 1
    x = torch.randn(n, c, h, w, generator=g).to(cuda)
    wt = (torch.randn(c, 1, k, k, generator=g) / k).to(cuda)
    scale, shift = (0.5 + torch.rand(c, generator=g)).to(cuda), (0.2 * torch.randn(c, generator=g)).to(cuda)
    if dt == A.BF16X2:
        xin = split_store(nhwc(x, c_p, A.F32))
    elif dt == A.F32_TO_BF16X2:
        xin = nhwc(x, c_p, A.F32)
    else:
        xin = nhwc(x, c_p, dt)
    lanes = P.identity_map(c, c_p)
    wk = torch.zeros(k * k, c_p, device=cuda)
    wk[:, :c] = wt.reshape(c, k * k).t()
    sc, sf = P.place_lanes(scale, lanes), P.place_lanes(shift, lanes)
    outs = []
    for forced in (None, bands):
        if forced is None:
            monkeypatch.setenv("EDE_DW_BANDS", "1")
        else:
            monkeypatch.setenv("EDE_DW_BANDS", str(forced))
        y = torch.zeros(n, oh, ow, (2 if dt != A.BF16 else 1) * c_p, device=cuda, dtype=torch.bfloat16)
        a = A.EdeDwConv(x=xin.data_ptr(), y=y.data_ptr(), w=wk.data_ptr(), scale=sc.data_ptr(), shift=sf.data_ptr(),
                        pool_sum=None, n=n, h=h, w_=w, c_p=c_p, x_cs=c_p, x_co=0, oh=oh, ow=ow, y_cs=c_p, y_co=0, k=k, s=s,
                        pt=pt, pl=pl, act=A.ACT_SWISH, dtype=dt)
        parts = int(lib.ede_dwconv2d_pool_parts(C.byref(a)))
        pool = torch.zeros(parts, n, c_p, device=cuda)
        a.pool_sum, a.pool_parts = pool.data_ptr(), parts
        A.check(lib.ede_dwconv2d(C.byref(a), stream()), "ede_dwconv2d")
        torch.cuda.synchronize()
        outs.append((y.clone(), pool.sum(0), parts))
    (y1, p1, parts1), (yb, pb_, partsb) = outs
    assert partsb == parts1 * bands, (parts1, partsb)
    assert torch.equal(y1, yb)                                     # same arithmetic per output, another walk order
    assert torch.allclose(p1, pb_, rtol=1e-5, atol=1e-3)


def test_se_scale_fp32_in_split_out(cuda, lib):
    """ede_se_scale with EDE_F32_TO_BF16X2: the gate multiply reads the depthwise kernel's fp32 lanes and writes the split
    lanes the project conv consumes (late MBConv blocks, where the gate is not folded into the weights)."""
    g = torch.Generator().manual_seed(5)
    n, hw, c, c_p = 3, 63, 44, 48
    x = torch.zeros(n, hw, c_p)
    x[..., :c] = torch.randn(n, hw, c, generator=g)
    gate = torch.zeros(n, c_p)
    gate[:, :c] = torch.rand(n, c, generator=g)
    xd, gd = x.to(cuda), gate.to(cuda)
    y = torch.full((n, hw, 2 * c_p), float("nan"), device=cuda, dtype=torch.bfloat16)
    a = A.EdeSeScale(x=xd.data_ptr(), y=y.data_ptr(), gate=gd.data_ptr(), n=n, hw=hw, c_p=c_p, x_cs=c_p, x_co=0,
                     y_cs=c_p, y_co=0, dtype=A.F32_TO_BF16X2)
    A.check(lib.ede_se_scale(C.byref(a), stream()), "ede_se_scale")
    torch.cuda.synchronize()
    got = split_load(y.view(n, hw, 1, 2 * c_p), c_p).view(n, hw, c_p).cpu()
    ref = x * gate.view(n, 1, c_p)
    assert bool(((got - ref).abs() <= 2e-5 * ref.abs() + 1e-7).all()), float((got - ref).abs().max())


@pytest.mark.parametrize("k,hw", [(5, (33, 41)), (3, (8, 70)), (5, (114, 152))])
def test_tapsum_one_channel_conv(cuda, lib, k, hw):
    """ede_tapsum: y[p] = act(scale * sum_t z[p + offset(t)][t] + shift), taps outside the map contribute zero -- the second
    half of a k x k conv to ONE channel (R.conv2, modules.py:285,296) after its 1x1 GEMM left the tap products in z."""
    g = torch.Generator().manual_seed(k)
    n, (h, w) = 2, hw
    z_cs = P.ceil8(k * k)
    z = torch.zeros(n, h, w, z_cs)
    z[..., :k * k] = torch.randn(n, h, w, k * k, generator=g)
    zd = z.to(cuda)
    scale, shift = torch.tensor([1.7] + [0.0] * 7).to(cuda), torch.tensor([-0.3] + [0.0] * 7).to(cuda)
    y = torch.full((n, h, w, 8), float("nan"), device=cuda)
    a = A.EdeTapSum(z=zd.data_ptr(), y=y.data_ptr(), scale=scale.data_ptr(), shift=shift.data_ptr(), n=n, h=h, w=w,
                    z_cs=z_cs, z_co=0, y_cs=8, y_co=0, y_dtype=A.F32, kh=k, kw=k, pt=k // 2, pl=k // 2, act=A.ACT_RELU)
    A.check(lib.ede_tapsum(C.byref(a), stream()), "ede_tapsum")
    torch.cuda.synchronize()
    zp = F.pad(z[..., :k * k].permute(0, 3, 1, 2), (k // 2,) * 4)          # [n, taps, h + k - 1, w + k - 1]
    ref = sum(zp[:, ty * k + tx, ty:ty + h, tx:tx + w] for ty in range(k) for tx in range(k))
    ref = torch.relu(1.7 * ref - 0.3)
    got = y.cpu()
    assert torch.allclose(got[..., 0], ref, rtol=1e-5, atol=1e-5), float((got[..., 0] - ref).abs().max())
    assert float(got[..., 1:].abs().sum()) == 0.0
    a.kh = a.kw = 7                                      # 49 taps: more than z holds / a block stages per pixel
    assert lib.ede_tapsum(C.byref(a), stream()) == -1


@pytest.mark.parametrize("n,c,csq,parts,fold", [(3, 150, 6, 1, False), (8, 672, 28, 3, False), (5, 96, 4, 2, True),
                                               (2, 1152, 48, 1, True)])
def test_se_fc_cluster_matches_one_block_per_image(cuda, lib, n, c, csq, parts, fold, monkeypatch):
    """ede_se_fc across a thread-block cluster (hidden units exchanged through distributed shared memory, gate channels
    and the weight fold split over the CTAs) == the one-block-per-image kernel, bit for bit: gate, cleared squeeze slabs
    and the per-image gate-scaled project weights; the gate also against an fp64 reference."""
    g = torch.Generator().manual_seed(n * 7 + csq)
    c_p = P.ceil8(c) + 8
    pool0 = torch.zeros(parts, n, c_p)
    pool0[:, :, :c] = torch.rand(parts, n, c, generator=g) * 50
    w1 = torch.zeros(csq, c_p); w1[:, :c] = torch.randn(csq, c, generator=g) / c ** 0.5
    w2t = torch.zeros(csq, c_p); w2t[:, :c] = torch.randn(csq, c, generator=g) / csq ** 0.5
    b1 = 0.1 * torch.randn(csq, generator=g)
    b2 = torch.zeros(c_p); b2[:c] = 0.1 * torch.randn(c, generator=g)
    rows, cols, passes = 32, 2 * c_p, 2                      # a split-input project conv: columns = doubled lanes
    w_src = (torch.randn(passes, rows, cols, generator=g) * 0.1).to(torch.bfloat16)
    dev = lambda t: t.to(cuda)                              # noqa: E731
    w1d, w2d, b1d, b2d, wsd = dev(w1), dev(w2t), dev(b1), dev(b2), dev(w_src)
    outs = []
    for knob in ("0", "8", "2"):
        monkeypatch.setenv("EDE_SE_FC_CLUSTER", knob)
        pool = dev(pool0.clone())
        gate = torch.full((n, c_p), float("nan"), device=cuda)
        w_dst = torch.full((n, passes, rows, cols), float("nan"), device=cuda, dtype=torch.bfloat16)
        f = A.EdeSeFc(pool_sum=pool.data_ptr(), w1=w1d.data_ptr(), b1=b1d.data_ptr(), w2=w2d.data_ptr(), b2=b2d.data_ptr(),
                      gate=gate.data_ptr(), n=n, c=c, c_p=c_p, csq=csq, inv_hw=1.0 / 266, clear_pool=1, w2_transposed=1,
                      pool_parts=parts, w_src=wsd.data_ptr() if fold else None, w_dst=w_dst.data_ptr() if fold else None,
                      w_rows=rows, w_cols=cols, w_passes=passes, w_split=1)
        A.check(lib.ede_se_fc(C.byref(f), stream()), "ede_se_fc")
        torch.cuda.synchronize()
        assert float(pool.abs().sum()) == 0.0               # slabs cleared for the next forward
        outs.append((gate.cpu(), w_dst.cpu() if fold else None))
    for gate_k, wd_k in outs[1:]:
        assert torch.equal(outs[0][0], gate_k)
        if fold:
            assert torch.equal(outs[0][1].view(torch.int16), wd_k.view(torch.int16))
    mean = pool0.double().sum(0)[:, :c] / 266
    hid = mean @ w1[:, :c].double().t() + b1.double()
    hid = hid * torch.sigmoid(hid)
    gref = torch.sigmoid(hid @ w2t[:, :c].double() + b2[:c].double())
    assert torch.allclose(outs[0][0][:, :c].double(), gref, rtol=1e-4, atol=1e-5) and float(outs[0][0][:, c:].abs().sum()) == 0
    if fold:                                                # column ci of the weights scaled by the gate of its lane
        lane = (torch.arange(cols) // 16) * 8 + torch.arange(cols) % 8
        gcol = outs[0][0][:, lane]                          # [n, cols]
        ref = (w_src[0].float() + w_src[1].float()).unsqueeze(0) * gcol.unsqueeze(1)
        got = outs[0][1][:, 0].float() + outs[0][1][:, 1].float()
        assert torch.allclose(got, ref, rtol=2e-5, atol=1e-7)
