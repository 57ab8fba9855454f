"""not gpu: the C-ABI library builds for sm_100a, loads, and exports every symbol include/ede.h declares."""
import ctypes
import os
import re
import subprocess

from efficientdepthestimation_b200 import _cabi as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "ede.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ede_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(lib):
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ede.h but not exported"
    assert sorted(A.EXPORTS) == names


def test_version_and_error_plumbing(lib):
    assert lib.ede_version() == 13
    assert lib.ede_fill_zero(None, -1, None) == -1            # EDE_E_ARG, no GPU touched
    assert b"ede_fill_zero" in lib.ede_last_error()
    assert lib.ede_plan_run(None, 0, None) == 0
    assert lib.ede_conv2d(None, None) == -1


def test_struct_sizes_match_header():
    """ctypes mirrors must have the layout the C compiler gives the header structs."""
    src = r'''
    #include <stdio.h>
    #include "ede.h"
    int main(){ printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(EdeConv), sizeof(EdeUmmaCfg),
      sizeof(EdeDwConv), sizeof(EdeSeFc), sizeof(EdeSeScale), sizeof(EdeResize), sizeof(EdeMaxPool),
      sizeof(EdeEltwise), sizeof(EdeLayout), sizeof(EdeOp), sizeof(EdeUpConv), sizeof(EdePrepackConv),
      sizeof(EdePackedConv), sizeof(EdeMeshFromDepth), sizeof(EdeRender)); return 0; }'''
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        sizes = [int(v) for v in subprocess.check_output([exe]).split()]
    mine = [ctypes.sizeof(t) for t in (A.EdeConv, A.EdeUmmaCfg, A.EdeDwConv, A.EdeSeFc, A.EdeSeScale, A.EdeResize,
                                       A.EdeMaxPool, A.EdeEltwise, A.EdeLayout, A.EdeOp, A.EdeUpConv, A.EdePrepackConv,
                                       A.EdePackedConv, A.EdeMeshFromDepth, A.EdeRender)]
    assert mine == sizes


def test_umma_config_is_pure_host(lib):
    a = A.EdeConv(x=8, y=8, scale=8, shift=8, w_umma=8, w_direct=8, n=4, h=114, w=152, cin_p=128, x_cs=128, x_co=0,
                  oh=114, ow=152, cout_p=128, y_cs=128, y_co=0, kh=5, kw=5, sh=1, sw=1, pt=2, pl=2, act=1,
                  act_c_end=128, x_dtype=A.BF16, y_dtype=A.BF16, x_layout=A.NHWC, w_umma_cout_p=128,
                  w_umma_cin_p=128, w_passes=1)
    cfg = A.umma_config(a)
    assert cfg.eligible == 1 and cfg.tmem_cols in (32, 64, 128, 256, 512) and cfg.smem_bytes <= 227 * 1024
    assert 0.5 < cfg.useful_frac <= 1.0
    a.x_dtype = A.F32
    assert A.umma_config(a).eligible == 0                      # fp32 activations take the exact SIMT path


def test_prepack_sizes_is_pure_host(lib):
    """ede_prepack_conv_sizes: the buffer sizes a C caller must allocate (no GPU touched)."""
    a = A.EdePrepackConv(cout=74, cin=74, kh=5, kw=5, cin_p=80, cout_p=80, passes=1, pack_g=1)
    b, d = (ctypes.c_int64 * 4)(), (ctypes.c_int32 * 2)()
    assert lib.ede_prepack_conv_sizes(ctypes.byref(a), b, d) == 0
    assert list(d) == [80, 128] and list(b) == [25 * 80 * 128 * 2, 25 * 80 * 80 * 4, 320, 320]
    a.pack_g, a.cout, a.cout_p, a.passes = 4, 16, 16, 2                  # 4 taps of a kernel row on the 128 M rows
    assert lib.ede_prepack_conv_sizes(ctypes.byref(a), b, d) == 0
    assert list(d) == [128, 128] and b[0] == 2 * 5 * 2 * 128 * 128 * 2
    a.cout_p = 40                                                        # does not fit 4 taps
    assert lib.ede_prepack_conv_sizes(ctypes.byref(a), b, d) == -1 and b"do not fit" in lib.ede_last_error()
    assert lib.ede_prepack_conv(None, None, None) == -1 and lib.ede_free(None) == 0


def _conv_desc(n, h, w, cin_p, cout_p, k, passes=1, x_dtype=None, y_dtype=None, pack=-1):
    x_dtype = A.BF16 if x_dtype is None else x_dtype
    y_dtype = A.BF16 if y_dtype is None else y_dtype
    return A.EdeConv(x=8, y=8, scale=8, shift=8, w_umma=8, w_direct=None, n=n, h=h, w=w, cin_p=cin_p, x_cs=cin_p, x_co=0,
                     oh=h, ow=w, cout_p=cout_p, y_cs=cout_p, y_co=0, kh=k, kw=k, sh=1, sw=1, pt=k // 2, pl=k // 2, act=1,
                     act_c_end=cout_p, x_dtype=x_dtype, y_dtype=y_dtype, x_layout=A.NHWC,
                     w_umma_cout_p=(cout_p + 15) // 16 * 16, w_umma_cin_p=(cin_p + 63) // 64 * 64, w_passes=passes,
                     w_pack_g=pack)


def test_tile_picker_offers_the_n_packed_form_to_narrow_convs(lib, monkeypatch):
    """ede_conv2d_umma_config (host only): narrow stride-1 k x k convs whose packed weights stay resident in shared
    memory get EDE_PACK_N | G (conv_npack_kernel: pixel pitch 32, G taps of a kernel row on the UMMA N axis); wide ones,
    1x1 convs and weight sets that do not fit keep the other forms.  The prepack sizes follow the answer."""
    PACK_N = 16
    for shape, g, n_pad in (((64, 114, 152, 32, 16, 3), 3, 48),          # conv1_2 of the narrow up-projections
                            ((64, 114, 152, 24, 32, 5), 5, 160),         # D.up4 / MFF.up1 on EfficientNet-B0 widths
                            ((64, 57, 76, 40, 48, 5), 5, 240),           # D.up3: 154 KB of weights, still resident
                            ((64, 114, 152, 80, 8, 5), 5, 80)):          # R.conv2: one real output channel
        cfg = A.umma_config(_conv_desc(*shape))
        assert cfg.eligible and cfg.pack_g == (PACK_N | g) and cfg.swap == 0 and cfg.a_mode == 0, (shape, cfg.pack_g)
        assert cfg.pitch == 32 and cfg.bw == 33 - shape[5] and cfg.bh == 4 * cfg.m_tiles and cfg.n_pad == n_pad
        assert cfg.a_stages >= 2 and cfg.b_stages >= 1 and cfg.smem_bytes <= 227 * 1024 and cfg.tmem_cols <= 512
        p = A.EdePrepackConv(cout=shape[4], cin=shape[3], kh=shape[5], kw=shape[5], cin_p=shape[3], cout_p=shape[4], passes=1,
                             pack_g=cfg.pack_g)
        b, d = (ctypes.c_int64 * 4)(), (ctypes.c_int32 * 2)()
        assert lib.ede_prepack_conv_sizes(ctypes.byref(p), b, d) == 0 and d[0] == n_pad
        # a conv that carries weights packed this way is only eligible in this form
        again = A.umma_config(_conv_desc(*shape, pack=cfg.pack_g))
        again_ok = _conv_desc(*shape, pack=cfg.pack_g)
        again_ok.w_umma_cout_p = n_pad
        assert again.eligible == 0 and A.umma_config(again_ok).pack_g == cfg.pack_g
    for shape in ((64, 114, 152, 80, 80, 5),                              # R.conv0: 384 KB of packed weights
                  (64, 114, 152, 64, 64, 5),                              # MFF.conv
                  (64, 57, 76, 48, 144, 1),                               # 1x1
                  (64, 14, 19, 160, 160, 5)):                             # Cout > 80
        cfg = A.umma_config(_conv_desc(*shape))
        assert cfg.eligible and not (cfg.pack_g & PACK_N), (shape, cfg.pack_g)
    monkeypatch.setenv("EDE_UMMA_NPACK", "0")
    assert not (A.umma_config(_conv_desc(64, 114, 152, 32, 16, 3)).pack_g & PACK_N)


def test_tile_picker_shares_the_pixel_box_between_weight_passes(lib, monkeypatch):
    """1x1 convs on a split-bf16 input with hi + lo weights: the weight stage holds both tiles (one pixel box per chunk)
    where three ring stages still fit, and only there."""
    deep = _conv_desc(64, 7, 9, 2304, 192, 1, passes=2, y_dtype=A.BF16X2)         # MBConv project 1152 -> 192
    cfg = A.umma_config(deep)
    assert cfg.eligible and cfg.a_mode == 1 and cfg.b_stage_bytes == 2 * cfg.n_pad * cfg.ck * 2 and cfg.b_stages >= 3
    wide = _conv_desc(64, 57, 76, 256, 256, 1, passes=2, y_dtype=A.BF16X2)        # ResNet conv3 128 -> 256: N = 256
    cfg = A.umma_config(wide)
    assert cfg.eligible and cfg.b_stage_bytes == cfg.n_pad * cfg.ck * 2
    monkeypatch.setenv("EDE_UMMA_SHARE_PASSES", "0")
    cfg = A.umma_config(deep)
    assert cfg.b_stage_bytes == cfg.n_pad * cfg.ck * 2


def test_dwconv_pool_parts_counts_row_bands(lib, monkeypatch):
    """ede_dwconv2d_pool_parts (host only): one squeeze slab per column strip and row band of the row-streaming kernel.
    Bands appear only where whole images cannot fill 148 SMs (ENB4-LR 192x192x48 at batch 8), not at batch 64."""
    def parts(n, h, w, c_p, k=3, s=1):
        a = A.EdeDwConv(x=8, y=8, w=8, scale=8, shift=8, pool_sum=None, n=n, h=h, w_=w, c_p=c_p, x_cs=c_p, x_co=0,
                        oh=h // s, ow=w // s, y_cs=c_p, y_co=0, k=k, s=s, pt=k // 2, pl=k // 2, act=A.ACT_SWISH,
                        dtype=A.F32_TO_BF16X2)
        return int(lib.ede_dwconv2d_pool_parts(ctypes.byref(a)))
    monkeypatch.setenv("EDE_DW_BANDS", "1")
    strips_small, strips_big = parts(8, 192, 192, 48), parts(64, 114, 152, 32)
    monkeypatch.delenv("EDE_DW_BANDS")
    assert parts(8, 192, 192, 48) > strips_small and parts(8, 192, 192, 48) % strips_small == 0
    assert parts(64, 114, 152, 32) == strips_big
    monkeypatch.setenv("EDE_DW_BANDS", "3")
    assert parts(64, 114, 152, 32) == 3 * strips_big
