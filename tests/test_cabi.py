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
