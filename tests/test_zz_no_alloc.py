"""ede_upconv_taps_static -- the allocation-free variant of the separable tap gather (its `rows` workspace is a
file-scope device array inside libede_sm100.so) -- against ede_upconv_taps with a caller-provided workspace, through
the C ABI, on two calls in a row.  Collected last on purpose."""
import ctypes as C

import pytest
import torch

from efficientdepthestimation_b200 import _cabi as A

pytestmark = pytest.mark.gpu


def _case(cuda, seed, n, h, w, oh, ow, cgrp, k, y_dtype):
    g = torch.Generator().manual_seed(seed)
    z_cs = k * k * cgrp
    z = torch.randn(n, h, w, z_cs, generator=g).to(cuda)
    scale = (0.5 + torch.rand(cgrp, generator=g)).to(cuda)
    shift = (0.2 * torch.randn(cgrp, generator=g)).to(cuda)
    dt = torch.float32 if y_dtype == A.F32 else torch.bfloat16
    ys = [torch.full((n, oh, ow, cgrp), float("nan"), dtype=dt, device=cuda) for _ in range(2)]
    rows = torch.empty(n * h * k * ow * cgrp, dtype=torch.float32, device=cuda)

    def args(y, rows_ptr):
        return A.EdeUpConv(z=z.data_ptr(), y=y.data_ptr(), scale=scale.data_ptr(), shift=shift.data_ptr(), rows=rows_ptr,
                           n=n, h=h, w=w, z_cs=z_cs, z_co=0, oh=oh, ow=ow, cgrp=cgrp, y_cs=cgrp, y_co=0,
                           y_dtype=y_dtype, kh=k, kw=k, pt=k // 2, pl=k // 2, act=A.ACT_RELU, act_c_end=cgrp // 2 & ~7)
    return ys, rows, args, (z, scale, shift)


def test_upconv_taps_static_matches_workspace_form(cuda, lib):
    st = torch.cuda.current_stream().cuda_stream
    limit = int(lib.ede_upconv_taps_static_max_bytes())
    assert 0 < limit <= 64 << 20
    # two calls in a row on the shared array: a 5x5 gather, then a smaller 3x3 one that finds the first call's contents
    for seed, shape in enumerate([(2, 14, 19, 57, 76, 32, 5, A.BF16), (1, 8, 10, 33, 41, 16, 3, A.F32)]):
        n, h, w, oh, ow, cgrp, k, y_dtype = shape
        assert n * h * k * ow * cgrp * 4 <= limit
        (y_ref, y_new), rows, args, keep = _case(cuda, seed, *shape)
        a = args(y_ref, rows.data_ptr())
        assert lib.ede_upconv_taps(C.byref(a), st) == 0, lib.ede_last_error()
        b = args(y_new, None)
        assert lib.ede_upconv_taps_static(C.byref(b), st) == 0, lib.ede_last_error()
        torch.cuda.synchronize()
        assert not torch.isnan(y_ref.float()).any() and y_ref.float().abs().sum() > 0
        # same kernels, same arithmetic, another workspace address: bit-exact
        assert torch.equal(y_ref, y_new), float((y_ref.float() - y_new.float()).abs().max())
        del keep


def test_upconv_taps_static_rejects_oversize(cuda, lib):
    st = torch.cuda.current_stream().cuda_stream
    limit = int(lib.ede_upconv_taps_static_max_bytes())
    # a rows workspace one column beyond the limit: rejected on the host before anything is launched or dereferenced
    n, h, k, cgrp = 64, 57, 5, 32
    ow = limit // (n * h * k * cgrp * 4) + 1
    dummy = torch.zeros(8, device=cuda)
    a = A.EdeUpConv(z=dummy.data_ptr(), y=dummy.data_ptr(), scale=dummy.data_ptr(), shift=dummy.data_ptr(), rows=None,
                    n=n, h=h, w=76, z_cs=k * k * cgrp, z_co=0, oh=114, ow=ow, cgrp=cgrp, y_cs=cgrp, y_co=0,
                    y_dtype=A.F32, kh=k, kw=k, pt=2, pl=2, act=A.ACT_NONE, act_c_end=0)
    assert lib.ede_upconv_taps_static(C.byref(a), st) == -1
    msg = lib.ede_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else msg
    assert "static scratch" in msg and str(limit) in msg, msg
