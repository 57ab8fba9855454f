"""Forward-plan compiler and runner (host logic of the B200 path).

A product module (``ReSIDE/models/*.py`` in this package) never executes torch ops in
``forward``.  Instead it *emits* its operators into a :class:`Graph` -- an ordered list of
``EdeOp`` records (include/ede.h) over NHWC activation buffers carved out of ONE arena --
which is compiled once per (module, input shapes, precision, device) and then replayed with
a single ``ede_plan_run`` call per forward.  There is NO fallback: without the CUDA library
or with a CPU tensor, ``forward`` raises.

Layout rules (include/ede.h): every activation view has a multiple-of-8 lane count; a
``View`` records which logical channel each lane carries (``-1`` = zero lane), so channel
concatenation (``torch.cat`` at ReSIDE/models/net.py:30, modules.py:265 of the reference) is
two producers writing adjacent slices of the same buffer, and the consumer's weights are
scattered to the lane order once at prepack time.
"""
import ctypes as C
import os
import threading
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import _cabi as A
from . import prepack as P

_PRECISION = os.environ.get("EDE_PRECISION", "bf16")
_ENCODER_PRECISION = os.environ.get("EDE_ENCODER_PRECISION", "bf16x2")
_ELT = {A.F32: 4, A.BF16: 2, A.BF16X2: 4}
_ALIGN = 1024
_SE_FOLD = os.environ.get("EDE_SE_FOLD", "1") != "0"   # SE gate folded into the project conv's per-image weights
_TAPSUM = os.environ.get("EDE_TAPSUM", "1") != "0"     # one-output-channel k x k convs as 1x1 GEMM + tap sum


def set_precision(mode: str, encoder: Optional[str] = None) -> None:
    """Precision of the forward plans compiled from now on.

    mode 'bf16': tcgen05 tensor-core path, fp32 accumulation, bf16 activations in the decoders.
        The ENCODER stage runs at `encoder` precision: 'bf16x2' (default; activations stored as split
        bf16 hi+lo pairs = 16 mantissa bits, weights as w_hi+w_lo, 4 MMAs per plain-bf16 MMA -- needed
        because randomly initialised BN+ReLU encoders amplify operand rounding ~100x)
        or 'bf16' (plain bf16 everywhere, fastest, ~3e-2 / 2e-1 relative L1 on ENB0-/RN50-HU).
    mode 'fp32': exact fp32 SIMT path for every operator (parity mode, 1e-3 tolerance)."""
    global _PRECISION, _ENCODER_PRECISION
    if mode not in ("bf16", "fp32"):
        raise ValueError(f"unknown precision mode '{mode}' (expected 'bf16' or 'fp32')")
    if encoder is not None:
        if encoder not in ("bf16x2", "bf16"):
            raise ValueError(f"unknown encoder precision '{encoder}' (expected 'bf16x2' or 'bf16')")
        _ENCODER_PRECISION = encoder
    _PRECISION = mode


def get_encoder_precision() -> str:
    return _ENCODER_PRECISION


def get_precision() -> str:
    return _PRECISION


class Buf:
    """One NHWC activation buffer ([n][h][w][cs] elements of `dtype`) inside the plan arena."""

    def __init__(self, n, h, w, cs, dtype):
        self.n, self.h, self.w, self.cs, self.dtype = n, h, w, cs, dtype
        self.first = self.last = -1
        self.offset = -1

    @property
    def nbytes(self):
        return self.n * self.h * self.w * self.cs * _ELT[self.dtype]


class View:
    """Channel slice [co, co+len(lanes)) of a buffer; lanes[j] = logical channel or -1 (zero)."""

    def __init__(self, buf: Buf, co: int, lanes: Sequence[int]):
        assert co % 8 == 0 and len(lanes) % 8 == 0
        self.buf, self.co, self.lanes = buf, co, list(lanes)

    n = property(lambda s: s.buf.n)
    h = property(lambda s: s.buf.h)
    w = property(lambda s: s.buf.w)
    c_p = property(lambda s: len(s.lanes))
    c = property(lambda s: sum(1 for v in s.lanes if v >= 0))
    dtype = property(lambda s: s.buf.dtype)

    @property
    def shape(self):          # logical NCHW shape, what the reference tensor would have
        return (self.n, self.c, self.h, self.w)

    def size(self, d):
        return self.shape[d]

    def slice(self, lo_p: int, n_p: int, relabel: bool = True) -> "View":
        lanes = self.lanes[lo_p:lo_p + n_p]
        if relabel:
            base = min([v for v in lanes if v >= 0], default=0)
            lanes = [v - base if v >= 0 else -1 for v in lanes]
        return View(self.buf, self.co + lo_p, lanes)


class Stage:
    """A region of the plan's weight blob.  `t` = CPU tensor uploaded into it (None: a region that a prepack kernel
    or the forward itself writes); `temp` regions live in the staging blob that is freed after prepacking."""
    __slots__ = ("t", "nbytes", "offset", "temp", "zero", "parts")

    def __init__(self, t, nbytes, temp=False, zero=False):
        self.t, self.nbytes, self.offset, self.temp, self.zero = t, int(nbytes), -1, temp, zero


def cat_views(views: Sequence[View]) -> View:
    """torch.cat(dim=1) of views that already sit side by side in one buffer."""
    b, co, lanes, base = views[0].buf, views[0].co, [], 0
    for v in views:
        if v.buf is not b or v.co != co + len(lanes):
            raise RuntimeError("cat_views: views are not adjacent slices of one buffer")
        lanes += [base + x if x >= 0 else -1 for x in v.lanes]
        base += v.c
    return View(b, co, lanes)


def _cpu(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """Prepacking arithmetic runs on HOST copies of the parameters (no eager GPU tensor ops at plan-compile time):
    the packed operands reach the device in one upload + one prepack kernel per conv (Plan.__init__)."""
    return None if t is None else t.detach().to("cpu")


def _bn_dict(bn):
    if bn is None:
        return None
    return dict(weight=_cpu(bn.weight), bias=_cpu(bn.bias), mean=_cpu(bn.running_mean), var=_cpu(bn.running_var),
                eps=bn.eps)


def _pad4(conv) -> Tuple[int, int, int, int]:
    """(top, left, bottom, right) zero padding of a conv module.  Conv2dStaticSamePadding of
    the product efficientnet_pytorch carries `static_pad` = (left, right, top, bottom)."""
    sp = getattr(conv, "static_pad", None)
    if sp is not None:
        l, r, t, b = sp
        return t, l, b, r
    ph, pw = conv.padding
    return ph, pw, ph, pw


class Graph:
    """Records operators; `finalize()` lays out the arena and returns a runnable Plan."""

    def __init__(self, device: torch.device, precision: Optional[str] = None):
        self.device = device
        self.precision = precision or _PRECISION
        self.encoder_precision = _ENCODER_PRECISION
        if self.precision == "fp32":
            self.decoder_dtype = self.encoder_dtype = A.F32
        else:
            self.decoder_dtype = A.BF16
            self.encoder_dtype = A.BF16X2 if self.encoder_precision == "bf16x2" else A.BF16
        self.act_dtype = self.decoder_dtype       # dtype of newly created activations (see encoder_scope)
        self.dry = device.type != "cuda"          # host-only compile (tests without a GPU)
        self.dry_no_lib = self.dry and not os.path.exists(A.LIB_PATH)
        self.ops: List[tuple] = []                # (kind, struct, [(field, Buf)], tag)
        self.bufs: List[Buf] = []
        self.stages: List[Stage] = []             # regions of the weight blob (host tensors, reserved outputs)
        self.prepacks: List[dict] = []            # one ede_prepack_conv_to call per dense conv (Plan.__init__)
        self.inputs: List[dict] = []              # {shape, targets: [(op index, pointer field)], lazy: to_nhwc op}
        self.outputs: List[tuple] = []            # (op index, shape)
        self.flops = 0.0                          # algorithmic 2*MAC of the reference graph's convs
        self.exec_flops = 0.0                     # 2*MAC actually issued to the tensor cores / CUDA cores
        self.conv_flops = 0.0                     # `flops` without the depthwise convs: the dense-conv roofline numerator
        self.conv_exec_flops = 0.0                # `exec_flops` of the dense convs (tensor-core MMAs + tap gathers + stem)
        self.count_ref_flops = True
        self.tags: List[str] = []

    # ------------------------------------------------------------------ helpers
    def encoder_scope(self):
        """Context manager: operators emitted inside run at the encoder precision."""
        g = self

        class _Scope:
            def __enter__(self_):
                self_.prev = g.act_dtype
                g.act_dtype = g.encoder_dtype

            def __exit__(self_, *exc):
                g.act_dtype = self_.prev
        return _Scope()

    def new_buf(self, n, h, w, cs, dtype=None) -> Buf:
        b = Buf(n, h, w, cs, self.act_dtype if dtype is None else dtype)
        self.bufs.append(b)
        return b

    def new_view(self, n, h, w, c, dtype=None) -> View:
        c_p = P.ceil8(c)
        return View(self.new_buf(n, h, w, c_p, dtype), 0, P.identity_map(c, c_p))

    def alloc_concat(self, n, h, w, channels: Sequence[int], dtype=None) -> List[View]:
        """Destination slices for producers whose outputs the reference concatenates."""
        cps = [P.ceil8(c) for c in channels]
        b = self.new_buf(n, h, w, sum(cps), dtype)
        out, co = [], 0
        for c, cp in zip(channels, cps):
            out.append(View(b, co, P.identity_map(c, cp)))
            co += cp
        return out

    def _dev(self, t: torch.Tensor, temp=False) -> Stage:
        """Host tensor -> a region of the weight blob (uploaded once in Plan.__init__)."""
        t = t.detach().to("cpu").contiguous()
        st = Stage(t, t.numel() * t.element_size(), temp=temp)
        self.stages.append(st)
        return st

    def _reserve(self, nbytes: int, zero=False) -> Stage:
        """Device-only region: written by a prepack kernel, or (zero=True) zero-initialised state of the forward."""
        st = Stage(None, nbytes, zero=zero)
        self.stages.append(st)
        return st

    def _add(self, kind, st, refs, tag, wrefs=()):
        st._wrefs = [(f, w) for f, w in wrefs if w is not None]      # weight-blob pointers, patched in Plan.__init__
        i = len(self.ops)
        for _, b in refs:
            if b.first < 0:
                b.first = i
            b.last = i
        self.ops.append((kind, st, refs, tag))
        return i

    # ------------------------------------------------------------------ operators
    def input_nchw(self, shape) -> View:
        """Network input: fp32 NCHW tensor supplied at run time -> NHWC view (EDE_OP_TO_NHWC)."""
        n, c, h, w = (int(s) for s in shape)
        v = self.new_view(n, h, w, c, self.encoder_dtype)     # inputs keep 16 mantissa bits (4 B/lane)
        st = A.EdeLayout(src=None, dst=None, n=n, c=c, h=h, w=w, c_p=v.c_p, cs=v.buf.cs, co=v.co, dtype=v.dtype)
        i = self._add(A.OP_TO_NHWC, st, [("dst", v.buf)], "to_nhwc")
        # the layout pass is dropped again in finalize() when only fused stem convs (EDE_OP_STEM) consumed the input
        self.inputs.append(dict(shape=(n, c, h, w), targets=[(i, "src")], lazy=i))
        v.input_index = len(self.inputs) - 1
        return v

    def output_nchw(self, v: View) -> None:
        """Module output: view -> fresh fp32 NCHW contiguous tensor (EDE_OP_TO_NCHW)."""
        lanes = [x for x in v.lanes if x >= 0]
        if lanes != list(range(len(lanes))) or v.lanes[:len(lanes)] != lanes:
            raise RuntimeError("output_nchw: view must carry its channels contiguously from lane 0")
        st = A.EdeLayout(src=None, dst=None, n=v.n, c=v.c, h=v.h, w=v.w, c_p=v.c_p, cs=v.buf.cs, co=v.co,
                         dtype=v.dtype)
        i = self._add(A.OP_TO_NCHW, st, [("src", v.buf)], "to_nchw")
        self.outputs.append((i, (v.n, v.c, v.h, v.w)))

    def conv(self, x: View, parts, stride=1, pad=(0, 0, 0, 0), act=A.ACT_NONE, res: Optional[View] = None,
             out: Optional[View] = None, relu_in=False, out_dtype=None, tag="conv", se_fold=None, hi_lo=False) -> View:
        """Dense conv + folded BN + activation (+ residual).  `parts` is a list of
        (weight OIHW, bn module|None, bias|None, activated: bool); the parts share the input and
        their outputs are laid side by side (each padded to 8 lanes) -- used to run the two 5x5
        convs of an _UpProjection (modules.py:35,37) as ONE GEMM with N = 2*Cout."""
        if relu_in and self.precision == "bf16":
            x = self.eltwise(x, act=A.ACT_RELU, tag=tag + ".relu_in")      # keeps the conv on the tcgen05 path
            relu_in = False
        kh, kw = int(parts[0][0].shape[2]), int(parts[0][0].shape[3])
        if (getattr(x, "input_index", None) is not None and len(parts) == 1 and res is None and not relu_in
                and out is None and kh == kw and not self.dry_no_lib
                and A.lib().ede_stem_supported(kh, stride, x.c, P.ceil8(int(parts[0][0].shape[0])))):
            return self._stem(x, parts[0], stride, pad, act, out_dtype, tag)
        pt, pl, pb, pr = pad
        oh = (x.h + pt + pb - kh) // stride + 1
        ow = (x.w + pl + pr - kw) // stride + 1
        cin = x.c
        if (_TAPSUM and self.precision == "bf16" and len(parts) == 1 and int(parts[0][0].shape[0]) == 1 and kh * kw >= 9
                and kh * kw <= 32 and stride == 1 and res is None and not relu_in and (oh, ow) == (x.h, x.w)
                and x.dtype != A.F32 and not tag.endswith(".z_gemm") and not self._npack_takes(x, 8, kh, kw, pad)):
            return self._conv_one_channel(x, parts[0], pad, act, out, out_dtype, tag)
        # a split-bf16 input is consumed as a plain bf16 tensor with twice the lanes (include/ede.h)
        split_in = x.dtype == A.BF16X2
        in_lanes = x.lanes
        if split_in:
            in_lanes = [v for g8 in range(0, x.c_p, 8) for v in x.lanes[g8:g8 + 8] * 2]
        mul = 2 if split_in else 1
        out_lanes, fulls, scales, shifts, act_end = [], [], [], [], 0
        seen_plain = False
        for (wt, bn, bias, activated) in parts:
            wt, bias = _cpu(wt), _cpu(bias)
            co = int(wt.shape[0])
            if int(wt.shape[1]) != cin:
                raise RuntimeError(f"{tag}: weight expects {int(wt.shape[1])} input channels, view has {cin}")
            omap = P.identity_map(co)
            fulls.append(P.scatter_weight(wt, in_lanes, omap))
            sc, sf = P.fold_bn(co, bn if isinstance(bn, dict) or bn is None else _bn_dict(bn), bias, "cpu")
            scales.append(P.place_lanes(sc, omap))
            shifts.append(P.place_lanes(sf, omap))
            base = sum(1 for v in out_lanes if v >= 0)
            out_lanes += [base + v if v >= 0 else -1 for v in omap]
            if activated and act != A.ACT_NONE:
                if seen_plain:
                    raise RuntimeError("conv: activated parts must precede plain ones")
                act_end = len(out_lanes)
            else:
                seen_plain = True
            if self.count_ref_flops:
                self.flops += 2.0 * x.n * oh * ow * co * cin * kh * kw
                self.conv_flops += 2.0 * x.n * oh * ow * co * cin * kh * kw
            self.exec_flops += 2.0 * x.n * oh * ow * co * cin * kh * kw
            self.conv_exec_flops += 2.0 * x.n * oh * ow * co * cin * kh * kw
        full = torch.cat(fulls, 0)                               # host fp32 [cout_p][cin lanes][kh][kw]
        cout_p = len(out_lanes)
        if out is None:
            out = View(self.new_buf(x.n, oh, ow, cout_p, out_dtype), 0, out_lanes)
        else:
            if (out.n, out.h, out.w, out.c_p) != (x.n, oh, ow, cout_p):
                raise RuntimeError(f"{tag}: destination slice {out.n, out.h, out.w, out.c_p} does not match "
                                   f"conv output {x.n, oh, ow, cout_p}")
        # hi+lo weights whenever the OUTPUT is kept at split precision (the accurate encoder path)
        # ... or when the caller keeps an fp32 output at that precision (hi_lo: the expand conv of an MBConv block)
        passes = 2 if ((out.dtype == A.BF16X2 or hi_lo) and self.precision != "fp32") else 1
        scale = self._dev(torch.cat(scales))
        shift = self._dev(torch.cat(shifts))
        if res is not None and res.dtype != out.dtype:
            res = self.resize(res, (res.h, res.w), out_dtype=out.dtype, force=True, tag=tag + ".res_cast")
        st = A.EdeConv(x=None, y=None, res=None, scale=1, shift=1, w_umma=1, w_direct=None,
                       n=x.n, h=x.h, w=x.w, cin_p=x.c_p * mul, x_cs=x.buf.cs * mul, x_co=x.co * mul,
                       oh=oh, ow=ow, cout_p=cout_p, y_cs=out.buf.cs, y_co=out.co,
                       res_cs=res.buf.cs if res is not None else 0, res_co=res.co if res is not None else 0,
                       kh=kh, kw=kw, sh=stride, sw=stride, pt=pt, pl=pl, act=act, act_c_end=act_end,
                       x_dtype=A.BF16 if split_in else x.dtype, y_dtype=out.dtype, x_layout=A.NHWC,
                       relu_in=1 if relu_in else 0, w_umma_cout_p=P.ceil_to(cout_p, 16),
                       w_umma_cin_p=P.ceil_to(x.c_p * mul, 64),
                       force_direct=1 if self.precision == "fp32" else 0, w_passes=passes, w_pack_g=-1,
                       w_per_image=kh * kw * passes if se_fold is not None else 0)
        # ask the tile picker whether the tcgen05 path takes this conv and how many kernel-row taps it wants on the
        # 128 M rows, then have ede_prepack_conv_to pack for exactly that (one kernel per conv, Plan.__init__)
        cfg = A.umma_config(st)
        pack_g = int(cfg.pack_g) if cfg.eligible and cfg.pack_g > 1 else 1
        job = dict(full=self._dev(full, temp=True), cout=cout_p, cin=int(full.shape[1]), kh=kh, kw=kw, passes=passes,
                   pack_g=pack_g, w_umma=None, w_direct=None, tag=tag)
        bytes4, dims2 = A.prepack_sizes(job)
        if cfg.eligible:
            job["w_umma"] = self._reserve(bytes4[0])
            st.w_umma_cout_p, st.w_umma_cin_p, st.w_pack_g = dims2[0], dims2[1], pack_g
        else:                                                    # exact SIMT path (fp32 mode, shapes tcgen05 cannot take)
            job["w_direct"] = self._reserve(bytes4[1])
            st.w_umma, st.w_pack_g = None, 1
        self.prepacks.append(job)
        w_used = job["w_umma"]
        if se_fold is not None:
            # the squeeze-excite gate of the input rides in the weights: ede_se_fc writes one gate-scaled copy of this
            # conv's packed weights per image (EdeSeFc.w_src / w_dst), and the conv indexes them by image
            if not cfg.eligible or kh * kw != 1 or pack_g != 1:
                raise RuntimeError(f"{tag}: the SE gate can only be folded into a 1x1 conv on the tcgen05 path")
            w_used = self._reserve(x.n * bytes4[0])
            fc = se_fold["st"]
            fc.w_rows, fc.w_cols, fc.w_passes, fc.w_split = dims2[0], dims2[1], passes, 1 if split_in else 0
            fc._wrefs += [("w_src", job["w_umma"]), ("w_dst", w_used)]
        refs = [("x", x.buf), ("y", out.buf)] + ([("res", res.buf)] if res is not None else [])
        self._add(A.OP_CONV, st, refs, tag, [("scale", scale), ("shift", shift), ("w_umma", w_used),
                                             ("w_direct", job["w_direct"])])
        return out

    def _npack_takes(self, x: View, cout_p: int, kh: int, kw: int, pad) -> bool:
        """Would the tile picker run this stride-1 conv in conv_npack_kernel (taps of a kernel row on the UMMA N axis)?
        For one output channel that beats the 1x1 GEMM + tap sum (R.conv2 of ENB0-HU batch 64: 0.101 ms against
        0.062 + 0.072): the tap products never leave TMEM."""
        if self.dry_no_lib or x.dtype == A.BF16X2:
            return False
        st = A.EdeConv(x=1, y=1, res=None, scale=1, shift=1, w_umma=1, w_direct=None, n=x.n, h=x.h, w=x.w, cin_p=x.c_p,
                       x_cs=x.buf.cs, x_co=x.co, oh=x.h, ow=x.w, cout_p=cout_p, y_cs=cout_p, y_co=0, res_cs=0, res_co=0,
                       kh=kh, kw=kw, sh=1, sw=1, pt=pad[0], pl=pad[1], act=A.ACT_NONE, act_c_end=0, x_dtype=x.dtype,
                       y_dtype=A.F32, x_layout=A.NHWC, relu_in=0, w_umma_cout_p=P.ceil_to(cout_p, 16),
                       w_umma_cin_p=P.ceil_to(x.c_p, 64), force_direct=0, w_passes=1, w_pack_g=-1, w_per_image=0)
        cfg = A.umma_config(st)
        return bool(cfg.eligible) and bool(cfg.pack_g & P.PACK_N)

    def _conv_one_channel(self, x: View, part, pad, act, out, out_dtype, tag) -> View:
        """k x k stride-1 conv to ONE channel (R.conv2, modules.py:296): a 128-row MMA would carry one useful row, so
        the Cin contraction of all kh*kw taps runs as ONE 1x1 GEMM (kh*kw output lanes, fp32) and EDE_OP_TAPSUM adds
        the taps at their pixel offsets (csrc/upconv.cu).  Exact in real arithmetic; z stays fp32."""
        wt, bn, bias, _ = part
        wt = _cpu(wt).float()
        cin, kh, kw = int(wt.shape[1]), int(wt.shape[2]), int(wt.shape[3])
        wz = wt[0].permute(1, 2, 0).reshape(kh * kw, cin, 1, 1)                     # lane ty*kw+tx
        prev = self.count_ref_flops
        self.count_ref_flops = False
        z = self.conv(x, [(wz, None, None, False)], 1, (0, 0, 0, 0), A.ACT_NONE, out_dtype=A.F32, tag=tag + ".z_gemm")
        self.count_ref_flops = prev
        fl = 2.0 * x.n * x.h * x.w * cin * kh * kw
        self.flops += fl
        self.conv_flops += fl
        if out is None:
            out = self.new_view(x.n, x.h, x.w, 1, out_dtype)
        sc, sf = P.fold_bn(1, bn if isinstance(bn, dict) or bn is None else _bn_dict(bn), _cpu(bias), "cpu")
        scale, shift = self._dev(P.place_lanes(sc, out.lanes)), self._dev(P.place_lanes(sf, out.lanes))
        st = A.EdeTapSum(z=None, y=None, scale=None, shift=None, n=x.n, h=x.h, w=x.w, z_cs=z.buf.cs, z_co=z.co,
                         y_cs=out.buf.cs, y_co=out.co, y_dtype=out.dtype, kh=kh, kw=kw, pt=pad[0], pl=pad[1], act=act)
        self._add(A.OP_TAPSUM, st, [("z", z.buf), ("y", out.buf)], tag + ".tapsum", [("scale", scale), ("shift", shift)])
        return out

    def _stem(self, x: View, part, stride, pad, act, out_dtype, tag) -> View:
        """First conv of the network straight from the fp32 NCHW input (EDE_OP_STEM, csrc/stem.cu)."""
        wt, bn, bias, _ = part
        wt, bias = _cpu(wt), _cpu(bias)
        co, cin, k = int(wt.shape[0]), int(wt.shape[1]), int(wt.shape[2])
        pt, pl, pb, pr = pad
        oh, ow = (x.h + pt + pb - k) // stride + 1, (x.w + pl + pr - k) // stride + 1
        out = self.new_view(x.n, oh, ow, co, out_dtype)
        w = torch.zeros(k * k * cin, out.c_p, dtype=torch.float32)
        w[:, :co] = wt.float().permute(2, 3, 1, 0).reshape(k * k * cin, co)
        sc, sf = P.fold_bn(co, bn if isinstance(bn, dict) or bn is None else _bn_dict(bn), bias, "cpu")
        w, sc, sf = self._dev(w), self._dev(P.place_lanes(sc, out.lanes)), self._dev(P.place_lanes(sf, out.lanes))
        st = A.EdeStem(x=None, y=None, w=None, scale=None, shift=None, n=x.n, cin=cin,
                       h=x.h, w_=x.w, oh=oh, ow=ow, cout_p=out.c_p, y_cs=out.buf.cs, y_co=out.co, k=k, s=stride,
                       pt=pt, pl=pl, act=act, y_dtype=out.dtype)
        i = self._add(A.OP_STEM, st, [("y", out.buf)], tag, [("w", w), ("scale", sc), ("shift", sf)])
        self.inputs[x.input_index]["targets"].append((i, "x"))
        fl = 2.0 * x.n * oh * ow * co * cin * k * k
        self.flops += fl
        self.exec_flops += fl
        self.conv_flops += fl
        self.conv_exec_flops += fl
        return out

    def conv_bn_act_add(self, pair: View, conv, bn, act, out: Optional[View] = None, tag="conv") -> View:
        """act(bn(conv(pair[:, :cin])) + pair[:, cin:]) -- the tail of an _UpProjection (modules.py:38-45 of the
        reference) -- as ONE GEMM without a residual read in the epilogue: the branch to add enters as `cout` extra
        input channels whose only non-zero weight is an exact 1.0 on the centre tap.  For that the BN scale moves
        from the epilogue into the conv weights (rounded to bf16 like the weights themselves).  The extra K is free
        while cin + cout still fits the channel chunk the conv would use anyway, so callers use it for narrow convs."""
        wt = _cpu(conv.weight)
        co, ci, kh, kw = (int(v) for v in wt.shape)
        if pair.c != ci + co or conv.stride[0] != 1 or kh % 2 == 0 or kw % 2 == 0:
            raise RuntimeError(f"{tag}: conv_bn_act_add needs [conv input | branch] channels and an odd stride-1 kernel")
        sc, sf = P.fold_bn(co, _bn_dict(bn), _cpu(conv.bias), "cpu")
        w = torch.zeros(co, ci + co, kh, kw, dtype=torch.float32)
        w[:, :ci] = wt.float() * sc.view(-1, 1, 1, 1)
        idx = torch.arange(co)
        w[idx, ci + idx, kh // 2, kw // 2] = 1.0
        before, before_c = self.flops, self.conv_flops
        y = self.conv(pair, [(w, None, sf, act != A.ACT_NONE)], 1, _pad4(conv), act, None, out, tag=tag)
        if self.count_ref_flops:                  # the reference does a cin-channel conv and an elementwise add
            self.flops = before + 2.0 * pair.n * y.h * y.w * co * ci * kh * kw
            self.conv_flops = before_c + 2.0 * pair.n * y.h * y.w * co * ci * kh * kw
        return y

    def conv_bn_act(self, x: View, conv, bn=None, act=A.ACT_NONE, res=None, out=None, relu_in=False,
                    out_dtype=None, tag="conv", se_fold=None, hi_lo=False) -> View:
        """nn.Conv2d (+ nn.BatchNorm2d) (+ activation) as one fused operator."""
        if conv.groups != 1:
            raise RuntimeError(f"{tag}: grouped convolution is not a dense conv")
        return self.conv(x, [(conv.weight, bn, conv.bias, act != A.ACT_NONE)], conv.stride[0], _pad4(conv), act,
                         res, out, relu_in, out_dtype, tag, se_fold=se_fold, hi_lo=hi_lo)

    def dwconv_bn_act(self, x: View, conv, bn, act, pool: bool, tag="dwconv", out_f32=False):
        """Depthwise conv + BN + activation; with pool=True also the per-(n,c) sums for SE.  out_f32: an fp32 input stays
        fp32 on the way out (the consumer is the gate multiply, which writes the split lanes the project conv reads)."""
        k, s = conv.kernel_size[0], conv.stride[0]
        pt, pl, pb, pr = _pad4(conv)
        oh, ow = (x.h + pt + pb - k) // s + 1, (x.w + pl + pr - k) // s + 1
        c = x.c
        if conv.groups != c or int(conv.weight.shape[0]) != c:
            raise RuntimeError(f"{tag}: not a depthwise convolution over {c} channels")
        # fp32 lanes in / split-bf16 lanes out (EDE_F32_TO_BF16X2) when the producer wrote fp32 inside a split-precision scope
        mixed = x.dtype == A.F32 and self.act_dtype == A.BF16X2 and not out_f32
        out = View(self.new_buf(x.n, oh, ow, x.c_p, A.BF16X2 if mixed else x.dtype), 0, x.lanes)
        wt = _cpu(conv.weight).float().reshape(c, k * k).t()               # [k*k][c]
        w = torch.zeros(k * k, x.c_p, dtype=torch.float32)
        idx = torch.tensor([j for j, v in enumerate(x.lanes) if v >= 0])
        src = torch.tensor([v for v in x.lanes if v >= 0])
        w[:, idx] = wt[:, src]
        sc, sf = P.fold_bn(c, _bn_dict(bn), _cpu(conv.bias), "cpu")
        w, sc, sf = self._dev(w), self._dev(P.place_lanes(sc, x.lanes)), self._dev(P.place_lanes(sf, x.lanes))
        st = A.EdeDwConv(x=None, y=None, w=None, scale=None, shift=None, pool_sum=None, n=x.n, h=x.h, w_=x.w,
                         c_p=x.c_p, x_cs=x.buf.cs, x_co=x.co,
                         oh=oh, ow=ow, y_cs=out.buf.cs, y_co=out.co, k=k, s=s, pt=pt, pl=pl, act=act,
                         dtype=A.F32_TO_BF16X2 if mixed else x.dtype)
        pool_t = None
        if pool:
            # one slab of squeeze sums per column strip of the depthwise kernel: plain stores, added in slab order by
            # the SE FC kernel -> a bit-reproducible forward.  Zeroed once at plan build (Plan.run re-zeroes the state
            # if a forward fails half way; EdeSeFc.clear_pool keeps the atomic fallback correct).
            st.pool_parts = 1 if self.dry_no_lib else int(A.lib().ede_dwconv2d_pool_parts(C.byref(st)))
            pool_t = self._reserve(st.pool_parts * x.n * x.c_p * 4, zero=True)
            pool_t.parts = st.pool_parts
        self._add(A.OP_DWCONV, st, [("x", x.buf), ("y", out.buf)], tag,
                  [("w", w), ("scale", sc), ("shift", sf), ("pool_sum", pool_t)])
        self.flops += 2.0 * x.n * oh * ow * c * k * k
        self.exec_flops += 2.0 * x.n * oh * ow * c * k * k
        return out, pool_t

    def expand_f32(self) -> bool:
        """Write the expanded tensor of an MBConv block as fp32 instead of split bf16? (EDE_EXPAND_F32=0 turns it off)"""
        return self.act_dtype == A.BF16X2 and self.precision == "bf16" and os.environ.get("EDE_EXPAND_F32", "1") != "0"

    def se_fold_pays(self, x: View, cout: int) -> bool:
        """Fold the gate into the consumer's weights (one scaled weight set per image) instead of a full read + write
        of the gated tensor?  Pays while the per-image weights are much smaller than the tensor: the early, high-
        resolution MBConv blocks.  tcgen05 path only."""
        return self.se_fold_pays_dims(x.h, x.w, x.c_p, x.dtype, cout)

    def se_fold_pays_dims(self, h: int, w: int, c_p: int, dtype: int, cout: int) -> bool:
        if not _SE_FOLD or self.precision != "bf16" or self.dry_no_lib:
            return False
        mul = 2 if dtype == A.BF16X2 else 1
        passes = 2 if self.act_dtype == A.BF16X2 else 1
        w_bytes = passes * P.ceil_to(P.ceil8(cout), 16) * P.ceil_to(c_p * mul, 64) * 2
        return 8 * w_bytes <= h * w * c_p * _ELT[dtype]

    def se_gate(self, x: View, pool_t, reduce_conv, expand_conv, tag="se", fold=False):
        """Squeeze-excite: gate = sigmoid(W2 swish(W1 mean + b1) + b2); returns x * gate -- or, with fold=True,
        (x, handle) for conv(..., se_fold=handle), which applies the gate through its weights."""
        c, csq = x.c, int(reduce_conv.weight.shape[0])
        idx = torch.tensor([j for j, v in enumerate(x.lanes) if v >= 0])
        src = torch.tensor([v for v in x.lanes if v >= 0])
        if idx.tolist() != list(range(c)):
            raise RuntimeError("se_gate: channels must occupy lanes 0..c-1")
        w1 = torch.zeros(csq, x.c_p, dtype=torch.float32)
        w1[:, idx] = _cpu(reduce_conv.weight).float().reshape(csq, c)[:, src]
        w2 = torch.zeros(csq, x.c_p, dtype=torch.float32)                          # [csq][c_p]: coalesced reads
        w2[:, idx] = _cpu(expand_conv.weight).float().reshape(c, csq)[src].t()
        b2 = torch.zeros(x.c_p, dtype=torch.float32)
        b2[idx] = _cpu(expand_conv.bias).float()[src]
        w1, w2, b2 = self._dev(w1), self._dev(w2), self._dev(b2)
        b1 = self._dev(_cpu(reduce_conv.bias).float())
        gate = self._reserve(x.n * x.c_p * 4, zero=True)
        st = A.EdeSeFc(pool_sum=None, w1=None, b1=None, w2=None, b2=None, gate=None, n=x.n, c=c, c_p=x.c_p, csq=csq,
                       inv_hw=1.0 / (x.h * x.w), clear_pool=1, w2_transposed=1, pool_parts=pool_t.parts)
        self._add(A.OP_SE_FC, st, [], tag + ".fc",
                  [("pool_sum", pool_t), ("w1", w1), ("b1", b1), ("w2", w2), ("b2", b2), ("gate", gate)])
        if fold:
            return x, dict(st=st, gate=gate)
        mixed = x.dtype == A.F32 and self.act_dtype == A.BF16X2          # fp32 from the depthwise kernel -> split lanes
        out = View(self.new_buf(x.n, x.h, x.w, x.c_p, A.BF16X2 if mixed else x.dtype), 0, x.lanes)
        st = A.EdeSeScale(x=None, y=None, gate=None, n=x.n, hw=x.h * x.w, c_p=x.c_p,
                          x_cs=x.buf.cs, x_co=x.co, y_cs=out.buf.cs, y_co=out.co,
                          dtype=A.F32_TO_BF16X2 if mixed else x.dtype)
        self._add(A.OP_SE_SCALE, st, [("x", x.buf), ("y", out.buf)], tag + ".scale", [("gate", gate)])
        return out

    def upconv(self, x: View, size, parts, pad, act, tag="upconv") -> View:
        """conv_kxk(bilinear_ac(x, size)) of several weight sets sharing the input, evaluated on the LOW-resolution
        grid (csrc/upconv.cu): a 1x1 tensor-core GEMM z = [W_tap]_tap x into fp32, then the tap gather.
        `parts` as in conv(); stride 1, symmetric `pad`."""
        oh, ow = int(size[0]), int(size[1])
        k = int(parts[0][0].shape[2])
        cin = x.c
        rows, scales, shifts, out_lanes, act_end = [], [], [], [], 0
        for (wt, bn, bias, activated) in parts:
            wt, bias = _cpu(wt), _cpu(bias)
            co = int(wt.shape[0])
            co_p = P.ceil8(co)
            w = torch.zeros(co_p, cin, k, k, dtype=torch.float32)
            w[:co] = wt.float()
            rows.append(w)
            sc, sf = P.fold_bn(co, _bn_dict(bn), bias, "cpu")
            omap = P.identity_map(co)
            scales.append(P.place_lanes(sc, omap))
            shifts.append(P.place_lanes(sf, omap))
            base = sum(1 for v in out_lanes if v >= 0)
            out_lanes += [base + v if v >= 0 else -1 for v in omap]
            if activated and act != A.ACT_NONE:
                act_end = len(out_lanes)
            self.flops += 2.0 * x.n * oh * ow * co * cin * k * k          # what the reference graph executes
            self.conv_flops += 2.0 * x.n * oh * ow * co * cin * k * k
        cgrp = len(out_lanes)
        wz = torch.cat(rows, 0)                                            # [cgrp][cin][k][k]
        wz = wz.permute(2, 3, 0, 1).reshape(k * k * cgrp, cin, 1, 1)       # lanes [(ty*k+tx)*cgrp + c]
        prev = self.count_ref_flops
        self.count_ref_flops = False
        # z is the largest tensor of the op (25 taps x cgrp lanes per low-resolution pixel): bf16 halves what the GEMM
        # writes and the gather reads; the sums of the gather stay fp32 (EDE_UPCONV_Z_BF16=0: fp32 z)
        z_dtype = A.BF16 if (self.precision == "bf16" and os.environ.get("EDE_UPCONV_Z_BF16", "1") != "0"
                             and os.environ.get("EDE_UPCONV_ROWS", "1") != "0" and x.h < oh) else A.F32
        z = self.conv(x, [(wz, None, None, False)], 1, (0, 0, 0, 0), A.ACT_NONE, out_dtype=z_dtype, tag=tag + ".z_gemm")
        self.count_ref_flops = prev
        out = View(self.new_buf(x.n, oh, ow, cgrp, self.decoder_dtype if self.act_dtype != A.F32 else A.F32), 0, out_lanes)
        scale, shift = self._dev(torch.cat(scales)), self._dev(torch.cat(shifts))
        # fp32 workspace of the separable two-pass gather: [n][h][k][ow][cgrp] (x-lerp of every tap on the low-res rows).
        # EDE_UPCONV_ROWS=0: the workspace-free one-pass streaming kernel instead (measured 20 % slower on B200: 240
        # registers per thread leave 8 warps per SM to hide the load latency; profiles/r02_upconv_stream.md)
        bufs = [("z", z.buf), ("y", out.buf)]
        if os.environ.get("EDE_UPCONV_ROWS", "1") != "0" and x.h < oh:
            bufs.append(("rows", self.new_buf(x.n, x.h, k * ow, cgrp, A.F32)))
        st = A.EdeUpConv(z=None, y=None, scale=None, shift=None, rows=None, n=x.n, h=x.h, w=x.w,
                         z_cs=z.buf.cs, z_co=z.co, oh=oh, ow=ow, cgrp=cgrp, y_cs=out.buf.cs, y_co=out.co,
                         y_dtype=out.dtype, kh=k, kw=k, pt=pad[0], pl=pad[1], act=act, act_c_end=act_end, z_dtype=z_dtype)
        self._add(A.OP_UPCONV, st, bufs, tag + ".taps",
                  [("scale", scale), ("shift", shift)])
        self.exec_flops += 2.0 * x.n * oh * ow * cgrp * k * k * 4
        self.conv_exec_flops += 2.0 * x.n * oh * ow * cgrp * k * k * 4
        return out

    def resize(self, x: View, size, out: Optional[View] = None, out_dtype=None, force=False, tag="resize") -> View:
        """F.interpolate(x, size, mode='bilinear', align_corners=True) (identity size = exact copy/cast)."""
        oh, ow = int(size[0]), int(size[1])
        if out is None:
            if (oh, ow) == (x.h, x.w) and not force and (out_dtype is None or out_dtype == x.dtype):
                return x                                   # identity resample
            out = View(self.new_buf(x.n, oh, ow, x.c_p, out_dtype), 0, x.lanes)
        st = A.EdeResize(x=None, y=None, n=x.n, h=x.h, w=x.w, c_p=x.c_p, x_cs=x.buf.cs, x_co=x.co, oh=oh, ow=ow,
                         y_cs=out.buf.cs, y_co=out.co, x_dtype=x.dtype, y_dtype=out.dtype)
        self._add(A.OP_RESIZE, st, [("x", x.buf), ("y", out.buf)], tag)
        return out

    def maxpool3x3s2(self, x: View, tag="maxpool") -> View:
        oh, ow = (x.h + 2 - 3) // 2 + 1, (x.w + 2 - 3) // 2 + 1
        out = View(self.new_buf(x.n, oh, ow, x.c_p, x.dtype), 0, x.lanes)
        st = A.EdeMaxPool(x=None, y=None, n=x.n, h=x.h, w=x.w, c_p=x.c_p, x_cs=x.buf.cs, x_co=x.co, oh=oh, ow=ow,
                          y_cs=out.buf.cs, y_co=out.co, dtype=x.dtype)
        self._add(A.OP_MAXPOOL, st, [("x", x.buf), ("y", out.buf)], tag)
        return out

    def eltwise(self, x: View, x2: Optional[View] = None, act=A.ACT_NONE, tag="eltwise") -> View:
        """y = act(x) (+ x2)."""
        if x2 is not None and (x2.lanes != x.lanes or x2.dtype != x.dtype):
            raise RuntimeError("eltwise: operands must share lane layout and dtype")
        out = View(self.new_buf(x.n, x.h, x.w, x.c_p, x.dtype), 0, x.lanes)
        st = A.EdeEltwise(x=None, x2=None, y=None, n_pix=x.n * x.h * x.w, c_p=x.c_p, x_cs=x.buf.cs, x_co=x.co,
                          x2_cs=x2.buf.cs if x2 is not None else 0, x2_co=x2.co if x2 is not None else 0,
                          y_cs=out.buf.cs, y_co=out.co, act=act, dtype=x.dtype)
        refs = [("x", x.buf), ("y", out.buf)] + ([("x2", x2.buf)] if x2 is not None else [])
        self._add(A.OP_ELTWISE, st, refs, tag)
        return out

    # ------------------------------------------------------------------ arena layout
    def finalize(self) -> "Plan":
        """Liveness-based first-fit placement of every buffer in one arena, then pointer patching."""
        dead = set()
        for inp in self.inputs:                    # NCHW -> NHWC passes whose output nobody reads
            i = inp["lazy"]
            dst = self.ops[i][2][0][1]
            if len(inp["targets"]) > 1 and not any(b is dst for j, op in enumerate(self.ops) if j != i for _, b in op[2]):
                dead.add(i)
                inp["targets"] = [t for t in inp["targets"] if t[0] != i]
        if dead:
            remap, keep = {}, []
            for j, op in enumerate(self.ops):
                if j not in dead:
                    remap[j] = len(keep)
                    keep.append(op)
            self.ops = keep
            for inp in self.inputs:
                inp["targets"] = [(remap[j], f) for j, f in inp["targets"]]
            self.outputs = [(remap[j], shp) for j, shp in self.outputs]
            for b in self.bufs:
                b.first = b.last = -1
            for j, (_, _, refs, _) in enumerate(self.ops):
                for _, b in refs:
                    if b.first < 0:
                        b.first = j
                    b.last = j
        live: List[Buf] = []
        top = 0
        for b in sorted((b for b in self.bufs if b.first >= 0), key=lambda b: (b.first, -b.nbytes)):
            live = [o for o in live if o.last >= b.first]
            live.sort(key=lambda o: o.offset)
            size = (b.nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
            pos = 0
            for o in live:
                o_size = (o.nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
                if pos + size <= o.offset:
                    break
                pos = max(pos, o.offset + o_size)
            b.offset = pos
            live.append(b)
            top = max(top, pos + size)
        return Plan(self, top)


class Plan:
    """A compiled forward: EdeOp array + arena + prepacked weights.  `run` is one C call (or one CUDA-graph launch).

    Device state: `arena` (activations, laid out by liveness) and `weights` (ONE blob: folded-BN vectors, depthwise /
    SE / stem weights, the conv operands written by ede_prepack_conv_to, and the zero-initialised squeeze-excite
    state).  Building it issues one host->device copy per blob and one prepack kernel per dense conv -- no eager
    tensor ops on the GPU."""

    def __init__(self, g: Graph, arena_bytes: int):
        self.device, self.precision, self.dry = g.device, g.precision, g.dry
        self.arena_bytes = arena_bytes
        self.flops = g.flops
        self.exec_flops = g.exec_flops
        self.conv_flops, self.conv_exec_flops = g.conv_flops, g.conv_exec_flops
        self.tags = [t for (_, _, _, t) in g.ops]
        self.kinds = [k for (k, _, _, _) in g.ops]
        self.inputs, self.outputs = g.inputs, g.outputs
        self.n_ops = len(g.ops)
        self.arena = self.weights = None
        self.graph_ok = True
        # ---- weight blob layout: uploaded regions first (one host image), then device-only regions
        live = [st for st in g.stages if not st.temp]
        temps = [st for st in g.stages if st.temp]
        top = 0
        for st in sorted(live, key=lambda st: st.t is None):
            st.offset = top
            top += (st.nbytes + 255) // 256 * 256
        self.weight_bytes = top
        upload_bytes = max([st.offset + st.nbytes for st in live if st.t is not None], default=0)
        t_top = 0
        for st in temps:
            st.offset = t_top
            t_top += (st.nbytes + 255) // 256 * 256
        self._zero_regions = [(st.offset, st.nbytes) for st in live if st.zero]
        base = wbase = 0
        if not self.dry:
            with torch.cuda.device(self.device):
                self.arena = torch.empty(arena_bytes + _ALIGN, dtype=torch.uint8, device=self.device)
                base = (self.arena.data_ptr() + _ALIGN - 1) // _ALIGN * _ALIGN
                self.weights = torch.empty(max(top, 256), dtype=torch.uint8, device=self.device)
                wbase = self.weights.data_ptr()
                self._upload(self.weights, [st for st in live if st.t is not None], upload_bytes)
                for off, nb in self._zero_regions:
                    A.check(A.lib().ede_fill_zero(wbase + off, nb, self._stream()), "ede_fill_zero")
                if g.prepacks:
                    stage = torch.empty(max(t_top, 256), dtype=torch.uint8, device=self.device)
                    self._upload(stage, temps, t_top)
                    for job in g.prepacks:
                        A.prepack_to(job, stage.data_ptr(), wbase, self._stream())
                    stage.record_stream(torch.cuda.current_stream(self.device))
                    del stage
        self.ops = (A.EdeOp * max(1, self.n_ops))()
        for i, (kind, st, refs, _) in enumerate(g.ops):
            for field, b in refs:
                setattr(st, field, base + b.offset if not self.dry else 1 + b.offset)
            for field, w in getattr(st, "_wrefs", ()):
                setattr(st, field, wbase + w.offset if not self.dry else 1 + w.offset)
            self.ops[i].kind = kind
            setattr(self.ops[i].u, A.OP_FIELD[kind], st)
        self.structs = [st for (_, st, _, _) in g.ops]      # host copies (introspection / tests)
        # (field, arena offset, bytes) of every activation buffer an op touches (debug tools: tools/replay_check.py)
        self.op_bufs = [[(f, b.offset, b.nbytes) for f, b in refs] for (_, _, refs, _) in g.ops]
        self.arena_base = base
        for st in g.stages:                                  # host tensors are not needed any more
            st.t = None
        self._lock = threading.Lock()
        self._last = None                                    # (stream id, event) of the previous run
        self._cg = None                                      # CUDA-graph replay state (enable_graph)

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    @staticmethod
    def _upload(blob: torch.Tensor, stages, nbytes: int) -> None:
        """One host image of the regions -> one host->device copy."""
        if nbytes <= 0 or not stages:
            return
        host = torch.zeros(nbytes, dtype=torch.uint8)
        for st in stages:
            host[st.offset:st.offset + st.nbytes] = st.t.reshape(-1).view(torch.uint8)
        blob[:nbytes].copy_(host)

    def conv_configs(self):
        """(tag, EdeConv, EdeUmmaCfg) of every dense conv -- which path each one takes."""
        out = []
        for k, st, tag in zip(self.kinds, self.structs, self.tags):
            if k == A.OP_CONV:
                out.append((tag, st, A.umma_config(st)))
        return out

    def launches_per_run(self) -> int:
        return self.n_ops

    def _rezero_state(self):
        """A forward that failed between a depthwise conv and its SE FC leaves dirty squeeze sums behind."""
        wbase = self.weights.data_ptr()
        for off, nb in self._zero_regions:
            A.lib().ede_fill_zero(wbase + off, nb, self._stream())

    def _guard_streams(self):
        """The arena and the SE state are per-plan: a run on another stream waits for the previous run."""
        cur = torch.cuda.current_stream(self.device)
        if self._last is not None and self._last[0] != cur.cuda_stream:
            cur.wait_event(self._last[1])
        return cur

    def _mark_done(self, cur):
        if self._last is None or self._last[0] != cur.cuda_stream:
            self._last = (cur.cuda_stream, torch.cuda.Event())
        self._last[1].record(cur)

    def _check_inputs(self, tensors):
        if self.dry:
            raise RuntimeError("efficientdepthestimation_b200: this plan was compiled without a CUDA device; "
                               "there is no CPU execution path")
        if len(tensors) != len(self.inputs):
            raise RuntimeError(f"plan expects {len(self.inputs)} input tensors, got {len(tensors)}")
        hold = []
        for inp, t in zip(self.inputs, tensors):
            if tuple(t.shape) != inp["shape"]:
                raise RuntimeError(f"plan compiled for input {inp['shape']}, got {tuple(t.shape)}")
            if t.dtype != torch.float32 or not t.is_contiguous():
                t = t.float().contiguous()
            hold.append(t)
        return hold

    def _enqueue(self, hold, outs):
        for inp, t in zip(self.inputs, hold):
            for i, field in inp["targets"]:
                setattr(getattr(self.ops[i].u, A.OP_FIELD[self.kinds[i]]), field, t.data_ptr())
        for (i, _), o in zip(self.outputs, outs):
            self.ops[i].u.layout.dst = o.data_ptr()
        rc = A.lib().ede_plan_run(self.ops, self.n_ops, self._stream())
        if rc != 0:
            msg = A.lib().ede_last_error().decode(errors="replace")
            self._rezero_state()
            raise RuntimeError(f"ede_plan_run failed (code {rc}): {msg}")

    def run(self, tensors: Sequence[torch.Tensor]) -> List[torch.Tensor]:
        hold = self._check_inputs(tensors)
        with self._lock, torch.cuda.device(self.device):
            cur = self._guard_streams()
            if self._cg is not None:
                outs = self._replay_graph(hold)
            else:
                outs = [torch.empty(shape, dtype=torch.float32, device=self.device) for _, shape in self.outputs]
                self._enqueue(hold, outs)
            self._mark_done(cur)
        return outs

    # ------------------------------------------------------------------ CUDA-graph replay (launch-bound small batches)
    def enable_graph(self) -> None:
        """Capture the plan's launches once into a CUDA graph over STATIC input / output buffers; `run` then costs one
        device-to-device input copy, one graph launch and one output copy.  Pays where the forward is launch-bound
        (batch 1: ~110 kernels of a few microseconds each)."""
        if self.dry or self._cg is not None:
            return
        with self._lock, torch.cuda.device(self.device):
            ins = [torch.zeros(inp["shape"], dtype=torch.float32, device=self.device) for inp in self.inputs]
            outs = [torch.empty(shape, dtype=torch.float32, device=self.device) for _, shape in self.outputs]
            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):
                self._enqueue(ins, outs)                       # warm-up outside capture (lazy attribute setting etc.)
                side.synchronize()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=side, capture_error_mode="thread_local"):
                    self._enqueue(ins, outs)
            torch.cuda.current_stream(self.device).wait_stream(side)
            self._cg = (graph, ins, outs)

    def _replay_graph(self, hold):
        graph, ins, outs = self._cg
        for dst, src in zip(ins, hold):
            dst.copy_(src, non_blocking=True)
        graph.replay()
        return [o.clone() for o in outs]


# ---------------------------------------------------------------------- module protocol
def emit(module, g: Graph, *args, **kw):
    """Emit `module` into `g`.  Product modules implement `_emit`; plain torch containers
    (nn.Sequential of conv/bn/relu/maxpool/blocks, as the reference builds them at
    modules.py:187-196 and lasinger2019.py:107-111) are pattern-matched here."""
    from torch import nn
    if hasattr(module, "_emit"):
        return module._emit(g, *args, **kw)
    if isinstance(module, nn.Sequential):
        (x,) = args
        mods = list(module.children())
        i = 0
        while i < len(mods):
            m = mods[i]
            if isinstance(m, nn.Conv2d):
                bn = mods[i + 1] if i + 1 < len(mods) and isinstance(mods[i + 1], nn.BatchNorm2d) else None
                j = i + (2 if bn is not None else 1)
                relu = j < len(mods) and isinstance(mods[j], nn.ReLU)
                nxt = mods[j + (1 if relu else 0)] if j + (1 if relu else 0) < len(mods) else None
                # EfficientNet stem -> first MBConv block (expand ratio 1: its first op is the depthwise kernel, a
                # CUDA-core consumer): fp32 lanes instead of split bf16, as between an expand conv and its depthwise conv
                f32 = (g.expand_f32() and nxt is not None and getattr(nxt, "_cfg", {}).get("expand") == 1
                       and hasattr(nxt, "_depthwise_conv"))
                x = g.conv_bn_act(x, m, bn, A.ACT_RELU if relu else A.ACT_NONE, tag=f"seq.{i}",
                                  out_dtype=A.F32 if f32 else None, hi_lo=f32)
                i = j + (1 if relu else 0)
                continue
            if isinstance(m, nn.MaxPool2d):
                k = m.kernel_size if isinstance(m.kernel_size, int) else m.kernel_size[0]
                s = m.stride if isinstance(m.stride, int) else m.stride[0]
                p = m.padding if isinstance(m.padding, int) else m.padding[0]
                if (k, s, p) != (3, 2, 1):
                    raise RuntimeError(f"unsupported MaxPool2d{(k, s, p)} on the B200 path")
                x = g.maxpool3x3s2(x)
            elif isinstance(m, nn.ReLU):
                x = g.eltwise(x, act=A.ACT_RELU)
            elif isinstance(m, nn.Identity):
                pass
            else:
                x = emit(m, g, x)
            i += 1
        return x
    if all(hasattr(module, a) for a in ("conv1", "bn1", "conv2", "bn2", "conv3", "bn3", "downsample")):
        return emit_bottleneck(module, g, *args)
    raise RuntimeError(f"efficientdepthestimation_b200: no B200 operator mapping for {type(module).__name__}")


def emit_bottleneck(m, g: Graph, x: View, tag="bottleneck") -> View:
    """ResNet bottleneck (ReSIDE/models/resnet.py:75-95; same attribute layout as torchvision's):
    1x1+BN+ReLU -> 3x3(stride)+BN+ReLU -> 1x1+BN (+ downsample 1x1(stride)+BN) -> add -> ReLU,
    the add and the final ReLU fused into the last conv's epilogue."""
    out = g.conv_bn_act(x, m.conv1, m.bn1, A.ACT_RELU, tag=tag + ".conv1")
    out = g.conv_bn_act(out, m.conv2, m.bn2, A.ACT_RELU, tag=tag + ".conv2")
    res = x
    if m.downsample is not None:
        ds = list(m.downsample.children())
        res = g.conv_bn_act(x, ds[0], ds[1], A.ACT_NONE, tag=tag + ".downsample")
    return g.conv_bn_act(out, m.conv3, m.bn3, A.ACT_RELU, res=res, tag=tag + ".conv3")


def _flatten(obj, tensors, spec):
    if isinstance(obj, torch.Tensor):
        tensors.append(obj)
        spec.append(("T", tuple(obj.shape)))
    elif isinstance(obj, (list, tuple)) and any(isinstance(o, torch.Tensor) for o in obj):
        spec.append(("L", len(obj)))
        for o in obj:
            _flatten(o, tensors, spec)
    else:
        spec.append(("V", tuple(int(v) for v in obj) if isinstance(obj, (list, tuple, torch.Size)) else obj))


def _rebuild(spec, it, views):
    kind, val = next(it)
    if kind == "T":
        return next(views)
    if kind == "L":
        return [_rebuild(spec, it, views) for _ in range(val)]
    return val


def _flatten_out(obj, out):
    if isinstance(obj, View):
        out.append(obj)
        return ("T",)
    return ("L", [_flatten_out(o, out) for o in obj], type(obj))


def _rebuild_out(s, it):
    if s[0] == "T":
        return next(it)
    items = [_rebuild_out(x, it) for x in s[1]]
    return tuple(items) if s[2] is tuple else items


def compile_module(module, args, device: torch.device, precision: Optional[str] = None):
    """Build the plan of `module(*args)`; tensors in `args` may be real or meta/CPU (dry)."""
    tensors, spec = [], []
    for a in args:
        _flatten(a, tensors, spec)
    g = Graph(device, precision)
    views = iter([g.input_nchw(t.shape) for t in tensors])
    it = iter(spec)
    call_args = [_rebuild(spec, it, views) for _ in args]
    with torch.no_grad():
        result = emit(module, g, *call_args)
        outs: List[View] = []
        out_spec = _flatten_out(result, outs)
        for v in outs:
            g.output_nchw(v)
    return g.finalize(), out_spec, tuple(spec)


_WEIGHT_EPOCH = [0]        # bumped by load_state_dict / .to() / .train() of ANY EdeModule: plans of parents go stale too


def _fingerprint(module):
    """(storage address, in-place version counter) of every parameter and buffer: catches in-place writes and
    re-assignments on any submodule (model.E.load_state_dict, p.data.copy_, optimiser steps) that no hook sees."""
    fp = []
    for t in module.parameters():
        fp.append((t.data_ptr(), t._version))
    for t in module.buffers():
        fp.append((t.data_ptr(), t._version))
    return tuple(fp)


class EdeModule:
    """Mixin giving an nn.Module the B200 `forward`: compile once per signature, then replay.

    Mirrors the reference's calling convention (fp32 NCHW in, fp32 NCHW out, eval mode).
    training=True, CPU tensors or a missing CUDA library raise -- there is no fallback.

    Compiled plans hold PREPACKED COPIES of the weights.  They are keyed by (input signature, precision, encoder
    precision, device) and validated on every call against (a) a process-wide epoch that any EdeModule's
    load_state_dict / .to() / .train() bumps and (b) a fingerprint of the module's parameters and buffers, so a
    weight change on any submodule recompiles instead of silently running old weights.  `freeze_plans()` skips (b)
    for latency-critical loops that own their weights; DataParallel replicas (fresh broadcast copies of the source
    module's weights on every call) are validated by epoch only."""

    def _ede_cache(self) -> dict:
        c = self.__dict__.get("_ede_plans")
        if c is None:
            c = {}
            self.__dict__["_ede_plans"] = c
        return c

    def invalidate_plans(self) -> None:
        """Drop compiled plans (weights changed: load_state_dict, .to(), .train())."""
        from torch import nn
        _WEIGHT_EPOCH[0] += 1
        # every EdeModule gets a (fresh) cache dict HERE, not lazily in its first forward: DataParallel replicas are
        # shallow copies made per call, so the dict must exist on the source module for the replicas to share it
        # (entries are keyed by device) -- otherwise each call would compile into a dict that dies with the replica
        for m in self.modules() if isinstance(self, nn.Module) else [self]:
            if isinstance(m, EdeModule):
                m.__dict__["_ede_plans"] = {}

    def freeze_plans(self, frozen: bool = True) -> None:
        """Promise that weights only change through load_state_dict / .to() / .train(): skip the per-call fingerprint."""
        self.__dict__["_ede_frozen"] = bool(frozen)

    def _apply(self, fn, *a, **kw):
        r = super()._apply(fn, *a, **kw)
        self.invalidate_plans()
        return r

    def load_state_dict(self, *a, **kw):
        r = super().load_state_dict(*a, **kw)
        self.invalidate_plans()
        return r

    def train(self, mode: bool = True):
        r = super().train(mode)
        if mode:
            self.invalidate_plans()
        return r

    def __deepcopy__(self, memo):           # plans hold raw device pointers: never copy them
        import copy
        cls = self.__class__
        new = cls.__new__(cls)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            new.__dict__[k] = {} if k == "_ede_plans" else copy.deepcopy(v, memo)
        return new

    def plan_for(self, *args, precision: Optional[str] = None):
        tensors, spec = [], []
        for a in args:
            _flatten(a, tensors, spec)
        if not tensors:
            raise RuntimeError(f"{type(self).__name__}.forward needs at least one tensor argument")
        dev = tensors[0].device
        key = (tuple(spec), precision or _PRECISION, _ENCODER_PRECISION, str(dev))
        cache = self._ede_cache()
        hit = cache.get(key)
        replica = bool(getattr(self, "_is_replica", False))
        check_fp = not replica and not self.__dict__.get("_ede_frozen", False)
        if hit is not None:
            stale = hit[3] != _WEIGHT_EPOCH[0] or (check_fp and hit[4] is not None and hit[4] != _fingerprint(self))
            if stale:
                hit = None
        if hit is None:
            plan, out_spec, in_spec = compile_module(self, args, dev, precision)
            # CUDA-graph replay for launch-bound plans -- not for DataParallel replicas: their forwards run in one host
            # thread per device at the same time, and a stream capture in one thread must not meet the plan build
            # (allocations, uploads) of another
            mode = os.environ.get("EDE_CUDA_GRAPH", "auto")
            if not plan.dry and not replica and (mode == "1" or (mode == "auto" and _graph_pays(plan))):
                plan.enable_graph()
            hit = (plan, out_spec, in_spec, _WEIGHT_EPOCH[0], None if replica else _fingerprint(self))
            cache[key] = hit
        return hit[0], hit[1], hit[2]

    def forward(self, *args):
        if self.training:
            raise RuntimeError(
                f"{type(self).__name__}: the B200 path implements the inference forward only "
                "(call .eval() first); there is no training-mode or CPU fallback")
        tensors, spec = [], []
        for a in args:
            _flatten(a, tensors, spec)
        for t in tensors:
            if t.device.type != "cuda":
                raise RuntimeError(
                    f"{type(self).__name__}.forward got a {t.device.type} tensor: efficientdepthestimation_b200 "
                    "runs on CUDA (sm_100a) only and has no CPU fallback")
        plan, out_spec, _ = self.plan_for(*args)
        outs = plan.run(tensors)
        return _rebuild_out(out_spec, iter(outs))


def _graph_pays(plan: "Plan") -> bool:
    """CUDA-graph replay for launch-bound plans: few images, so every kernel runs for only a few microseconds."""
    n = max((inp["shape"][0] for inp in plan.inputs), default=0)
    px = max((inp["shape"][0] * inp["shape"][2] * inp["shape"][3] for inp in plan.inputs), default=0)
    return n > 0 and px <= 4 * 384 * 384
