"""ctypes binding of libede_sm100.so (the C ABI declared in include/ede.h).

No torch types cross this boundary: pointers are ``tensor.data_ptr()`` integers, the
stream is ``torch.cuda.current_stream().cuda_stream``.  There is NO fallback: if the
shared library is missing or a call fails, a RuntimeError is raised.
"""
import ctypes as C
import os
from ctypes import POINTER, Structure, Union, c_double, c_float, c_int32, c_int64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libede_sm100.so")

F32, BF16, BF16X2 = 0, 1, 2
F32_TO_BF16X2 = 3      # EdeDwConv.dtype / EdeSeScale.dtype only: fp32 lanes in, split-bf16 lanes out
ACT_NONE, ACT_RELU, ACT_SWISH, ACT_SIGMOID = 0, 1, 2, 3
NHWC, NCHW = 0, 1
(OP_CONV, OP_DWCONV, OP_SE_FC, OP_SE_SCALE, OP_RESIZE, OP_MAXPOOL, OP_ELTWISE, OP_TO_NHWC, OP_TO_NCHW,
 OP_ZERO, OP_UPCONV, OP_STEM, OP_TAPSUM) = range(1, 14)


def _ints(*names):
    return [(n, c_int32) for n in names]


class EdeConv(Structure):
    _fields_ = [("x", c_void_p), ("y", c_void_p), ("res", c_void_p), ("scale", c_void_p), ("shift", c_void_p),
                ("w_umma", c_void_p), ("w_direct", c_void_p)] + _ints(
        "n", "h", "w", "cin_p", "x_cs", "x_co", "oh", "ow", "cout_p", "y_cs", "y_co", "res_cs", "res_co",
        "kh", "kw", "sh", "sw", "pt", "pl", "act", "act_c_end", "x_dtype", "y_dtype", "x_layout", "relu_in",
        "w_umma_cout_p", "w_umma_cin_p", "force_direct", "w_passes", "w_pack_g", "w_per_image")


class EdeUmmaCfg(Structure):
    _fields_ = _ints("eligible", "a_mode", "ck", "swizzle", "bh", "bw", "pitch", "box_rows", "m_tiles", "n_pad",
                     "n_tiles", "a_stages", "b_stages", "a_stage_bytes", "b_stage_bytes", "smem_bytes",
                     "tmem_cols", "tiles_w", "tiles_h", "flat") + [("useful_frac", c_double)] + _ints("swap", "pack_g")


class EdeDwConv(Structure):
    _fields_ = [("x", c_void_p), ("y", c_void_p), ("w", c_void_p), ("scale", c_void_p), ("shift", c_void_p),
                ("pool_sum", c_void_p)] + _ints(
        "n", "h", "w_", "c_p", "x_cs", "x_co", "oh", "ow", "y_cs", "y_co", "k", "s", "pt", "pl", "act", "dtype",
        "pool_parts")


class EdeSeFc(Structure):
    _fields_ = [("pool_sum", c_void_p), ("w1", c_void_p), ("b1", c_void_p), ("w2", c_void_p), ("b2", c_void_p),
                ("gate", c_void_p)] + _ints("n", "c", "c_p", "csq") + [("inv_hw", c_float)] + _ints("clear_pool", "w2_transposed", "pool_parts") + [
        ("w_src", c_void_p), ("w_dst", c_void_p)] + _ints("w_rows", "w_cols", "w_passes", "w_split")


class EdeSeScale(Structure):
    _fields_ = [("x", c_void_p), ("y", c_void_p), ("gate", c_void_p)] + _ints(
        "n", "hw", "c_p", "x_cs", "x_co", "y_cs", "y_co", "dtype")


class EdeResize(Structure):
    _fields_ = [("x", c_void_p), ("y", c_void_p)] + _ints(
        "n", "h", "w", "c_p", "x_cs", "x_co", "oh", "ow", "y_cs", "y_co", "x_dtype", "y_dtype")


class EdeMaxPool(Structure):
    _fields_ = [("x", c_void_p), ("y", c_void_p)] + _ints(
        "n", "h", "w", "c_p", "x_cs", "x_co", "oh", "ow", "y_cs", "y_co", "dtype")


class EdeEltwise(Structure):
    _fields_ = [("x", c_void_p), ("x2", c_void_p), ("y", c_void_p)] + _ints(
        "n_pix", "c_p", "x_cs", "x_co", "x2_cs", "x2_co", "y_cs", "y_co", "act", "dtype")


class EdeLayout(Structure):
    _fields_ = [("src", c_void_p), ("dst", c_void_p)] + _ints("n", "c", "h", "w", "c_p", "cs", "co", "dtype")


class EdeUpConv(Structure):
    _fields_ = [("z", c_void_p), ("y", c_void_p), ("scale", c_void_p), ("shift", c_void_p), ("rows", c_void_p)] + _ints(
        "n", "h", "w", "z_cs", "z_co", "oh", "ow", "cgrp", "y_cs", "y_co", "y_dtype", "kh", "kw", "pt", "pl", "act",
        "act_c_end", "z_dtype")


class EdeStem(Structure):
    _fields_ = [("x", c_void_p), ("y", c_void_p), ("w", c_void_p), ("scale", c_void_p), ("shift", c_void_p)] + _ints(
        "n", "cin", "h", "w_", "oh", "ow", "cout_p", "y_cs", "y_co", "k", "s", "pt", "pl", "act", "y_dtype")


class EdeTapSum(Structure):
    _fields_ = [("z", c_void_p), ("y", c_void_p), ("scale", c_void_p), ("shift", c_void_p)] + _ints(
        "n", "h", "w", "z_cs", "z_co", "y_cs", "y_co", "y_dtype", "kh", "kw", "pt", "pl", "act")


class EdeDepthEpilogue(Structure):
    _fields_ = [("pred", c_void_p), ("label", c_void_p), ("resized", c_void_p), ("depth_mm", c_void_p),
                ("sums16", c_void_p), ("edges4", c_void_p)] + _ints("n", "h", "w", "oh", "ow") + [
        ("edge_thr", c_float), ("cap", c_float), ("mm_scale", c_float)]


class EdePreprocess(Structure):
    _fields_ = [("frames", c_void_p), ("y", c_void_p), ("xb", c_void_p), ("xk", c_void_p), ("yb", c_void_p),
                ("yk", c_void_p)] + _ints("n", "ih", "iw", "ch", "cw", "xks", "yks", "band", "span") + [
        ("mean", c_float * 3), ("std", c_float * 3)]


class EdePrepackConv(Structure):
    _fields_ = [("w_oihw", c_void_p), ("in_map", c_void_p), ("out_map", c_void_p), ("bn_weight", c_void_p),
                ("bn_bias", c_void_p), ("bn_mean", c_void_p), ("bn_var", c_void_p), ("bias", c_void_p),
                ("w_umma", c_void_p), ("w_direct", c_void_p), ("scale", c_void_p), ("shift", c_void_p),
                ("bn_eps", c_float)] + _ints("cout", "cin", "kh", "kw", "cin_p", "cout_p", "passes", "pack_g")


class EdePackedConv(Structure):
    _fields_ = [("w_umma", c_void_p), ("w_direct", c_void_p), ("scale", c_void_p), ("shift", c_void_p)] + _ints(
        "w_umma_cout_p", "w_umma_cin_p", "w_pack_g", "w_passes") + [("device_bytes", c_int64)]


class EdeMeshFromDepth(Structure):
    _fields_ = [("depth", c_void_p), ("vertices", c_void_p), ("uv", c_void_p), ("indices", c_void_p)] + _ints(
        "h", "w", "depth_stride", "density") + [("displacement", c_float)]


class EdeRender(Structure):
    _fields_ = [("vertices", c_void_p), ("uv", c_void_p), ("indices", c_void_p), ("mvp", c_void_p),
                ("texture", c_void_p), ("frames", c_void_p), ("workspace", c_void_p)] + _ints(
        "n_vertices", "n_triangles", "n_frames", "tex_h", "tex_w", "height", "width", "flip_y")


class _Zero(Structure):
    _fields_ = [("p", c_void_p), ("bytes", c_int64)]


class _OpU(Union):
    _fields_ = [("conv", EdeConv), ("dw", EdeDwConv), ("se_fc", EdeSeFc), ("se_scale", EdeSeScale),
                ("resize", EdeResize), ("maxpool", EdeMaxPool), ("eltwise", EdeEltwise), ("layout", EdeLayout),
                ("upconv", EdeUpConv), ("stem", EdeStem), ("tapsum", EdeTapSum), ("zero", _Zero)]


class EdeOp(Structure):
    _fields_ = [("kind", c_int32), ("_pad", c_int32), ("u", _OpU)]


OP_FIELD = {OP_CONV: "conv", OP_DWCONV: "dw", OP_SE_FC: "se_fc", OP_SE_SCALE: "se_scale", OP_RESIZE: "resize",
            OP_MAXPOOL: "maxpool", OP_ELTWISE: "eltwise", OP_TO_NHWC: "layout", OP_TO_NCHW: "layout",
            OP_ZERO: "zero", OP_UPCONV: "upconv", OP_STEM: "stem", OP_TAPSUM: "tapsum"}

# every symbol include/ede.h declares (checked by tests/test_cabi.py without a GPU)
EXPORTS = ["ede_conv2d", "ede_conv2d_umma_config", "ede_upconv_taps", "ede_stem_conv", "ede_stem_supported", "ede_dwconv2d", "ede_se_fc", "ede_se_scale",
           "ede_resize_bilinear_ac", "ede_resize_planes_f32", "ede_maxpool3x3s2", "ede_eltwise", "ede_nchw_to_nhwc", "ede_nhwc_to_nchw",
           "ede_fill_zero", "ede_sobel", "ede_sobel_loss_metrics", "ede_edge_counts", "ede_depth_epilogue", "ede_resample_coeffs", "ede_preprocess_frames", "ede_plan_run",
           "ede_launch_count", "ede_version", "ede_last_error", "ede_umma_available", "ede_debug_umma_prof",
           "ede_prepack_conv_sizes", "ede_prepack_conv_to", "ede_prepack_conv", "ede_free", "ede_dwconv2d_pool_parts",
           "ede_mesh_from_depth", "ede_render_workspace_bytes", "ede_render_mesh", "ede_tapsum",
           "ede_upconv_taps_static", "ede_upconv_taps_static_max_bytes"]

_lib = None


def lib():
    """The loaded library; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(efficientdepthestimation_b200 has no CPU or PyTorch fallback)")
    L = C.CDLL(LIB_PATH)
    for name, st in (("ede_conv2d", EdeConv), ("ede_dwconv2d", EdeDwConv), ("ede_se_fc", EdeSeFc),
                     ("ede_se_scale", EdeSeScale), ("ede_resize_bilinear_ac", EdeResize),
                     ("ede_maxpool3x3s2", EdeMaxPool), ("ede_eltwise", EdeEltwise),
                     ("ede_nchw_to_nhwc", EdeLayout), ("ede_nhwc_to_nchw", EdeLayout),
                     ("ede_upconv_taps", EdeUpConv), ("ede_upconv_taps_static", EdeUpConv), ("ede_stem_conv", EdeStem), ("ede_tapsum", EdeTapSum),
                     ("ede_depth_epilogue", EdeDepthEpilogue), ("ede_preprocess_frames", EdePreprocess)):
        f = getattr(L, name)
        f.argtypes = [POINTER(st), c_void_p]
        f.restype = c_int32
    L.ede_upconv_taps_static_max_bytes.argtypes = []
    L.ede_upconv_taps_static_max_bytes.restype = c_int64
    L.ede_mesh_from_depth.argtypes = [POINTER(EdeMeshFromDepth), c_void_p]
    L.ede_mesh_from_depth.restype = c_int32
    L.ede_render_workspace_bytes.argtypes = [POINTER(EdeRender)]
    L.ede_render_workspace_bytes.restype = c_int64
    L.ede_render_mesh.argtypes = [POINTER(EdeRender), c_void_p]
    L.ede_render_mesh.restype = c_int32
    L.ede_dwconv2d_pool_parts.argtypes = [POINTER(EdeDwConv)]
    L.ede_dwconv2d_pool_parts.restype = c_int32
    L.ede_stem_supported.argtypes = [c_int32, c_int32, c_int32, c_int32]
    L.ede_stem_supported.restype = c_int32
    L.ede_conv2d_umma_config.argtypes = [POINTER(EdeConv), POINTER(EdeUmmaCfg)]
    L.ede_conv2d_umma_config.restype = c_int32
    L.ede_fill_zero.argtypes = [c_void_p, c_int64, c_void_p]
    L.ede_fill_zero.restype = c_int32
    L.ede_resize_planes_f32.argtypes = [c_void_p, c_void_p, c_int64, c_int32, c_int32, c_int32, c_int32, c_void_p]
    L.ede_resize_planes_f32.restype = c_int32
    L.ede_sobel.argtypes = [c_void_p, c_void_p, c_int32, c_int32, c_int32, c_void_p]
    L.ede_sobel.restype = c_int32
    L.ede_sobel_loss_metrics.argtypes = [c_void_p, c_void_p, c_int32, c_int32, c_int32, c_void_p, c_void_p]
    L.ede_sobel_loss_metrics.restype = c_int32
    L.ede_edge_counts.argtypes = [c_void_p, c_void_p, c_int32, c_int32, c_int32, c_float, c_void_p, c_void_p]
    L.ede_edge_counts.restype = c_int32
    L.ede_resample_coeffs.argtypes = [c_int32, c_int32, c_int32, c_int32, c_void_p, c_void_p]
    L.ede_resample_coeffs.restype = c_int32
    L.ede_plan_run.argtypes = [POINTER(EdeOp), c_int32, c_void_p]
    L.ede_plan_run.restype = c_int32
    L.ede_launch_count.argtypes = [c_int32]
    L.ede_launch_count.restype = c_int64
    L.ede_version.restype = c_int32
    L.ede_last_error.restype = C.c_char_p
    L.ede_umma_available.restype = c_int32
    L.ede_prepack_conv_sizes.argtypes = [POINTER(EdePrepackConv), POINTER(c_int64), POINTER(c_int32)]
    L.ede_prepack_conv_sizes.restype = c_int32
    L.ede_prepack_conv_to.argtypes = [POINTER(EdePrepackConv), c_void_p]
    L.ede_prepack_conv_to.restype = c_int32
    L.ede_prepack_conv.argtypes = [POINTER(EdePrepackConv), POINTER(POINTER(EdePackedConv)), c_void_p]
    L.ede_prepack_conv.restype = c_int32
    L.ede_free.argtypes = [POINTER(EdePackedConv)]
    L.ede_free.restype = c_int32
    L.ede_debug_umma_prof.argtypes = [c_void_p, c_int64]
    L.ede_debug_umma_prof.restype = c_int64
    _lib = L
    return L


def check(rc: int, what: str = "libede"):
    if rc != 0:
        raise RuntimeError(f"{what} failed (code {rc}): {lib().ede_last_error().decode(errors='replace')}")


def umma_config(conv: EdeConv) -> EdeUmmaCfg:
    cfg = EdeUmmaCfg()
    check(lib().ede_conv2d_umma_config(C.byref(conv), C.byref(cfg)), "ede_conv2d_umma_config")
    return cfg


def _prepack_struct(job) -> EdePrepackConv:
    """engine job (already lane-scattered host weight [cout_p][cin_p][kh][kw]) -> EdePrepackConv with identity maps."""
    return EdePrepackConv(cout=job["cout"], cin=job["cin"], kh=job["kh"], kw=job["kw"], cin_p=job["cin"],
                          cout_p=job["cout"], passes=job["passes"], pack_g=job["pack_g"], bn_eps=0.0)


def prepack_sizes(job):
    """(bytes of {w_umma, w_direct, scale, shift}, (w_umma_cout_p, w_umma_cin_p)) -- host function, no GPU."""
    b, d = (c_int64 * 4)(), (c_int32 * 2)()
    check(lib().ede_prepack_conv_sizes(C.byref(_prepack_struct(job)), b, d), "ede_prepack_conv_sizes")
    return list(b), list(d)


def prepack_to(job, stage_base: int, weight_base: int, stream) -> None:
    """One ede_prepack_conv_to launch: staged fp32 weight -> w_umma / w_direct regions of the weight blob."""
    a = _prepack_struct(job)
    a.w_oihw = stage_base + job["full"].offset
    if job["w_umma"] is not None:
        a.w_umma = weight_base + job["w_umma"].offset
    if job["w_direct"] is not None:
        a.w_direct = weight_base + job["w_direct"].offset
    check(lib().ede_prepack_conv_to(C.byref(a), stream), f"ede_prepack_conv_to({job.get('tag', '')})")
