/* ede.h -- C ABI of libede_sm100.so: the B200 (sm_100a) kernels behind the ReSIDE
 * depth-estimation forward path.
 *
 * The reference (AnthonyDickson/EfficientDepthEstimation) has no FFI of its own: its
 * hot path is a tree of torch.nn modules whose arithmetic is executed by ATen/cuDNN
 * (SURVEY.md 2.1).  Each entry point below therefore replaces ONE library operator as
 * the reference calls it; the citation after "replaces:" is the reference call site
 * (paths relative to the reference repository root).
 *
 * Conventions: plain C, raw DEVICE pointers, explicit shapes, caller-provided
 * cudaStream_t passed as void*; every call is asynchronous on that stream, allocates
 * nothing, never synchronises, never throws.  Return 0 on success, negative EDE_E_* on
 * error (text via ede_last_error(), thread-local).  Activations are NHWC ("pixel-major"):
 * element (n,h,w,c) of a view lives at base[((n*H+h)*W+w)*cs + co + c]; cs (channel
 * stride) and co (channel offset) are multiples of 8 elements so that every pixel slice
 * is 16-byte aligned, lanes between the real channel count and the next multiple of 8
 * hold zeros.  dtype: EDE_F32, EDE_BF16 or EDE_BF16X2.
 * EDE_BF16X2 ("split bf16") keeps 16 mantissa bits per value: every group of 8 lanes is stored as
 * [8 x bf16 hi][8 x bf16 lo] (4 bytes per lane, value = hi + lo).  The same bytes seen as a plain
 * EDE_BF16 view with twice the lanes (cs, co, c_p all doubled) are a valid tensor-core operand: with
 * weights duplicated over the hi/lo lanes one K=16 MMA computes (hi+lo).w.  ede_conv2d therefore takes a
 * split INPUT as x_dtype = EDE_BF16 with doubled lane numbers, and w_passes = 2 adds the low half of the
 * weights (w_umma holds [w_hi ; w_lo] as 2*kh*kw taps) -- the accurate path of the encoders.
 */
#ifndef EDE_H_
#define EDE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EDE_ABI_VERSION 13

enum { EDE_F32 = 0, EDE_BF16 = 1, EDE_BF16X2 = 2,
       EDE_F32_TO_BF16X2 = 3 };   /* EdeDwConv.dtype / EdeSeScale.dtype only: x lanes are fp32, y lanes split bf16 (same
                                     4 bytes per lane) */
enum { EDE_ACT_NONE = 0, EDE_ACT_RELU = 1, EDE_ACT_SWISH = 2, EDE_ACT_SIGMOID = 3 };
enum { EDE_OK = 0, EDE_E_ARG = -1, EDE_E_CUDA = -2, EDE_E_UNSUPPORTED = -3, EDE_E_DRIVER = -4 };
enum { EDE_NHWC = 0, EDE_NCHW = 1 };
enum { EDE_PACK_N = 16 };   /* flag in w_pack_g / pack_g: kernel-row taps packed on the UMMA N axis (see EdeConv.w_pack_g) */

/* One dense convolution + folded BatchNorm + activation (+ residual).
 * replaces: nn.Conv2d -> nn.BatchNorm2d(eval) -> ReLU/swish -> `out += residual`
 *   ReSIDE/models/resnet.py:78-93 (Bottleneck), ReSIDE/models/modules.py:35-39 (_UpProjection),
 *   modules.py:229,265-266,288-296 (D.conv, MFF.conv, R.*), ReSIDE/models/lasinger2019.py:153-161,
 *   288-293, and MBConv expand/project 1x1 of efficientnet-pytorch 0.6.3.
 *   y[n,oh,ow,k] = act_k( scale[k] * sum_{r,s,c} x[n, oh*sh-pt+r, ow*sw-pl+s, c] * w[k,c,r,s]
 *                         + shift[k] + res[n,oh,ow,k] ),  act applied to channels k < act_c_end. */
typedef struct EdeConv {
    const void* x;          /* input view base (dtype x_dtype)                                  */
    void* y;                /* output view base (dtype y_dtype)                                 */
    const void* res;        /* optional residual view (dtype y_dtype), NULL if none             */
    const float* scale;     /* [cout_p] fp32, folded BN gamma/sqrt(var+eps) (1 if no BN)        */
    const float* shift;     /* [cout_p] fp32, folded BN beta-mean*scale (+ conv bias)           */
    const void* w_umma;     /* bf16 [kh*kw][cout_p16][cin_p64] K-major packed weights, or NULL   */
    const float* w_direct;  /* fp32 [kh*kw][cin_p][cout_p] packed weights (always present)       */
    int32_t n, h, w, cin_p, x_cs, x_co;          /* cin_p: visible input lanes (multiple of 8)   */
    int32_t oh, ow, cout_p, y_cs, y_co;          /* cout_p: written output lanes (multiple of 8) */
    int32_t res_cs, res_co;
    int32_t kh, kw, sh, sw, pt, pl;              /* bottom/right padding is implied by oh/ow     */
    int32_t act, act_c_end;
    int32_t x_dtype, y_dtype;
    int32_t x_layout;       /* EDE_NCHW only for a 3-channel fp32 network input (stem convs)     */
    int32_t relu_in;        /* apply ReLU to x while loading (lasinger2019.py:153), direct path  */
    int32_t w_umma_cout_p, w_umma_cin_p;         /* padded dims of w_umma                        */
    int32_t force_direct;   /* 1: never take the tcgen05 path (fp32 parity mode / debugging)     */
    int32_t w_passes;       /* 1, or 2: w_umma holds 2*kh*kw taps, [w_hi taps ; w_lo taps]          */
    int32_t w_pack_g;       /* how w_umma packs kernel-row taps on its Cout axis: 0/1 = one tap per tile
                               (layout above); G = 2/4: tiles [kh*ceil(kw/G)][128][cin_p64] with row
                               j*(128/G)+c = w[c][.][ky][gx*G+j] (zero rows for taps >= kw), needs
                               cout_p <= 128/G; EDE_PACK_N | G (G = 2..5): the G taps sit on the UMMA N axis of the
                               pixels-on-M form instead -- tiles [kh*ceil(kw/G)][G*cout_p16][cin_p64] with row
                               j*cout_p16+c (cout_p16 = cout_p rounded up to 16, G*cout_p16 <= 256); -1 only for
                               ede_conv2d_umma_config: 'which packing should I prepack for?' (answer in
                               EdeUmmaCfg.pack_g)                                                       */
    int32_t w_per_image;    /* T > 0: w_umma holds n weight sets of T taps each (image i uses taps [i*T, (i+1)*T)): the
                               squeeze-excite gate folded into the project conv's weights by ede_se_fc (w_dst); tiles
                               then never span two images */
} EdeConv;
int ede_conv2d(const EdeConv* a, void* stream);

/* Weight prepacking: reference layout -> the operand layouts of ede_conv2d, one pass on the device.
 * replaces: nothing the reference executes per forward -- cuDNN consumes OIHW weights and the reference evaluates
 *   BatchNorm as a separate module (ReSIDE/models/modules.py:21-29, resnet.py:64-71); this is the load-time step
 *   between `model.load_state_dict` (ReSIDE/test.py:22, Benchmark/reside_model.py:44) and the first forward.
 * w_oihw: DEVICE fp32 [cout][cin][kh][kw] (nn.Conv2d.weight).  in_map / out_map: DEVICE int32 [cin_p] / [cout_p],
 *   lane -> source channel, -1 = zero lane; NULL = identity (lanes beyond cin / cout are zero).  bn_*: DEVICE fp32
 *   [cout] running statistics + affine (all four or none), bias: conv bias or NULL.  Outputs (DEVICE, any may be
 *   NULL to skip): w_umma bf16 (layout of EdeConv.w_umma for `pack_g`, `passes`; dims from ede_prepack_conv_sizes),
 *   w_direct fp32 [kh*kw][cin_p][cout_p], scale / shift fp32 [cout_p] with
 *   scale = gamma / sqrt(var + eps), shift = beta - mean * scale + scale * bias (fp64 arithmetic, pad lanes 0).
 * ede_prepack_conv_sizes (HOST, no GPU): byte sizes of {w_umma, w_direct, scale, shift}; dims2 = {w_umma_cout_p,
 *   w_umma_cin_p}.  ede_prepack_conv_to: packs into caller-provided buffers (asynchronous, no allocation).
 * ede_prepack_conv: allocates ONE device block for the four buffers and packs into it; ede_free releases it. */
typedef struct EdePrepackConv {
    const float* w_oihw; const int32_t* in_map; const int32_t* out_map;
    const float* bn_weight; const float* bn_bias; const float* bn_mean; const float* bn_var; const float* bias;
    void* w_umma; float* w_direct; float* scale; float* shift;
    float bn_eps;
    int32_t cout, cin, kh, kw, cin_p, cout_p, passes, pack_g;
} EdePrepackConv;
typedef struct EdePackedConv {
    void* w_umma; float* w_direct; float* scale; float* shift;
    int32_t w_umma_cout_p, w_umma_cin_p, w_pack_g, w_passes;
    int64_t device_bytes;
} EdePackedConv;
int ede_prepack_conv_sizes(const EdePrepackConv* a, int64_t* bytes4, int32_t* dims2);
int ede_prepack_conv_to(const EdePrepackConv* a, void* stream);
int ede_prepack_conv(const EdePrepackConv* a, EdePackedConv** handle, void* stream);
int ede_free(EdePackedConv* handle);

/* Tile configuration the tcgen05 path would use for `a` (pure host function, no GPU). */
typedef struct EdeUmmaCfg {
    int32_t eligible;       /* 0: ede_conv2d would take the direct path                          */
    int32_t a_mode;         /* 0: halo tile resident in smem, taps = shifted UMMA windows;
                               1: one TMA box per (tap, channel chunk)                            */
    int32_t ck, swizzle;    /* channels per K chunk (16/32/64) and smem swizzle bytes            */
    int32_t bh, bw, pitch, box_rows;             /* output tile, smem pixel pitch, input rows    */
    int32_t m_tiles, n_pad, n_tiles;             /* UMMA M=128 tiles per CTA, UMMA N, grid.y      */
    int32_t a_stages, b_stages, a_stage_bytes, b_stage_bytes, smem_bytes, tmem_cols;
    int32_t tiles_w, tiles_h, flat;              /* flat: N*H*W folded into one pixel axis (1x1)  */
    double useful_frac;     /* real MACs / issued tensor MACs                                    */
    int32_t swap;           /* 1: weights are the M-side operand (Cout <= 128), 256 pixels on N  */
    int32_t pack_g;         /* kernel-row taps per MMA: 1, 2 or 4 (swapped form) or EDE_PACK_N | G, see w_pack_g */
} EdeUmmaCfg;
int ede_conv2d_umma_config(const EdeConv* a, EdeUmmaCfg* out);

/* Depthwise convolution + folded BN + swish, accumulating per-(n,c) sums for squeeze-excite.
 * replaces: MBConvBlock._depthwise_conv (Conv2dStaticSamePadding, groups=C) -> _bn1 -> swish and
 *   F.adaptive_avg_pool2d(x, 1) of efficientnet-pytorch 0.6.3, reached
 *   from ReSIDE/models/modules.py:157-159.  w: fp32 [kh*kw][c_p].  pool_sum: fp32 [n][c_p],
 *   must be zeroed by the caller (ede_fill_zero) before the call; NULL to skip. */
typedef struct EdeDwConv {
    const void* x; void* y; const float* w; const float* scale; const float* shift; float* pool_sum;
    int32_t n, h, w_, c_p, x_cs, x_co, oh, ow, y_cs, y_co;
    int32_t k, s, pt, pl, act, dtype;            /* dtype of x and y; EDE_F32_TO_BF16X2: fp32 in (as the expand conv of an
                                                    MBConv block can write it), split bf16 out (tensor-core operand) */
    int32_t pool_parts;     /* 0/1: pool_sum is fp32 [n][c_p], accumulated with atomics (order-dependent rounding).
                               P > 1: pool_sum is fp32 [P][n][c_p]; when the kernel runs with <= P column strips each
                               strip WRITES its own slab (no atomics: bit-reproducible forward), else it falls back
                               to atomics on slab 0.  ede_se_fc sums the slabs in order. */
} EdeDwConv;
int ede_dwconv2d(const EdeDwConv* a, void* stream);
/* HOST function, no GPU: the pool_parts value with which ede_dwconv2d writes (instead of atomically adds) the squeeze
 * sums of this shape -- its number of column strips; 1 when the shape takes the atomic path anyway. */
int ede_dwconv2d_pool_parts(const EdeDwConv* a);

/* Squeeze-excite gate: gate[n,c] = sigmoid(W2 . swish(W1 . (pool_sum[n,:]/hw) + b1) + b2).
 * replaces: _se_reduce -> swish -> _se_expand -> sigmoid (App. A.3).  w1: fp32 [csq][c_p],
 *   w2: fp32 [c_p][csq] (rows >= c hold zeros), gate: fp32 [n][c_p]. */
typedef struct EdeSeFc {
    float* pool_sum;        /* read; zeroed afterwards when clear_pool = 1 */
    const float* w1; const float* b1; const float* w2; const float* b2; float* gate;
    int32_t n, c, c_p, csq; float inv_hw;
    int32_t clear_pool;     /* 1: write zeros back to pool_sum after reading it, so that the next forward's
                               ede_dwconv2d accumulates from zero without a separate ede_fill_zero launch */
    int32_t w2_transposed;  /* 1: w2 is fp32 [csq][c_p] (coalesced reads), 0: [c_p][csq]            */
    int32_t pool_parts;     /* slabs of pool_sum (see EdeDwConv.pool_parts), summed in slab order; 0 = 1   */
    /* optional: fold the gate into the weights of the 1x1 conv that consumes the gated tensor (x * g) . W == x . (g * W):
     * w_src = that conv's w_umma (bf16 [w_passes][w_rows][w_cols], hi then lo halves), w_dst = bf16 [n][w_passes][w_rows]
     * [w_cols] with column ci scaled by gate[n][lane(ci)] and re-split into hi / lo; w_split = 1 when the conv reads a
     * split-bf16 input (columns = doubled lanes: groups of 16 = [8 hi][8 lo] of 8 channels).  NULL = off. */
    const void* w_src; void* w_dst;
    int32_t w_rows, w_cols, w_passes, w_split;
} EdeSeFc;
int ede_se_fc(const EdeSeFc* a, void* stream);

/* y[n,h,w,c] = x[n,h,w,c] * gate[n,c]  (the `torch.sigmoid(s) * x` of App. A.3). */
typedef struct EdeSeScale {
    const void* x; void* y; const float* gate;
    int32_t n, hw, c_p, x_cs, x_co, y_cs, y_co, dtype;
} EdeSeScale;
int ede_se_scale(const EdeSeScale* a, void* stream);

/* Bilinear resize, align_corners=True, into a channel slice of the destination.
 * replaces: F.interpolate(x, size, mode='bilinear', align_corners=True) at
 *   ReSIDE/models/modules.py:34, ReSIDE/models/lasinger2019.py:244,289, ReSIDE/train.py:273-276. */
typedef struct EdeResize {
    const void* x; void* y;
    int32_t n, h, w, c_p, x_cs, x_co, oh, ow, y_cs, y_co, x_dtype, y_dtype;
} EdeResize;
int ede_resize_bilinear_ac(const EdeResize* a, void* stream);

/* Same resize on fp32 planes ([planes][h][w] == NCHW): the post-forward depth-map resize of
 * ReSIDE/train.py:273-276, test.py:49-52, test_nyu.py:85, inference_benchmark.py:162. */
int ede_resize_planes_f32(const float* x, float* y, int64_t planes, int32_t h, int32_t w, int32_t oh, int32_t ow,
                          void* stream);

/* Tap gather of "conv(k x k, stride 1) of a bilinearly upsampled map" evaluated on the low-resolution grid.
 * replaces: F.interpolate(x, size, 'bilinear', align_corners=True) followed by conv1/conv2 of _UpProjection
 *   (ReSIDE/models/modules.py:34-37) where the projection is narrow (MFF, modules.py:259-263).
 * z: NHWC [n,h,w] of z_dtype (EDE_F32, or EDE_BF16 with the `rows` workspace: half the bytes of the largest tensor of the
 *   op), lanes [(ty*kw+tx)*cgrp + c] = (W_tap x)[c], produced by ede_conv2d as a 1x1 GEMM.
 *   y[n,oy,ox,c] = act_c( scale[c] * sum_taps inside(oy+ty-pt, ox+tx-pl) * bilinear_ac(z[.., tap, c])(oy+ty-pt, ox+tx-pl)
 *                         + shift[c] ),  act applied to lanes c < act_c_end.
 * rows: optional fp32 workspace [n][h][kh][ow][cgrp]; when given the gather runs in its separable form (x-lerp of
 *   every tap on the h low-resolution rows into `rows`, then the y-lerp), NULL = one-pass 4-corner gather. */
typedef struct EdeUpConv {
    const void* z; void* y; const float* scale; const float* shift; float* rows;
    int32_t n, h, w, z_cs, z_co, oh, ow, cgrp, y_cs, y_co, y_dtype;
    int32_t kh, kw, pt, pl, act, act_c_end;
    int32_t z_dtype;        /* EDE_F32 (0, default) or EDE_BF16 */
} EdeUpConv;
int ede_upconv_taps(const EdeUpConv* a, void* stream);

/* Allocation-free variant of the separable form of ede_upconv_taps: the `rows` workspace is a file-scope device array
 * inside the library (a->rows is ignored), so the call needs no cudaMalloc by anyone.
 * No counterpart in the reference: added for callers that cannot allocate (code
 *   under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Limit: n*h*kh*ow*cgrp*4 bytes <= ede_upconv_taps_static_max_bytes() (16 MiB, held by every GPU that loads the
 *   library until the process exits); larger shapes return EDE_E_ARG with the byte counts in ede_last_error().  Needs an
 *   upsampled map (h < oh) like the separable form.  The first call on a device looks the array's address up
 *   (cudaGetSymbolAddress, that device current); no call allocates or frees device memory.
 * All calls share the one array: it keeps the previous call's contents (pass 1 rewrites every element it later reads,
 *   nothing needs clearing), and two calls that overlap on the GPU (different streams) race on it. */
int ede_upconv_taps_static(const EdeUpConv* a, void* stream);
int64_t ede_upconv_taps_static_max_bytes(void);

/* k x k stride-1 convolution to ONE output channel, second half: sum of the per-tap products that ede_conv2d left in z.
 * replaces: the 5x5 conv to a single channel + bias that ends the Hu decoder (R.conv2, ReSIDE/models/modules.py:285,296).
 * z: fp32 NHWC [n,h,w], lane ty*kw+tx = (W[0,:,ty,tx] . x) at that pixel (a 1x1 GEMM with kh*kw output lanes);
 *   y[n,oy,ox, lane 0] = act(scale[0] * sum_{ty,tx} inside(oy+ty-pt, ox+tx-pl) * z[n, oy+ty-pt, ox+tx-pl, ty*kw+tx] + shift[0]),
 *   lanes 1..7 of the 8-lane output group are written as zeros.  kh*kw <= 32. */
typedef struct EdeTapSum {
    const float* z; void* y; const float* scale; const float* shift;
    int32_t n, h, w, z_cs, z_co, y_cs, y_co, y_dtype, kh, kw, pt, pl, act;
} EdeTapSum;
int ede_tapsum(const EdeTapSum* a, void* stream);

/* 3x3 stride-2 pad-1 max pooling. replaces: nn.MaxPool2d(3,2,1), ReSIDE/models/resnet.py:107. */
typedef struct EdeMaxPool {
    const void* x; void* y;
    int32_t n, h, w, c_p, x_cs, x_co, oh, ow, y_cs, y_co, dtype;
} EdeMaxPool;
int ede_maxpool3x3s2(const EdeMaxPool* a, void* stream);

/* Elementwise y = act(x) (+ x2): the pre-activation ReLU of lasinger2019.py:153 and the
 * `input_from_decoder + res_block1(...)` add of lasinger2019.py:241. x2 may be NULL. */
typedef struct EdeEltwise {
    const void* x; const void* x2; void* y;
    int32_t n_pix, c_p, x_cs, x_co, x2_cs, x2_co, y_cs, y_co, act, dtype;
} EdeEltwise;
int ede_eltwise(const EdeEltwise* a, void* stream);

/* Layout changes at the model boundary (the reference is NCHW fp32 end to end).
 * nchw_to_nhwc: src fp32 [n][c][h][w] -> dst view (dst_dtype), lanes c..c_p-1 zeroed.
 * nhwc_to_nchw: src view -> dst fp32 [n][c][h][w] contiguous. */
typedef struct EdeLayout {
    const void* src; void* dst;
    int32_t n, c, h, w, c_p, cs, co, dtype;
} EdeLayout;
int ede_nchw_to_nhwc(const EdeLayout* a, void* stream);
int ede_nhwc_to_nchw(const EdeLayout* a, void* stream);

int ede_fill_zero(void* p, int64_t bytes, void* stream);

/* The network's first convolution straight from the fp32 NCHW image (CUDA cores, fp32 exact), fused with the
 * folded BN, the activation and the NCHW -> NHWC layout change.
 * replaces: conv1 7x7 s2 p3 + bn1 + ReLU of ReSIDE/models/resnet.py:103-106 and the EfficientNet stem
 *   (conv_stem 3x3 s2 static-same padding + bn0; modules.py:185 keeps no swish after it).
 * x: fp32 [n][cin][h][w] (cin <= 4); w: fp32 [k*k*cin][cout_p] with row (ky*k+kx)*cin+ci; y: NHWC view.
 * Supported shapes: ede_stem_supported(k, s, cin, cout_p) (3x3 s2 -> 32/48/64 lanes, 7x7 s2 -> 64 lanes). */
typedef struct EdeStem {
    const float* x; void* y; const float* w; const float* scale; const float* shift;
    int32_t n, cin, h, w_, oh, ow, cout_p, y_cs, y_co, k, s, pt, pl, act, y_dtype;
} EdeStem;
int ede_stem_conv(const EdeStem* a, void* stream);
int ede_stem_supported(int32_t k, int32_t s, int32_t cin, int32_t cout_p);

/* Sobel gradients. replaces: Sobel.forward, ReSIDE/sobel.py:19-23 ([n,1,h,w] -> [n,2,h,w], fp32 NCHW). */
int ede_sobel(const float* x, float* y, int32_t n, int32_t h, int32_t w, void* stream);

/* Fused Sobel depth/gradient/normal loss + depth metrics: ONE pass over (pred, label).
 * replaces: ReSIDE/train.py:216-236 (loss) and MetricsTracker.update, ReSIDE/util.py:39-92.
 * pred/label: fp32 [n,1,h,w].  out: 16 doubles, ACCUMULATED (zero it first):
 *  [0] sum log(|o-d|+.5)  [1] sum |1-cos|  [2] sum log(|odx-ddx|+.5)  [3] sum log(|ody-ddy|+.5)
 *  [4] sum |o-d|  [5] sum (o-d)^2  [6] sum rel (masked)  [7] sum |log10 o - log10 d| (masked)
 *  [8..10] #(max(o/d,d/o) <= 1.25^k)  [11] #non-NaN labels  [12] #pixels  [13] #images [14..15] 0 */
int ede_sobel_loss_metrics(const float* pred, const float* label, int32_t n, int32_t h, int32_t w,
                           double* out16, void* stream);

/* Edge-detection counts of ReSIDE/test.py:56-76,94-102: sqrt(gx^2+gy^2) > thr for both maps.
 * out: 4 doubles accumulated: [0] #pred edges [1] #label edges [2] #both [3] #equal. */
int ede_edge_counts(const float* pred, const float* label, int32_t n, int32_t h, int32_t w,
                    float thr, double* out4, void* stream);

/* Fused post-forward epilogue: bilinear align_corners resize of the depth map to the label / image
 * size, consumed in the same pass -- nothing but the requested outputs is written.
 * replaces: F.interpolate(output, size, 'bilinear', align_corners=True) + MetricsTracker.update at
 *   ReSIDE/train.py:273-278, ReSIDE/test.py:49-54; edge_detection + thresholds at ReSIDE/test.py:56-76,94-102;
 *   the cap / millimetre / uint16 conversion at ReSIDE/test_nyu.py:85-88; the loss of train.py:216-236 when the
 *   label is at the resized size.
 * pred: fp32 [n,1,h,w].  Optional (NULL = skip): label fp32 [n,1,oh,ow]; resized fp32 [n,1,oh,ow]; depth_mm
 * uint16 [n,oh,ow] = (uint16)(int32)((r > cap ? 0 : r) * mm_scale); sums16 (layout of ede_sobel_loss_metrics,
 * accumulated) and edges4 (layout of ede_edge_counts, accumulated) both need `label`. */
typedef struct EdeDepthEpilogue {
    const float* pred; const float* label;
    float* resized; uint16_t* depth_mm; double* sums16; double* edges4;
    int32_t n, h, w, oh, ow;
    float edge_thr, cap, mm_scale;
} EdeDepthEpilogue;
int ede_depth_epilogue(const EdeDepthEpilogue* a, void* stream);

/* Video-frame preprocessing in front of the forward path, one kernel, bit-exact against Pillow:
 * uint8 HWC RGB frames [n,ih,iw,3] -> Scale to (sw,sh) with Image.BILINEAR -> CenterCrop (cw,ch) -> /255 ->
 * (x-mean)/std -> fp32 NCHW [n,3,ch,cw].
 * replaces: transforms.Compose([Scale([320,240]), CenterCrop([304,228]), ToTensor(), Normalize(mean,std)]) at
 *   ReSIDE/inference_benchmark.py:46-52, ReSIDE/demo.py / demo_transform.py:23-127 (PIL + torch on the CPU).
 * ede_resample_coeffs (HOST function, no GPU): Pillow's fixed-point coefficient table for output samples
 *   [first, first+count) of an in_size -> out_size BILINEAR resample; returns ksize (> 0), or < 0 on error; with
 *   NULL tables it only returns ksize.  bounds[2*i] = first input sample, bounds[2*i+1] = tap count,
 *   kk[i*ksize + t] = weights scaled by 2^22.
 * ede_preprocess_frames: xb/xk = DEVICE copies of the tables of the cropped columns, yb/yk of the cropped rows;
 *   band = output rows per block; span >= input rows any band touches (shared memory = span*cw*3 bytes). */
typedef struct EdePreprocess {
    const uint8_t* frames; float* y;
    const int32_t* xb; const int32_t* xk; const int32_t* yb; const int32_t* yk;
    int32_t n, ih, iw, ch, cw, xks, yks, band, span;
    float mean[3], std[3];
} EdePreprocess;
int ede_resample_coeffs(int32_t in_size, int32_t out_size, int32_t first, int32_t count, int32_t* bounds, int32_t* kk);
int ede_preprocess_frames(const EdePreprocess* a, void* stream);

/* Depth-mesh rendering as a CUDA rasteriser (no graphics API).
 * replaces: Mesh.from_texture (DepthRenderer/DepthRenderer/render.py:465-546, + the displacement of
 *   Benchmark/benchmark.py:949), the vertex / fragment shaders (DepthRenderer/DepthRenderer/shaders/shader.vert:1-14,
 *   shader.frag:1-8) and the fixed-function pipeline they run in (render.py:357-360 GL_LINEAR + GL_CLAMP, :631-634 cull
 *   back faces + depth test + clear to opaque black, :787 glReadPixels RGBA8), driven per frame by
 *   MeshRenderer.draw (render.py:800-823) from Benchmark/benchmark.py:845-1039 (create_rendered_images).
 * ede_mesh_from_depth: depth uint8 [h][w] (element stride depth_stride: 1, or the channel count when the map is HxWxC
 *   and channel 0 is used) -> vertices fp32 [(2^d+1)^2][3], uv fp32 [..][2], indices uint32 [2 * 4^d][3].
 * ede_render_mesh: n_frames camera poses (mvp fp32 [n_frames][16], ROW-major as render.py:812 uploads them with
 *   transpose = GL_TRUE) of ONE mesh -> frames uint8 [n_frames][height][width][4]; row 0 = bottom of the window as
 *   glReadPixels returns it, or the top when flip_y = 1.  texture: uint8 [tex_h][tex_w][4], row 0 = v 0.
 *   workspace: DEVICE scratch of ede_render_workspace_bytes() bytes (HOST function).  Rasterisation rules (sub-pixel
 *   snapping, fill rule, depth format, texel rounding): oracle/render_cpu.py, matched bit for bit. */
typedef struct EdeMeshFromDepth {
    const uint8_t* depth; float* vertices; float* uv; uint32_t* indices;
    int32_t h, w, depth_stride, density;
    float displacement;
} EdeMeshFromDepth;
int ede_mesh_from_depth(const EdeMeshFromDepth* a, void* stream);
typedef struct EdeRender {
    const float* vertices; const float* uv; const uint32_t* indices; const float* mvp; const uint8_t* texture;
    uint8_t* frames; void* workspace;
    int32_t n_vertices, n_triangles, n_frames, tex_h, tex_w, height, width, flip_y;
} EdeRender;
int64_t ede_render_workspace_bytes(const EdeRender* a);
int ede_render_mesh(const EdeRender* a, void* stream);

/* Run a pre-built operator list ("forward plan") back to back on one stream. */
enum { EDE_OP_CONV = 1, EDE_OP_DWCONV = 2, EDE_OP_SE_FC = 3, EDE_OP_SE_SCALE = 4, EDE_OP_RESIZE = 5,
       EDE_OP_MAXPOOL = 6, EDE_OP_ELTWISE = 7, EDE_OP_TO_NHWC = 8, EDE_OP_TO_NCHW = 9, EDE_OP_ZERO = 10,
       EDE_OP_UPCONV = 11, EDE_OP_STEM = 12, EDE_OP_TAPSUM = 13 };
typedef struct EdeOp {
    int32_t kind; int32_t _pad;
    union {
        EdeConv conv; EdeDwConv dw; EdeSeFc se_fc; EdeSeScale se_scale; EdeResize resize;
        EdeMaxPool maxpool; EdeEltwise eltwise; EdeLayout layout; EdeUpConv upconv; EdeStem stem; EdeTapSum tapsum;
        struct { void* p; int64_t bytes; } zero;
    } u;
} EdeOp;
int ede_plan_run(const EdeOp* ops, int32_t n_ops, void* stream);
/* number of kernel launches issued by this library on the calling thread since the last reset */
int64_t ede_launch_count(int32_t reset);

/* debug hook (env EDE_UMMA_PROF=1): per-CTA globaltimer stamps of the last tcgen05 conv launch, 8 x u64 per
 * CTA [entry, mainloop start, first operands landed, last MMA issued, accumulator ready, epilogue done,
 * CTA done, -]; synchronises the device; returns the CTA count (0 when profiling is off). */
int64_t ede_debug_umma_prof(unsigned long long* host, int64_t max_ctas);

int ede_version(void);
const char* ede_last_error(void);
/* 1 when the tcgen05/TMA path is usable (driver entry point for tensor maps resolved, sm_100). */
int ede_umma_available(void);

#ifdef __cplusplus
}
#endif
#endif /* EDE_H_ */
