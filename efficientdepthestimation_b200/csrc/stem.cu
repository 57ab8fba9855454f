// stem.cu -- the network's first convolution, straight from the fp32 NCHW input image.
// replaces: ResNet conv1 7x7 s2 p3 (3 -> 64) + bn1 + ReLU (ReSIDE/models/resnet.py:103-106) and the EfficientNet
//   stem Conv2dStaticSamePadding 3x3 s2 (3 -> 32/48) + bn0, and with them the NCHW -> NHWC
//   layout pass that used to precede them.
// With 3 input channels the GEMM K is 27 / 147: far too thin for the tensor cores (the implicit-GEMM form padded
// every pixel to a 128-byte row and was TMA-bound), so this is a CUDA-core kernel shaped for bandwidth:
//   * a block stages its input patch (3 x rows x cols, zero padded, columns de-interleaved by stride phase so that
//     stride-2 reads are bank-conflict free) and all weights in shared memory;
//   * a thread owns two output pixels and ALL output channels (2 x COUT fp32 accumulators): per tap it reads two
//     inputs and COUT/4 broadcast weight vectors for 2*COUT FMAs;
//   * folded BN + activation, then each pixel's channels leave as contiguous 32-byte vectors in NHWC.
// fp32 inputs and weights, fp32 accumulation: exact in every precision mode.
#include <algorithm>

#include "common.cuh"

namespace ede {

struct StemTile {
    int lw, th;            // column pairs per tile row (tile width = 2*lw), tile rows; lw*th <= 256
    int idx_w;             // staged columns per stride phase
    int ih;                // staged input rows
};

template <typename TY, int K, int S, int COUT>
__global__ void __launch_bounds__(256) stem_conv_kernel(const EdeStem a, const StemTile tl) {
    extern __shared__ float sm[];
    const int cin = a.cin;
    float* const ws = sm;                                        // [K*K*cin][COUT]
    float* const xs = sm + K * K * cin * COUT;                   // [cin][ih][S][idx_w]
    const int t = threadIdx.x;
    const int n = blockIdx.z, oh0 = blockIdx.y * tl.th, ow0 = blockIdx.x * (2 * tl.lw);
    pdl_trigger();

    for (int i = t; i < K * K * cin * COUT; i += 256) {
        const int r = i / COUT, c = i - r * COUT;
        ws[i] = c < a.cout_p ? a.w[(long long)r * a.cout_p + c] : 0.f;
    }
    pdl_wait();                                                  // (the weights above are static; x and y are not)
    const int iw_span = tl.idx_w * S;
    const int ih_base = oh0 * S - a.pt, iw_base = ow0 * S - a.pl;
    for (int i = t; i < cin * tl.ih * iw_span; i += 256) {
        const int il = i % iw_span;
        const int r = i / iw_span;
        const int hl = r % tl.ih, ci = r / tl.ih;
        const int ihh = ih_base + hl, iww = iw_base + il;
        float v = 0.f;
        if (ihh >= 0 && ihh < a.h && iww >= 0 && iww < a.w_)
            v = a.x[(((long long)n * cin + ci) * a.h + ihh) * a.w_ + iww];
        xs[((ci * tl.ih + hl) * S + (il % S)) * tl.idx_w + il / S] = v;
    }
    __syncthreads();

    const int row = t / tl.lw, cp = t - row * tl.lw;
    if (row >= tl.th) return;
    float acc0[COUT], acc1[COUT];
#pragma unroll
    for (int c = 0; c < COUT; ++c) { acc0[c] = 0.f; acc1[c] = 0.f; }
    for (int ci = 0; ci < cin; ++ci) {
#pragma unroll
        for (int ky = 0; ky < K; ++ky) {
            const float* xr = xs + ((ci * tl.ih + row * S + ky) * S) * tl.idx_w;
#pragma unroll
            for (int kx = 0; kx < K; ++kx) {
                const float* xp = xr + (kx % S) * tl.idx_w + kx / S;
                const float x0 = xp[cp], x1 = xp[cp + tl.lw];
                const float4* wv = reinterpret_cast<const float4*>(ws + ((ky * K + kx) * cin + ci) * COUT);
#pragma unroll
                for (int c4 = 0; c4 < COUT / 4; ++c4) {
                    const float4 w = wv[c4];
                    acc0[4 * c4] = fmaf(x0, w.x, acc0[4 * c4]); acc0[4 * c4 + 1] = fmaf(x0, w.y, acc0[4 * c4 + 1]);
                    acc0[4 * c4 + 2] = fmaf(x0, w.z, acc0[4 * c4 + 2]); acc0[4 * c4 + 3] = fmaf(x0, w.w, acc0[4 * c4 + 3]);
                    acc1[4 * c4] = fmaf(x1, w.x, acc1[4 * c4]); acc1[4 * c4 + 1] = fmaf(x1, w.y, acc1[4 * c4 + 1]);
                    acc1[4 * c4 + 2] = fmaf(x1, w.z, acc1[4 * c4 + 2]); acc1[4 * c4 + 3] = fmaf(x1, w.w, acc1[4 * c4 + 3]);
                }
            }
        }
    }
    const int oh = oh0 + row;
    if (oh >= a.oh) return;
#pragma unroll
    for (int px = 0; px < 2; ++px) {
        const int ow = ow0 + cp + px * tl.lw;
        if (ow >= a.ow) continue;
        TY* yrow = reinterpret_cast<TY*>(a.y) + (((long long)n * a.oh + oh) * a.ow + ow) * a.y_cs + a.y_co;
#pragma unroll
        for (int c = 0; c < COUT; c += 8) {
            if (c >= a.cout_p) break;
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                v[j] = apply_act(fmaf(px ? acc1[c + j] : acc0[c + j], __ldg(a.scale + c + j), __ldg(a.shift + c + j)), a.act);
            store8<TY>(yrow + c, v);
        }
    }
}

template <typename TY, int K, int S, int COUT>
static int stem_launch_t(const EdeStem& a, cudaStream_t st) {
    StemTile tl;
    const int wt = ceil_div(a.ow, 64);                           // tiles across: widest tile is 64 columns
    tl.lw = std::min(32, ceil_div(ceil_div(a.ow, wt), 2));
    tl.th = std::max(1, std::min(256 / tl.lw, a.oh));
    tl.idx_w = 2 * tl.lw + (K - 1) / S + 1;
    tl.ih = tl.th * S + K - S;
    const size_t smem = ((size_t)K * K * a.cin * COUT + (size_t)a.cin * tl.ih * S * tl.idx_w) * sizeof(float);
    if (smem > 200 * 1024) return 1;
    static bool attr_set[64] = {};
    int dev = 0;
    (void)cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!attr_set[dev]) {
        if (cudaFuncSetAttribute(stem_conv_kernel<TY, K, S, COUT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 200 * 1024) != cudaSuccess) {
            (void)cudaGetLastError();
            return 1;
        }
        attr_set[dev] = true;
    }
    dim3 grid((unsigned)ceil_div(a.ow, 2 * tl.lw), (unsigned)ceil_div(a.oh, tl.th), (unsigned)a.n);
    if (grid.y > 65535u || grid.z > 65535u) return 1;
    if (launch_pdl(stem_conv_kernel<TY, K, S, COUT>, grid, dim3(256), smem, st, a, tl) != cudaSuccess) {
        (void)cudaGetLastError();
        return 1;
    }
    return 0;
}

template <typename TY>
static int stem_launch(const EdeStem& a, cudaStream_t st) {
    const int key = a.k * 1000 + a.s * 100 + round_up(a.cout_p, 16);
    switch (key) {
        case 3232: return stem_launch_t<TY, 3, 2, 32>(a, st);    // EfficientNet-B0 stem
        case 3248: return stem_launch_t<TY, 3, 2, 48>(a, st);    // EfficientNet-B4 stem
        case 3264: return stem_launch_t<TY, 3, 2, 64>(a, st);
        case 7264: return stem_launch_t<TY, 7, 2, 64>(a, st);    // ResNet conv1
        default: return 1;
    }
}

}  // namespace ede

using namespace ede;

extern "C" {

int ede_stem_supported(int32_t k, int32_t s, int32_t cin, int32_t cout_p) {
    const int key = k * 1000 + s * 100 + (cout_p + 15) / 16 * 16;
    return cin >= 1 && cin <= 4 && (key == 3232 || key == 3248 || key == 3264 || key == 7264) ? 1 : 0;
}

int ede_stem_conv(const EdeStem* a, void* stream) {
    EDE_REQUIRE(a && a->x && a->y && a->w && a->scale && a->shift, "ede_stem_conv: null argument");
    EDE_REQUIRE(a->n > 0 && a->h > 0 && a->w_ > 0 && a->oh > 0 && a->ow > 0, "ede_stem_conv: bad shape");
    EDE_REQUIRE(((a->cout_p | a->y_cs | a->y_co) & 7) == 0, "ede_stem_conv: channel counts must be multiples of 8");
    EDE_REQUIRE(ede_stem_supported(a->k, a->s, a->cin, a->cout_p), "ede_stem_conv: unsupported stem %dx%d s%d %d->%d",
                a->k, a->k, a->s, a->cin, a->cout_p);
    int rc = 1;
    EDE_DISPATCH_DTYPE(a->y_dtype, TY, (rc = stem_launch<TY>(*a, (cudaStream_t)stream)));
    EDE_REQUIRE(rc == 0, "ede_stem_conv: launch configuration failed");
    EDE_LAUNCH_CHECK("stem_conv_kernel");
    return EDE_OK;
}

}  // extern "C"
