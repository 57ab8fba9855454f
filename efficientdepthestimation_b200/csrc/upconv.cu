// upconv.cu -- "convolution of a bilinearly upsampled map" evaluated on the LOW-resolution grid.
//
// The Hu decoder's MFF up-projections (ReSIDE/models/modules.py:33-37,259-263) upsample an encoder map
// x[Cin,h,w] to the 114x152 decoder grid and then run two 5x5 convolutions with only 16 outputs each.
// Both steps are linear, so   conv5x5(up(x))[p] = sum_t up(W_t x)[p + t]   (zero outside the upsampled
// grid): the Cin contraction can run ONCE per low-resolution pixel as a 1x1 GEMM z = [W_t]_t x with
// 25*32 output lanes (tensor cores, N = 200..256 per tile instead of N = 32), and this kernel gathers the
// 25 taps: for every output pixel and tap it bilinearly samples (align_corners=True) the tap's z-plane at
// the tap-shifted position, sums, applies the folded BatchNorm / ReLU and stores NHWC.  Exact in real
// arithmetic; fp32 z keeps it at least as accurate as convolving a bf16 upsampled map.
#include "common.cuh"

namespace ede {

template <int K, typename TY>
__global__ void __launch_bounds__(256) upconv_taps_kernel(const EdeUpConv a, float scale_h, float scale_w) {
    const int ngrp = a.cgrp >> 3;
    const long long total = (long long)a.n * a.oh * a.ow * ngrp;
    const float* __restrict__ z = a.z;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % ngrp) << 3;
        long long pix = i / ngrp;
        const int ox = (int)(pix % a.ow);
        pix /= a.ow;
        const int oy = (int)(pix % a.oh);
        const int n = (int)(pix / a.oh);
        // per-axis sampling tables of the K tap rows / columns (ATen upsample_bilinear2d, align_corners)
        int y0[K], y1[K], x0[K], x1[K];
        float wy0[K], wy1[K], wx0[K], wx1[K];
#pragma unroll
        for (int t = 0; t < K; ++t) {
            const int qy = oy + t - a.pt, qx = ox + t - a.pl;
            const bool vy = qy >= 0 && qy < a.oh, vx = qx >= 0 && qx < a.ow;
            const float sy = scale_h * (vy ? qy : 0), sx = scale_w * (vx ? qx : 0);
            y0[t] = (int)sy; x0[t] = (int)sx;
            y1[t] = y0[t] + (y0[t] < a.h - 1 ? 1 : 0);
            x1[t] = x0[t] + (x0[t] < a.w - 1 ? 1 : 0);
            const float ly = sy - y0[t], lx = sx - x0[t];
            wy0[t] = vy ? 1.f - ly : 0.f; wy1[t] = vy ? ly : 0.f;       // taps in the zero padding contribute 0
            wx0[t] = vx ? 1.f - lx : 0.f; wx1[t] = vx ? lx : 0.f;
        }
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        const long long img = (long long)n * a.h;
#pragma unroll
        for (int ty = 0; ty < K; ++ty) {
            const float* r0 = z + (img + y0[ty]) * a.w * (long long)a.z_cs + a.z_co + c;
            const float* r1 = z + (img + y1[ty]) * a.w * (long long)a.z_cs + a.z_co + c;
#pragma unroll
            for (int tx = 0; tx < K; ++tx) {
                const int lane = (ty * K + tx) * a.cgrp;
                float v00[8], v01[8], v10[8], v11[8];
                load8<float>(r0 + (long long)x0[tx] * a.z_cs + lane, v00);
                load8<float>(r0 + (long long)x1[tx] * a.z_cs + lane, v01);
                load8<float>(r1 + (long long)x0[tx] * a.z_cs + lane, v10);
                load8<float>(r1 + (long long)x1[tx] * a.z_cs + lane, v11);
                const float w00 = wy0[ty] * wx0[tx], w01 = wy0[ty] * wx1[tx];
                const float w10 = wy1[ty] * wx0[tx], w11 = wy1[ty] * wx1[tx];
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    acc[j] = fmaf(w00, v00[j], fmaf(w01, v01[j], fmaf(w10, v10[j], fmaf(w11, v11[j], acc[j]))));
            }
        }
        float sc[8], sf[8];
        load8<float>(a.scale + c, sc);
        load8<float>(a.shift + c, sf);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            acc[j] = fmaf(acc[j], sc[j], sf[j]);
            if (c < a.act_c_end) acc[j] = apply_act(acc[j], a.act);
        }
        store8<TY>(reinterpret_cast<TY*>(a.y) + (((long long)n * a.oh + oy) * a.ow + ox) * a.y_cs + a.y_co + c, acc);
    }
}

}  // namespace ede

using namespace ede;

extern "C" int ede_upconv_taps(const EdeUpConv* a, void* stream) {
    EDE_REQUIRE(a && a->z && a->y && a->scale && a->shift, "ede_upconv_taps: null argument");
    EDE_REQUIRE(((a->cgrp | a->z_cs | a->z_co | a->y_cs | a->y_co | a->act_c_end) & 7) == 0,
                "ede_upconv_taps: channel counts must be multiples of 8");
    EDE_REQUIRE(a->n > 0 && a->h > 0 && a->w > 0 && a->oh > 0 && a->ow > 0 && a->cgrp > 0, "ede_upconv_taps: bad shape");
    EDE_REQUIRE(a->kh == a->kw && (a->kh == 3 || a->kh == 5), "ede_upconv_taps: only 3x3 and 5x5 taps are supported");
    const float sh = a->oh > 1 ? (float)(a->h - 1) / (float)(a->oh - 1) : 0.f;
    const float sw = a->ow > 1 ? (float)(a->w - 1) / (float)(a->ow - 1) : 0.f;
    const long long total = (long long)a->n * a->oh * a->ow * (a->cgrp >> 3);
    long long g = (total + 255) / 256;
    if (g > 148LL * 64) g = 148LL * 64;
    cudaStream_t st = (cudaStream_t)stream;
    if (a->kh == 5) {
        EDE_DISPATCH_DTYPE(a->y_dtype, TY, (upconv_taps_kernel<5, TY><<<(unsigned)g, 256, 0, st>>>(*a, sh, sw)));
    } else {
        EDE_DISPATCH_DTYPE(a->y_dtype, TY, (upconv_taps_kernel<3, TY><<<(unsigned)g, 256, 0, st>>>(*a, sh, sw)));
    }
    EDE_LAUNCH_CHECK("upconv_taps_kernel");
    return EDE_OK;
}
