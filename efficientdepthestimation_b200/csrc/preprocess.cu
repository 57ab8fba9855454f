// preprocess.cu -- video-frame preprocessing in front of the forward path.
//
// uint8 HWC frame -> Scale (Pillow BILINEAR resample: antialiased triangle filter, 22-bit fixed point, horizontal
// pass rounded to uint8, then vertical pass) -> CenterCrop -> /255 -> (x - mean) / std -> fp32 NCHW, in ONE kernel.
// Only the cropped window is ever computed; the horizontally resampled rows a band of output rows needs live in
// shared memory as uint8, so the frame is read once and nothing but the network input is written.
//
// The arithmetic is integer / byte work until the last two fp32 operations, and is bit-exact against Pillow:
//   coefficient tables (host, double):   precompute_coeffs + normalize_coeffs_8bpc of Pillow's Resample.c as used by
//                                        Image.resize(size, Image.BILINEAR) at ReSIDE/demo_transform.py:31-34
//   per-pixel:                           ss = 1 << 21; ss += px * k (int32); out = clip8(ss >> 22)
//   ToTensor / Normalize:                float(u8) / 255 -> sub mean -> div std, each rounded to fp32
//                                        (ReSIDE/demo_transform.py:72-127; IEEE division, no FMA contraction)
#include <math.h>

#include "common.cuh"

namespace ede {

constexpr int PRECISION_BITS = 32 - 8 - 2;

// Pillow's coefficient table for output samples [first, first + count) of an in_size -> out_size resample with the
// triangle (BILINEAR, support 1) filter.  Returns ksize; fills bounds[2*i] = xmin, bounds[2*i+1] = tap count and
// kk[i*ksize + t] = fixed-point weights (when the pointers are non-null).
static int resample_coeffs(int in_size, int out_size, int first, int count, int32_t* bounds, int32_t* kk) {
    const double scale = (double)in_size / (double)out_size;
    const double filterscale = scale < 1.0 ? 1.0 : scale;
    const double support = 1.0 * filterscale;
    const int ksize = (int)ceil(support) * 2 + 1;
    if (!bounds || !kk) return ksize;
    const double ss = 1.0 / filterscale;
    double* k = (double*)malloc(sizeof(double) * (size_t)ksize);
    if (!k) return -1;
    for (int i = 0; i < count; ++i) {
        const int xx = first + i;
        const double center = (xx + 0.5) * scale;
        double ww = 0.0;
        int xmin = (int)(center - support + 0.5);
        if (xmin < 0) xmin = 0;
        int xmax = (int)(center + support + 0.5);
        if (xmax > in_size) xmax = in_size;
        xmax -= xmin;
        for (int x = 0; x < xmax; ++x) {
            double t = (x + xmin - center + 0.5) * ss;
            if (t < 0.0) t = -t;
            const double w = t < 1.0 ? 1.0 - t : 0.0;
            k[x] = w;
            ww += w;
        }
        for (int x = 0; x < xmax; ++x)
            if (ww != 0.0) k[x] /= ww;
        for (int x = 0; x < ksize; ++x) {
            const double v = x < xmax ? k[x] : 0.0;
            kk[(size_t)i * ksize + x] = v < 0 ? (int32_t)(-0.5 + v * (1 << PRECISION_BITS)) : (int32_t)(0.5 + v * (1 << PRECISION_BITS));
        }
        bounds[2 * i] = xmin;
        bounds[2 * i + 1] = xmax;
    }
    free(k);
    return ksize;
}

__device__ __forceinline__ int clip8(int v) {
    v >>= PRECISION_BITS;
    return v < 0 ? 0 : (v > 255 ? 255 : v);
}

// grid = (bands, n); block = 256 threads = 8 warps.  Dynamic smem: span * cw * 3 bytes (uint8 rows after the
// horizontal pass).  No per-element integer division: warps own rows, lanes sweep columns; the two fp32 operations of
// ToTensor/Normalize are a 3 x 256-entry table per block (the input of the float math is one byte).
__global__ void __launch_bounds__(256) preprocess_kernel(const EdePreprocess a) {
    extern __shared__ unsigned char rows[];                      // [span][cw][3]
    __shared__ float lut[3][256];
    const int n = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r0 = blockIdx.x * a.band, r1 = min(r0 + a.band, a.ch);
    for (int i = threadIdx.x; i < 768; i += 256) {
        const int ch = i >> 8;
        const float m = ch == 0 ? a.mean[0] : (ch == 1 ? a.mean[1] : a.mean[2]);
        const float sd = ch == 0 ? a.std[0] : (ch == 1 ? a.std[1] : a.std[2]);
        lut[ch][i & 255] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)(i & 255), 255.f), m), sd);
    }
    // input rows this band touches
    const int in0 = a.yb[2 * r0];
    const int nrows = a.yb[2 * (r1 - 1)] + a.yb[2 * (r1 - 1) + 1] - in0;
    const unsigned char* __restrict__ f = a.frames + ((long long)n * a.ih + in0) * a.iw * 3;
    // horizontal pass: a lane keeps its column's window, the warp's rows stream through it
    for (int c = lane; c < a.cw; c += 32) {
        const int xmin = a.xb[2 * c], cnt = a.xb[2 * c + 1];
        const int32_t* __restrict__ k = a.xk + c * a.xks;
        for (int r = warp; r < nrows; r += 8) {
            const unsigned char* __restrict__ p = f + ((long long)r * a.iw + xmin) * 3;
            int s0 = 1 << (PRECISION_BITS - 1), s1 = s0, s2 = s0;
            for (int t = 0; t < cnt; ++t) {
                const int kv = k[t];
                s0 += p[3 * t + 0] * kv;
                s1 += p[3 * t + 1] * kv;
                s2 += p[3 * t + 2] * kv;
            }
            unsigned char* o = rows + (r * a.cw + c) * 3;
            o[0] = (unsigned char)clip8(s0);
            o[1] = (unsigned char)clip8(s1);
            o[2] = (unsigned char)clip8(s2);
        }
    }
    __syncthreads();
    // vertical pass + ToTensor + Normalize: warps own (channel, row) pairs, lanes sweep the columns of a plane row
    const int nr = r1 - r0;
    for (int q = warp; q < nr * 3; q += 8) {
        const int ch = q / nr, r = q - ch * nr;
        const int oy = r0 + r;
        const int ymin = a.yb[2 * oy] - in0, cnt = a.yb[2 * oy + 1];
        const int32_t* __restrict__ k = a.yk + oy * a.yks;
        float* __restrict__ y = a.y + (((long long)n * 3 + ch) * a.ch + oy) * a.cw;
        const unsigned char* __restrict__ src = rows + ymin * a.cw * 3 + ch;
        for (int c = lane; c < a.cw; c += 32) {
            int s = 1 << (PRECISION_BITS - 1);
            for (int t = 0; t < cnt; ++t) s += src[(t * a.cw + c) * 3] * k[t];
            y[c] = lut[ch][clip8(s)];
        }
    }
}

}  // namespace ede

using namespace ede;

extern "C" {

int ede_resample_coeffs(int32_t in_size, int32_t out_size, int32_t first, int32_t count, int32_t* bounds, int32_t* kk) {
    EDE_REQUIRE(in_size > 0 && out_size > 0 && first >= 0 && count >= 0 && first + count <= out_size,
                "ede_resample_coeffs: bad arguments");
    const int ks = resample_coeffs(in_size, out_size, first, count, bounds, kk);
    EDE_REQUIRE(ks > 0, "ede_resample_coeffs: out of memory");
    return ks;
}

int ede_preprocess_frames(const EdePreprocess* a, void* stream) {
    EDE_REQUIRE(a && a->frames && a->y && a->xb && a->xk && a->yb && a->yk, "ede_preprocess_frames: null argument");
    EDE_REQUIRE(a->n > 0 && a->ih > 0 && a->iw > 0 && a->ch > 0 && a->cw > 0 && a->xks > 0 && a->yks > 0 && a->band > 0 && a->span > 0,
                "ede_preprocess_frames: bad shape");
    const size_t smem = (size_t)a->span * a->cw * 3;
    EDE_REQUIRE(smem <= 200 * 1024, "ede_preprocess_frames: band of %d rows x %d columns needs %zu bytes of shared memory", a->span, a->cw, smem);
    if (smem > 48 * 1024) {
        int rc = check_cuda(cudaFuncSetAttribute(preprocess_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                            "cudaFuncSetAttribute(preprocess_kernel)");
        if (rc) return rc;
    }
    dim3 grid(ceil_div(a->ch, a->band), a->n);
    preprocess_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(*a);
    EDE_LAUNCH_CHECK("preprocess_kernel");
    return EDE_OK;
}

}  // extern "C"
