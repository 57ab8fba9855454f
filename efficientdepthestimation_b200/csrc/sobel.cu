// sobel.cu -- Sobel gradients and the fused depth/gradient/normal loss + metric reduction.
//
// One pass over (pred, label): each block stages a (TH+2)x(TW+2) zero-padded tile of both
// maps in shared memory, every thread evaluates the two 3x3 stencils for its pixels and the
// 13 partial sums are reduced warp-shuffle -> shared -> one double atomicAdd per block.
#include "common.cuh"

namespace ede {

constexpr int STW = 64, STH = 16;       // tile of output pixels per block (256 threads x 4 pixels)

__device__ __forceinline__ void sobel_at(const float (*t)[STW + 2], int y, int x, float& gx, float& gy) {
    // cross-correlation with ReSIDE/sobel.py:9-10, (y,x) are tile coords of the centre + 1
    const float a = t[y - 1][x - 1], b = t[y - 1][x], c = t[y - 1][x + 1];
    const float d = t[y][x - 1], e = t[y][x], f = t[y][x + 1];
    const float g = t[y + 1][x - 1], h = t[y + 1][x], i = t[y + 1][x + 1];
    // the zero taps stay in the sum: Conv2d multiplies them too, so a NaN/Inf anywhere in the 3x3 window
    // poisons both gradients exactly as in the reference (0 * NaN = NaN); for finite inputs they add +0
    gx = (a - c) + 2.f * (d - f) + (g - i) + (0.f * b + 0.f * e + 0.f * h);
    gy = (a + 2.f * b + c) - (g + 2.f * h + i) + (0.f * d + 0.f * e + 0.f * f);
}

__device__ __forceinline__ void load_tile(float (*t)[STW + 2], const float* img, int h, int w, int y0, int x0) {
    for (int i = threadIdx.x; i < (STH + 2) * (STW + 2); i += blockDim.x) {
        const int ty = i / (STW + 2), tx = i - ty * (STW + 2);
        const int y = y0 + ty - 1, x = x0 + tx - 1;
        t[ty][tx] = (y >= 0 && y < h && x >= 0 && x < w) ? img[(long long)y * w + x] : 0.f;
    }
}

__global__ void __launch_bounds__(256) sobel_kernel(const float* __restrict__ x, float* __restrict__ y, int h, int w) {
    __shared__ float t[STH + 2][STW + 2];
    const int n = blockIdx.z, y0 = blockIdx.y * STH, x0 = blockIdx.x * STW;
    load_tile(t, x + (long long)n * h * w, h, w, y0, x0);
    __syncthreads();
    for (int i = threadIdx.x; i < STH * STW; i += blockDim.x) {
        const int ty = i / STW, tx = i - ty * STW;
        const int py = y0 + ty, px = x0 + tx;
        if (py < h && px < w) {
            float gx, gy;
            sobel_at(t, ty + 1, tx + 1, gx, gy);
            y[(((long long)n * 2 + 0) * h + py) * w + px] = gx;
            y[(((long long)n * 2 + 1) * h + py) * w + px] = gy;
        }
    }
}

constexpr int NSUM = 13;

__global__ void __launch_bounds__(256) sobel_loss_metrics_kernel(const float* __restrict__ pred,
                                                                 const float* __restrict__ label, int h, int w,
                                                                 double* __restrict__ out) {
    __shared__ float tp[STH + 2][STW + 2];
    __shared__ float tl[STH + 2][STW + 2];
    __shared__ double red[8][NSUM];
    const int n = blockIdx.z, y0 = blockIdx.y * STH, x0 = blockIdx.x * STW;
    load_tile(tp, pred + (long long)n * h * w, h, w, y0, x0);
    load_tile(tl, label + (long long)n * h * w, h, w, y0, x0);
    __syncthreads();
    double s[NSUM];
#pragma unroll
    for (int k = 0; k < NSUM; ++k) s[k] = 0.0;
    for (int i = threadIdx.x; i < STH * STW; i += blockDim.x) {
        const int ty = i / STW, tx = i - ty * STW;
        if (y0 + ty >= h || x0 + tx >= w) continue;
        const float o = tp[ty + 1][tx + 1], d = tl[ty + 1][tx + 1];
        float ox, oy, dx, dy;
        sobel_at(tp, ty + 1, tx + 1, ox, oy);
        sobel_at(tl, ty + 1, tx + 1, dx, dy);
        // loss terms, ReSIDE/train.py:231-234
        s[0] += (double)logf(fabsf(o - d) + 0.5f);
        const float dot = ox * dx + oy * dy + 1.f;
        const float no = sqrtf(ox * ox + oy * oy + 1.f), nd = sqrtf(dx * dx + dy * dy + 1.f);
        s[1] += (double)fabsf(1.f - dot / (no * nd));
        s[2] += (double)logf(fabsf(ox - dx) + 0.5f);
        s[3] += (double)logf(fabsf(oy - dy) + 0.5f);
        // metric sums, ReSIDE/util.py:46-73
        const bool isnan_l = d != d;
        const bool invalid = !(d > 0.f);
        const float r = o - d;
        s[4] += (double)fabsf(r);
        s[5] += (double)r * (double)r;
        if (!isnan_l && !invalid) {
            s[6] += (double)(fabsf(r) / d);
            s[7] += (double)fabsf(log10f(o) - log10f(d));
        }
        const float q1 = o / d, q2 = d / o;
        if (q1 == q1 && q2 == q2) {                    // torch.max propagates NaN -> comparison false
            const float ratio = fmaxf(q1, q2);
            s[8] += ratio <= 1.25f ? 1.0 : 0.0;
            s[9] += ratio <= 1.5625f ? 1.0 : 0.0;
            s[10] += ratio <= 1.953125f ? 1.0 : 0.0;
        }
        s[11] += isnan_l ? 0.0 : 1.0;
        s[12] += 1.0;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < NSUM; ++k) {
        const double v = warp_sum(s[k]);
        if (lane == 0) red[warp][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < NSUM) {
        double v = 0.0;
        for (int wi = 0; wi < 8; ++wi) v += red[wi][threadIdx.x];
        atomicAdd(out + threadIdx.x, v);
    }
    if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) atomicAdd(out + 13, 1.0);
}

__global__ void __launch_bounds__(256) edge_counts_kernel(const float* __restrict__ pred,
                                                          const float* __restrict__ label, int h, int w, float thr,
                                                          double* __restrict__ out) {
    __shared__ float tp[STH + 2][STW + 2];
    __shared__ float tl[STH + 2][STW + 2];
    __shared__ double red[8][4];
    const int n = blockIdx.z, y0 = blockIdx.y * STH, x0 = blockIdx.x * STW;
    load_tile(tp, pred + (long long)n * h * w, h, w, y0, x0);
    load_tile(tl, label + (long long)n * h * w, h, w, y0, x0);
    __syncthreads();
    double s[4] = {0.0, 0.0, 0.0, 0.0};
    for (int i = threadIdx.x; i < STH * STW; i += blockDim.x) {
        const int ty = i / STW, tx = i - ty * STW;
        if (y0 + ty >= h || x0 + tx >= w) continue;
        float ox, oy, dx, dy;
        sobel_at(tp, ty + 1, tx + 1, ox, oy);
        sobel_at(tl, ty + 1, tx + 1, dx, dy);
        const bool ep = sqrtf(ox * ox + oy * oy) > thr, el = sqrtf(dx * dx + dy * dy) > thr;
        s[0] += ep; s[1] += el; s[2] += (ep && el); s[3] += (ep == el);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const double v = warp_sum(s[k]);
        if (lane == 0) red[warp][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        double v = 0.0;
        for (int wi = 0; wi < 8; ++wi) v += red[wi][threadIdx.x];
        atomicAdd(out + threadIdx.x, v);
    }
}

}  // namespace ede

using namespace ede;

extern "C" {

int ede_sobel(const float* x, float* y, int32_t n, int32_t h, int32_t w, void* stream) {
    EDE_REQUIRE(x && y && n > 0 && h > 0 && w > 0, "ede_sobel: bad arguments");
    dim3 grid(ceil_div(w, STW), ceil_div(h, STH), n);
    sobel_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, y, h, w);
    EDE_LAUNCH_CHECK("sobel_kernel");
    return EDE_OK;
}

int ede_sobel_loss_metrics(const float* pred, const float* label, int32_t n, int32_t h, int32_t w, double* out16,
                           void* stream) {
    EDE_REQUIRE(pred && label && out16 && n > 0 && h > 0 && w > 0, "ede_sobel_loss_metrics: bad arguments");
    dim3 grid(ceil_div(w, STW), ceil_div(h, STH), n);
    sobel_loss_metrics_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(pred, label, h, w, out16);
    EDE_LAUNCH_CHECK("sobel_loss_metrics_kernel");
    return EDE_OK;
}

int ede_edge_counts(const float* pred, const float* label, int32_t n, int32_t h, int32_t w, float thr, double* out4,
                    void* stream) {
    EDE_REQUIRE(pred && label && out4 && n > 0 && h > 0 && w > 0, "ede_edge_counts: bad arguments");
    dim3 grid(ceil_div(w, STW), ceil_div(h, STH), n);
    edge_counts_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(pred, label, h, w, thr, out4);
    EDE_LAUNCH_CHECK("edge_counts_kernel");
    return EDE_OK;
}

}  // extern "C"
