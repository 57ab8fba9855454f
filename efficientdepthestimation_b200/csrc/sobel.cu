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

// per-pixel terms of the loss (ReSIDE/train.py:231-234) and the metric sums (ReSIDE/util.py:46-73)
// Accumulated in fp32 over the (at most four) pixels a thread owns in one tile, then added to the thread's fp64 sums:
// the reference sums everything in fp32, so this is still the more accurate side, and it keeps the per-pixel work
// free of fp32->fp64 conversions (quarter-rate) and fp64 adds.
__device__ __forceinline__ void accumulate_loss_metrics(float* s, float o, float d, float ox, float oy, float dx,
                                                        float dy) {
    // loss terms, ReSIDE/train.py:231-234
    s[0] += logf(fabsf(o - d) + 0.5f);
    const float dot = ox * dx + oy * dy + 1.f;
    const float no = sqrtf(ox * ox + oy * oy + 1.f), nd = sqrtf(dx * dx + dy * dy + 1.f);
    s[1] += fabsf(1.f - dot / (no * nd));
    s[2] += logf(fabsf(ox - dx) + 0.5f);
    s[3] += logf(fabsf(oy - dy) + 0.5f);
    // metric sums, ReSIDE/util.py:46-73
    const bool isnan_l = d != d;
    const bool invalid = !(d > 0.f);
    const float r = o - d;
    s[4] += fabsf(r);
    s[5] += r * r;
    if (!isnan_l && !invalid) {
        s[6] += fabsf(r) / d;
        s[7] += fabsf(log10f(o) - log10f(d));
    }
    const float q1 = o / d, q2 = d / o;
    if (q1 == q1 && q2 == q2) {                    // torch.max propagates NaN -> comparison false
        const float ratio = fmaxf(q1, q2);
        s[8] += ratio <= 1.25f ? 1.f : 0.f;
        s[9] += ratio <= 1.5625f ? 1.f : 0.f;
        s[10] += ratio <= 1.953125f ? 1.f : 0.f;
    }
    s[11] += isnan_l ? 0.f : 1.f;
    s[12] += 1.f;
}

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
    float sf[NSUM];
#pragma unroll
    for (int k = 0; k < NSUM; ++k) sf[k] = 0.f;
    for (int i = threadIdx.x; i < STH * STW; i += blockDim.x) {
        const int ty = i / STW, tx = i - ty * STW;
        if (y0 + ty >= h || x0 + tx >= w) continue;
        const float o = tp[ty + 1][tx + 1], d = tl[ty + 1][tx + 1];
        float ox, oy, dx, dy;
        sobel_at(tp, ty + 1, tx + 1, ox, oy);
        sobel_at(tl, ty + 1, tx + 1, dx, dy);
        accumulate_loss_metrics(sf, o, d, ox, oy, dx, dy);
    }
#pragma unroll
    for (int k = 0; k < NSUM; ++k) s[k] += (double)sf[k];
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

// Label-free half of the epilogue (test_nyu.py:85-88, inference_benchmark.py:162): resize and convert, four output
// pixels of one row per thread, nothing staged -- the low-resolution map is L1/L2 resident.
__global__ void __launch_bounds__(256) depth_resize_convert_kernel(const EdeDepthEpilogue a, float scale_h, float scale_w,
                                                                   int quads_per_row) {
    const long long total = (long long)a.n * a.oh * quads_per_row;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const unsigned q = (unsigned)(i % quads_per_row);
        const unsigned row = (unsigned)(i / quads_per_row);              // n * oh + oy
        const unsigned n = row / (unsigned)a.oh, oy = row - n * (unsigned)a.oh;
        const float* __restrict__ p = a.pred + (long long)n * a.h * a.w;
        const int x0 = (int)q * 4;
        float v[4];
        const LerpTap ty = lerp_tap_ac(scale_h, (float)oy, a.h);         // one row per thread: the y taps are shared
        const float fx0 = (float)x0;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            v[j] = x0 + j < a.ow ? bilerp_taps(p, a.w, ty, lerp_tap_ac(scale_w, fx0 + (float)j, a.w)) : 0.f;
        const long long base = (long long)row * a.ow + x0;
        const bool full = x0 + 4 <= a.ow;
        if (a.resized) {
            if (full && ((a.ow & 3) == 0)) *reinterpret_cast<float4*>(a.resized + base) = make_float4(v[0], v[1], v[2], v[3]);
            else
                for (int j = 0; j < 4 && x0 + j < a.ow; ++j) a.resized[base + j] = v[j];
        }
        if (a.depth_mm) {
            unsigned short u[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) u[j] = (unsigned short)(int)__fmul_rn(v[j] > a.cap ? 0.f : v[j], a.mm_scale);
            if (full && ((a.ow & 3) == 0)) *reinterpret_cast<uint2*>(a.depth_mm + base) = make_uint2(u[0] | ((unsigned)u[1] << 16), u[2] | ((unsigned)u[3] << 16));
            else
                for (int j = 0; j < 4 && x0 + j < a.ow; ++j) a.depth_mm[base + j] = u[j];
        }
    }
}

// Fused post-forward epilogue: the low-resolution depth map is resized (bilinear, align_corners) tile by tile into
// shared memory and consumed in place -- metric / loss sums and edge counts against the label, the fp32 resized map
// and the 16-bit millimetre map are all optional outputs of the same pass, so the full-resolution prediction never
// has to exist in HBM when only the sums or the 16-bit map are wanted.
__global__ void __launch_bounds__(256, 3) depth_epilogue_kernel(const EdeDepthEpilogue a, float scale_h, float scale_w,
                                                             int tiles_x, int tiles_y) {
    __shared__ float tp[STH + 2][STW + 2];
    __shared__ float tl[STH + 2][STW + 2];
    __shared__ double red[8][NSUM + 4];
    double s[NSUM + 4];
#pragma unroll
    for (int k = 0; k < NSUM + 4; ++k) s[k] = 0.0;
    // persistent blocks walk the tiles: one set of 17 atomics per BLOCK, not per tile (they all hit the same lines)
    const int per_img = tiles_x * tiles_y, ntiles = per_img * a.n;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int n = tile / per_img, rem = tile - n * per_img;
        const int by = rem / tiles_x, bx = rem - by * tiles_x;
        const int y0 = by * STH, x0 = bx * STW;
        const float* __restrict__ p = a.pred + (long long)n * a.h * a.w;
        for (int i = threadIdx.x; i < (STH + 2) * (STW + 2); i += blockDim.x) {
            const int ty = i / (STW + 2), tx = i - ty * (STW + 2);
            const int y = y0 + ty - 1, x = x0 + tx - 1;
            // zeros outside the map: the Sobel Conv2d zero-pads the RESIZED map (ReSIDE/sobel.py:8)
            tp[ty][tx] = (y >= 0 && y < a.oh && x >= 0 && x < a.ow) ? bilerp_ac_plane(p, a.h, a.w, scale_h, scale_w, y, x) : 0.f;
        }
        if (a.label) load_tile(tl, a.label + (long long)n * a.oh * a.ow, a.oh, a.ow, y0, x0);
        __syncthreads();
        float sf[NSUM + 4];
#pragma unroll
        for (int k = 0; k < NSUM + 4; ++k) sf[k] = 0.f;
        for (int i = threadIdx.x; i < STH * STW; i += blockDim.x) {
            const int ty = i / STW, tx = i - ty * STW;
            const int py = y0 + ty, px = x0 + tx;
            if (py >= a.oh || px >= a.ow) continue;
            const float o = tp[ty + 1][tx + 1];
            const long long idx = ((long long)n * a.oh + py) * a.ow + px;
            if (a.resized) a.resized[idx] = o;
            if (a.depth_mm) {
                // test_nyu.py:87-88: outputs[outputs > 10.0] = 0; (outputs * 1000.0).astype(np.uint16)  (C cast via int32)
                const float v = o > a.cap ? 0.f : o;
                a.depth_mm[idx] = (unsigned short)(int)__fmul_rn(v, a.mm_scale);
            }
            if (a.label) {
                const float d = tl[ty + 1][tx + 1];
                float ox, oy, dx, dy;
                sobel_at(tp, ty + 1, tx + 1, ox, oy);
                sobel_at(tl, ty + 1, tx + 1, dx, dy);
                if (a.sums16) accumulate_loss_metrics(sf, o, d, ox, oy, dx, dy);
                if (a.edges4) {
                    const bool ep = sqrtf(ox * ox + oy * oy) > a.edge_thr, el = sqrtf(dx * dx + dy * dy) > a.edge_thr;
                    sf[NSUM + 0] += ep; sf[NSUM + 1] += el; sf[NSUM + 2] += (ep && el); sf[NSUM + 3] += (ep == el);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < NSUM + 4; ++k) s[k] += (double)sf[k];
        __syncthreads();
    }
    if (!a.label) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < NSUM + 4; ++k) {
        const double v = warp_sum(s[k]);
        if (lane == 0) red[warp][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < NSUM + 4) {
        double v = 0.0;
        for (int wi = 0; wi < 8; ++wi) v += red[wi][threadIdx.x];
        if (threadIdx.x < NSUM) { if (a.sums16) atomicAdd(a.sums16 + threadIdx.x, v); }
        else if (a.edges4) atomicAdd(a.edges4 + (threadIdx.x - NSUM), v);
    }
    if (a.sums16 && threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(a.sums16 + 13, (double)a.n);
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

int ede_depth_epilogue(const EdeDepthEpilogue* a, void* stream) {
    EDE_REQUIRE(a && a->pred, "ede_depth_epilogue: null argument");
    EDE_REQUIRE(a->n > 0 && a->h > 0 && a->w > 0 && a->oh > 0 && a->ow > 0, "ede_depth_epilogue: bad shape");
    EDE_REQUIRE(a->h < (1 << 22) && a->w < (1 << 22) && (long long)a->h * a->w < 0x7fffffffLL, "ede_depth_epilogue: map too large");
    EDE_REQUIRE(a->label || (!a->sums16 && !a->edges4), "ede_depth_epilogue: sums / edge counts need a label");
    EDE_REQUIRE(a->resized || a->depth_mm || a->sums16 || a->edges4, "ede_depth_epilogue: no output requested");
    const float sh = a->oh > 1 ? (float)(a->h - 1) / (float)(a->oh - 1) : 0.f;
    const float sw = a->ow > 1 ? (float)(a->w - 1) / (float)(a->ow - 1) : 0.f;
    if (!a->label) {
        const int quads = ceil_div(a->ow, 4);
        const long long total = (long long)a->n * a->oh * quads;
        EDE_REQUIRE((long long)a->n * a->oh <= 0x7fffffffLL, "ede_depth_epilogue: too many rows");
        depth_resize_convert_kernel<<<(unsigned)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16), 256, 0, (cudaStream_t)stream>>>(*a, sh, sw, quads);
        EDE_LAUNCH_CHECK("depth_resize_convert_kernel");
        return EDE_OK;
    }
    const int tiles_x = ceil_div(a->ow, STW), tiles_y = ceil_div(a->oh, STH);
    const long long ntiles = (long long)tiles_x * tiles_y * a->n;
    EDE_REQUIRE(ntiles <= 0x7fffffffLL, "ede_depth_epilogue: too many tiles");
    static const char* knob = getenv("EDE_EPILOGUE_BLOCKS");              // tuning knob: blocks in the persistent grid
    const long long cap = knob ? atoll(knob) : 148 * 3;
    const int grid = (int)(ntiles < cap ? ntiles : cap);
    depth_epilogue_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(*a, sh, sw, tiles_x, tiles_y);
    EDE_LAUNCH_CHECK("depth_epilogue_kernel");
    return EDE_OK;
}

}  // extern "C"
