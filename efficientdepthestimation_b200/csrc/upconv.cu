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
#include <algorithm>
#include <mutex>

#include "common.cuh"

namespace ede {

template <int K, typename TY>
__global__ void __launch_bounds__(256) upconv_taps_kernel(const EdeUpConv a, float scale_h, float scale_w) {
    const int ngrp = a.cgrp >> 3;
    const long long total = (long long)a.n * a.oh * a.ow * ngrp;
    const float* __restrict__ z = reinterpret_cast<const float*>(a.z);
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

// ---- separable form (used when the caller provides the `rows` workspace): bilinear interpolation is a lerp
// along x followed by a lerp along y, and the zero-padding mask of a tap is a product of an x and a y mask, so
//   rows[n][y][ty][px][c] = sum_tx inside_x(px+tx-pl) * lerp_x(z[n][y][.][(ty,tx),c], px+tx-pl)      (pass 1)
//   y[n][py][px][c]       = act(scale * sum_ty inside_y(py+ty-pt) * lerp_y(rows[n][.][ty][px][c], py+ty-pt) + shift)
// Pass 1 runs on the h low-resolution rows only; pass 2 reads 2*K vectors per output instead of 4*K*K.
// grid = (n * h * K low-resolution (row, tap row) pairs, column blocks): no 64-bit index arithmetic per thread
template <int K, typename TZ>
__global__ void __launch_bounds__(256) upconv_rows_kernel(const EdeUpConv a, float scale_w) {
    const int ngrp = a.cgrp >> 3;
    pdl_trigger();
    pdl_wait();
    const int t = blockIdx.y * 256 + threadIdx.x;            // (output column, 8-lane group)
    if (t >= a.ow * ngrp) return;
    const int px = t / ngrp, c = (t - px * ngrp) << 3;
    const int r = blockIdx.x;                                // (n * h + y) * K + ty
    const int ny = r / K, ty = r - ny * K;
    const TZ* zrow = reinterpret_cast<const TZ*>(a.z) + (long long)ny * a.w * a.z_cs + a.z_co + ty * K * a.cgrp + c;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
    for (int tx = 0; tx < K; ++tx) {
        const int qx = px + tx - a.pl;
        if (qx < 0 || qx >= a.ow) continue;                  // tap column in the zero padding
        const float sx = scale_w * qx;
        const int x0 = (int)sx, x1 = x0 + (x0 < a.w - 1 ? 1 : 0);
        const float lx = sx - x0, hx = 1.f - lx;
        float v0[8], v1[8];
        load8<TZ>(zrow + x0 * a.z_cs + tx * a.cgrp, v0);
        load8<TZ>(zrow + x1 * a.z_cs + tx * a.cgrp, v1);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(hx, v0[j], fmaf(lx, v1[j], acc[j]));
    }
    float* dst = a.rows + ((long long)r * a.ow * ngrp + t) * 8;     // [n*h + y][ty][px][cgrp]
    *reinterpret_cast<float4*>(dst) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
}

// Pass 2.  A thread finishes R vertically adjacent outputs of one (column, 8-lane group): for one tap row ty their
// 2R lerp operands are only three distinct `rows` vectors (the map is upsampled >= 3x for R = 4, >= 1x for R = 2),
// loaded once and combined with per-output weights (zero where a vector does not take part).  The weights depend on
// the band of output rows only: a block (= one band of one image) computes them once into shared memory.
// grid = (n * bands, column blocks).
template <int K, typename TY, int R>
__global__ void __launch_bounds__(256) upconv_cols_kernel(const EdeUpConv a, float scale_h) {
    __shared__ float w_s[K][R][3];
    __shared__ int yb_s[K];                                  // first low-resolution row of tap row ty, -1 = all padding
    const int ngrp = a.cgrp >> 3;
    const int bands = (a.oh + R - 1) / R;
    const int n = blockIdx.x / bands, oy0 = (blockIdx.x - n * bands) * R;
    pdl_trigger();
    if (threadIdx.x < K * R) {
        const int ty = threadIdx.x / R, r = threadIdx.x - ty * R;
        const int q0 = oy0 + ty - a.pt;
        const int qlo = max(q0, 0), qhi = min(q0 + R - 1, a.oh - 1);
        const int yb = qlo > qhi ? -1 : (int)(scale_h * qlo);
        if (r == 0) yb_s[ty] = yb;
        const int q = q0 + r;
        float wa = 0.f, wb = 0.f, wc = 0.f;
        if (yb >= 0 && q >= 0 && q < a.oh) {                 // else: tap row in the zero padding
            const float sy = scale_h * q;
            const int y0 = (int)sy, y1 = y0 + (y0 < a.h - 1 ? 1 : 0);
            const float ly = sy - y0, hy = 1.f - ly;
            const int d0 = y0 - yb, d1 = y1 - yb;            // 0..2 by the choice of R
            wa = (d0 == 0 ? hy : 0.f) + (d1 == 0 ? ly : 0.f);
            wb = (d0 == 1 ? hy : 0.f) + (d1 == 1 ? ly : 0.f);
            wc = (d0 == 2 ? hy : 0.f) + (d1 == 2 ? ly : 0.f);
        }
        w_s[ty][r][0] = wa; w_s[ty][r][1] = wb; w_s[ty][r][2] = wc;
    }
    __syncthreads();
    pdl_wait();
    const int t = blockIdx.y * 256 + threadIdx.x;            // (output column, 8-lane group)
    if (t >= a.ow * ngrp) return;
    const int ox = t / ngrp, c = (t - ox * ngrp) << 3;
    const long long ty_stride = (long long)a.ow * a.cgrp, y_stride = ty_stride * K;
    const float* base = a.rows + (long long)n * a.h * y_stride + (long long)t * 8;
    float acc[R][8];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[r][j] = 0.f;
#pragma unroll
    for (int ty = 0; ty < K; ++ty) {
        const int yb = yb_s[ty];
        if (yb < 0) continue;                                // the whole band's tap row lies in the zero padding
        float va[8], vb[8], vc[8];
        const float* p = base + ty * ty_stride;
        load8<float>(p + yb * y_stride, va);
        load8<float>(p + min(yb + 1, a.h - 1) * y_stride, vb);
        load8<float>(p + min(yb + 2, a.h - 1) * y_stride, vc);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float wa = w_s[ty][r][0], wb = w_s[ty][r][1], wc = w_s[ty][r][2];
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[r][j] = fmaf(wa, va[j], fmaf(wb, vb[j], fmaf(wc, vc[j], acc[r][j])));
        }
    }
    float sc[8], sf[8];
    load8<float>(a.scale + c, sc);
    load8<float>(a.shift + c, sf);
    TY* yout = reinterpret_cast<TY*>(a.y) + (((long long)n * a.oh + oy0) * a.ow + ox) * a.y_cs + a.y_co + c;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (oy0 + r >= a.oh) break;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            acc[r][j] = fmaf(acc[r][j], sc[j], sf[j]);
            if (c < a.act_c_end) acc[r][j] = apply_act(acc[r][j], a.act);
        }
        store8<TY>(yout + (long long)r * a.ow * a.y_cs, acc[r]);
    }
}


// ---- one-pass streaming form (no workspace): the two lerps fused.  A thread owns one (output column, 8-lane group)
// and walks DOWN a band of output rows.  For every tap row ty it keeps the two x-lerped low-resolution rows that
// bracket the current sample position in registers,
//   A[ty] = sum_tx inside_x * lerp_x(z[n][y0][.][(ty,tx),c]),   B[ty] = the same for row y0+1,
// so an output row costs K y-lerps (2 FMAs per lane each).  The map is upsampled (scale_h < 1), so from one output row
// to the next y0 either stays or moves down by ONE: then B becomes A and only one new row is x-lerped.  All conditions
// depend on (output row, tap row) only -- uniform over the block.  z is read once from DRAM (neighbouring bands re-read
// their 2*K boundary rows through L2); the `rows` workspace of the separable form (1.2x the bytes of z for 28x38 ->
// 114x152) is neither written nor read.
template <int V> __device__ __forceinline__ void ldz(const float* p, float (&v)[V]) {
    if constexpr (V == 8) {
        uint32_t u[8];
        ldg256(p, u);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(u[j]);
    } else {
        const float4 f = *reinterpret_cast<const float4*>(p);
        v[0] = f.x; v[1] = f.y; v[2] = f.z; v[3] = f.w;
    }
}
template <int K, int V>
__device__ __forceinline__ void upconv_xlerp(const float* __restrict__ zrow, const int (&off0)[K], const int (&off1)[K],
                                             const float (&hx)[K], const float (&lx)[K], float (&out)[V]) {
    float u0[K][V], u1[K][V];
#pragma unroll
    for (int tx = 0; tx < K; ++tx) {                          // 2K independent loads in flight
        ldz<V>(zrow + off0[tx], u0[tx]);
        ldz<V>(zrow + off1[tx], u1[tx]);
    }
#pragma unroll
    for (int j = 0; j < V; ++j) out[j] = 0.f;
#pragma unroll
    for (int tx = 0; tx < K; ++tx)
#pragma unroll
        for (int j = 0; j < V; ++j) out[j] = fmaf(hx[tx], u0[tx][j], fmaf(lx[tx], u1[tx][j], out[j]));
}
template <int K>
__device__ __forceinline__ void upconv_prefetch(const float* zrow, const int (&off0)[K], const int (&off1)[K]) {
#pragma unroll
    for (int tx = 0; tx < K; ++tx) {
        asm volatile("prefetch.global.L1 [%0];" ::"l"(zrow + off0[tx]));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(zrow + off1[tx]));
    }
}

template <int K, typename TY, int V, int THREADS, int MINB, bool PF>
__global__ void __launch_bounds__(THREADS, MINB) upconv_stream_kernel(const EdeUpConv a, float scale_h, float scale_w, int band) {
    const int ngrp = a.cgrp / V;
    const int bands = (a.oh + band - 1) / band;
    const int n = blockIdx.y / bands, oy0 = (blockIdx.y - n * bands) * band;
    const int oy1 = min(oy0 + band, a.oh);
    pdl_trigger();
    const int t = blockIdx.x * THREADS + threadIdx.x;        // (output column, V-lane group)
    if (t >= a.ow * ngrp) return;
    const int px = t / ngrp, c = (t - px * ngrp) * V;
    int off0[K], off1[K];
    float hx[K], lx[K];
#pragma unroll
    for (int tx = 0; tx < K; ++tx) {
        const int qx = px + tx - a.pl;
        const bool vx = qx >= 0 && qx < a.ow;                // else: tap column in the zero padding
        const float sx = scale_w * (vx ? qx : 0);
        const int x0 = (int)sx, x1 = x0 + (x0 < a.w - 1 ? 1 : 0);
        const float l = sx - x0;
        hx[tx] = vx ? 1.f - l : 0.f;
        lx[tx] = vx ? l : 0.f;
        off0[tx] = x0 * a.z_cs + tx * a.cgrp;
        off1[tx] = x1 * a.z_cs + tx * a.cgrp;
    }
    float sc[V], sf[V];
#pragma unroll
    for (int j = 0; j < V; ++j) { sc[j] = a.scale[c + j]; sf[j] = a.shift[c + j]; }
    const bool activated = c < a.act_c_end;
    const float* zimg = reinterpret_cast<const float*>(a.z) + (long long)n * a.h * a.w * a.z_cs + a.z_co + c;
    const int zrow_stride = a.w * a.z_cs;
    // row base of the output pixel (lane 0 of the view) -- store4 wants the row and the lane separately
    TY* yout = reinterpret_cast<TY*>(a.y) + (((long long)n * a.oh + oy0) * a.ow + px) * a.y_cs + a.y_co;
    float A[K][V], B[K][V];
    int ya[K];
#pragma unroll
    for (int ty = 0; ty < K; ++ty) ya[ty] = -2;
    pdl_wait();
    for (int oy = oy0; oy < oy1; ++oy) {
        float acc[V];
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] = 0.f;
#pragma unroll
        for (int ty = 0; ty < K; ++ty) {
            const int q = oy + ty - a.pt;
            if (q < 0 || q >= a.oh) continue;                // tap row in the zero padding
            const float sy = scale_h * q;
            const int y0 = (int)sy;
            const float ly = sy - y0, hy = 1.f - ly;
            if (ya[ty] != y0) {
                const float* zr = zimg + (long long)y0 * zrow_stride + ty * K * a.cgrp;
                if (ya[ty] == y0 - 1) {
#pragma unroll
                    for (int j = 0; j < V; ++j) A[ty][j] = B[ty][j];
                } else {
                    upconv_xlerp<K, V>(zr, off0, off1, hx, lx, A[ty]);
                }
                ya[ty] = y0;
                if (y0 < a.h - 1) {
                    upconv_xlerp<K, V>(zr + zrow_stride, off0, off1, hx, lx, B[ty]);
                    if (PF && y0 < a.h - 2) upconv_prefetch<K>(zr + 2 * zrow_stride, off0, off1);
                } else {
#pragma unroll
                    for (int j = 0; j < V; ++j) B[ty][j] = A[ty][j];
                }
            }
#pragma unroll
            for (int j = 0; j < V; ++j) acc[j] = fmaf(hy, A[ty][j], fmaf(ly, B[ty][j], acc[j]));
        }
#pragma unroll
        for (int j = 0; j < V; ++j) {
            acc[j] = fmaf(acc[j], sc[j], sf[j]);
            if (activated) acc[j] = apply_act(acc[j], a.act);
        }
        if constexpr (V == 8) store8<TY>(yout + c, acc);
        else store4<TY>(yout, c, acc);
        yout += (long long)a.ow * a.y_cs;
    }
}

}  // namespace ede

namespace ede {
// ------------------------------------------------------------------ k x k conv with ONE output channel as GEMM + tap sum
// y[n,oy,ox] = act(scale * sum_{ty,tx} z[n, oy+ty-pt, ox+tx-pl, ty*kw+tx] + shift)   (taps outside the map contribute 0)
// z = (W_tap . x) for every tap, produced by ede_conv2d as a 1x1 GEMM with kh*kw output lanes: a 5x5 conv to one channel
// (R.conv2 of the Hu decoder, ReSIDE/models/modules.py:296) would fill one of 128 tensor-core rows; as a GEMM it fills 25.
// A block owns 8 x 32 output pixels: the (8+kh-1) x (32+kw-1) pixel vectors of z it needs are staged once in shared
// memory (coalesced 16-byte loads; the kh*kw tap lanes of a pixel are packed at an ODD pitch so that neighbouring
// pixels' reads of one lane hit different banks), then every thread adds the kh*kw taps of its output pixel.
// 43 KB of shared memory per block -> 5 blocks per SM hide the load phase of one block behind the sums of the others.
constexpr int TS_TH = 8, TS_TW = 32;
template <typename TY>
__global__ void __launch_bounds__(256) tapsum_kernel(const EdeTapSum a) {
    extern __shared__ float zt[];                               // [(TS_TH+kh-1) * (TS_TW+kw-1)][pitch]
    pdl_trigger();
    pdl_wait();
    const int n = blockIdx.z, oy0 = blockIdx.y * TS_TH, ox0 = blockIdx.x * TS_TW;
    const int th = TS_TH + a.kh - 1, tw = TS_TW + a.kw - 1;
    const int taps = a.kh * a.kw, pitch = taps | 1;
    const int quads = (taps + 3) >> 2;                          // float4 groups of one pixel vector that hold taps
    const float* zn = a.z + (long long)n * a.h * a.w * a.z_cs + a.z_co;
    // 8 lanes per pixel (one float4 each), 32 pixels per pass; (ty, tx) advance without divisions
    const int q = threadIdx.x & 7;
    int p = threadIdx.x >> 3, ty = p / tw, tx = p - ty * tw;
    for (; p < th * tw; p += 32) {
        if (q < quads) {
            const int iy = oy0 + ty - a.pt, ix = ox0 + tx - a.pl;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);         // outside the map: the conv's zero padding
            if (iy >= 0 && iy < a.h && ix >= 0 && ix < a.w)
                v = __ldg(reinterpret_cast<const float4*>(zn + ((long long)iy * a.w + ix) * a.z_cs) + q);
            float* d = zt + p * pitch + 4 * q;
            const int left = taps - 4 * q;
            d[0] = v.x;
            if (left > 1) d[1] = v.y;
            if (left > 2) d[2] = v.z;
            if (left > 3) d[3] = v.w;
        }
        tx += 32;
        while (tx >= tw) { tx -= tw; ++ty; }
    }
    __syncthreads();
    const int lx = threadIdx.x & 31, ly = threadIdx.x >> 5;
    const int oy = oy0 + ly, ox = ox0 + lx;
    if (oy >= a.h || ox >= a.w) return;
    float acc = 0.f;
    for (int ky = 0; ky < a.kh; ++ky) {
        const float* row = zt + ((ly + ky) * tw + lx) * pitch + ky * a.kw;
        for (int kx = 0; kx < a.kw; ++kx) acc += row[kx * pitch + kx];
    }
    float v[8] = {apply_act(fmaf(acc, a.scale[0], a.shift[0]), a.act), 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    TY* y = reinterpret_cast<TY*>(a.y) + (((long long)n * a.h + oy) * a.w + ox) * a.y_cs + a.y_co;
    store8<TY>(y, v);
}
}  // namespace ede

using namespace ede;

extern "C" int ede_upconv_taps(const EdeUpConv* a, void* stream) {
    EDE_REQUIRE(a && a->z && a->y && a->scale && a->shift, "ede_upconv_taps: null argument");
    EDE_REQUIRE(((a->cgrp | a->z_cs | a->z_co | a->y_cs | a->y_co | a->act_c_end) & 7) == 0,
                "ede_upconv_taps: channel counts must be multiples of 8");
    EDE_REQUIRE(a->n > 0 && a->h > 0 && a->w > 0 && a->oh > 0 && a->ow > 0 && a->cgrp > 0, "ede_upconv_taps: bad shape");
    EDE_REQUIRE(a->kh == a->kw && (a->kh == 3 || a->kh == 5), "ede_upconv_taps: only 3x3 and 5x5 taps are supported");
    EDE_REQUIRE(a->z_dtype == EDE_F32 || (a->z_dtype == EDE_BF16 && a->rows != nullptr),
                "ede_upconv_taps: z must be fp32, or bf16 with the rows workspace (separable form)");
    const float sh = a->oh > 1 ? (float)(a->h - 1) / (float)(a->oh - 1) : 0.f;
    const float sw = a->ow > 1 ? (float)(a->w - 1) / (float)(a->ow - 1) : 0.f;
    const long long total = (long long)a->n * a->oh * a->ow * (a->cgrp >> 3);
    long long g = (total + 255) / 256;
    if (g > 148LL * 64) g = 148LL * 64;
    cudaStream_t st = (cudaStream_t)stream;
    if (a->rows) {                                           // separable two-pass form
        EDE_REQUIRE(sh < 1.f, "ede_upconv_taps: the separable form needs an upsampled map (h < oh)");
        const unsigned col_blocks = (unsigned)((a->ow * (a->cgrp >> 3) + 255) / 256);
        const long long r1 = (long long)a->n * a->h * a->kh;
        EDE_REQUIRE(r1 < 0x7fffffffLL && col_blocks <= 65535u, "ede_upconv_taps: map too large");
        const dim3 g1((unsigned)r1, col_blocks);
        if (a->z_dtype == EDE_BF16) {
            if (a->kh == 5) (void)launch_pdl(upconv_rows_kernel<5, __nv_bfloat16>, g1, dim3(256), 0, st, *a, sw);
            else (void)launch_pdl(upconv_rows_kernel<3, __nv_bfloat16>, g1, dim3(256), 0, st, *a, sw);
        } else {
            if (a->kh == 5) (void)launch_pdl(upconv_rows_kernel<5, float>, g1, dim3(256), 0, st, *a, sw);
            else (void)launch_pdl(upconv_rows_kernel<3, float>, g1, dim3(256), 0, st, *a, sw);
        }
        EDE_LAUNCH_CHECK("upconv_rows_kernel");
        // output rows per thread: 4 when the map is upsampled >= 3x (three low-res rows cover a band), else 2
        const int R = sh <= 1.f / 3.f ? 4 : 2;
        const dim3 g2((unsigned)(a->n * ((a->oh + R - 1) / R)), col_blocks);
        if (a->kh == 5) {
            if (R == 4) { EDE_DISPATCH_DTYPE(a->y_dtype, TY, ((void)launch_pdl(upconv_cols_kernel<5, TY, 4>, g2, dim3(256), 0, st, *a, sh))); }
            else { EDE_DISPATCH_DTYPE(a->y_dtype, TY, ((void)launch_pdl(upconv_cols_kernel<5, TY, 2>, g2, dim3(256), 0, st, *a, sh))); }
        } else {
            if (R == 4) { EDE_DISPATCH_DTYPE(a->y_dtype, TY, ((void)launch_pdl(upconv_cols_kernel<3, TY, 4>, g2, dim3(256), 0, st, *a, sh))); }
            else { EDE_DISPATCH_DTYPE(a->y_dtype, TY, ((void)launch_pdl(upconv_cols_kernel<3, TY, 2>, g2, dim3(256), 0, st, *a, sh))); }
        }
        EDE_LAUNCH_CHECK("upconv_cols_kernel");
        return EDE_OK;
    }
    static const char* stream_knob = getenv("EDE_UPCONV_STREAM");         // 0: keep the 4-corner gather for rows == NULL
    if (sh < 1.f && !(stream_knob && stream_knob[0] == '0')) {              // upsampled map: one-pass streaming form
        static const char* band_knob = getenv("EDE_UPCONV_BAND");
        const unsigned col_blocks = (unsigned)((a->ow * (a->cgrp >> 3) + 255) / 256);
        // rows per band: long bands amortise the 2*K rows x-lerped at a band's start, short ones fill the machine;
        // aim at >= 6 waves of 148 SMs (one 256-thread block each: 240 registers per thread), never below 16 rows
        int band = band_knob ? atoi(band_knob) : 0;
        if (band <= 0) {
            const long long want_blocks = 148LL * 6;
            long long bands = (want_blocks + (long long)a->n * col_blocks - 1) / ((long long)a->n * col_blocks);
            bands = std::max(1LL, std::min(bands, (long long)(a->oh + 15) / 16));
            band = (int)((a->oh + bands - 1) / bands);
        }
        band = std::max(1, std::min(band, (int)a->oh));
        const long long gy = (long long)a->n * ((a->oh + band - 1) / band);
        EDE_REQUIRE(gy <= 65535 && col_blocks < 0x7fffffffu, "ede_upconv_taps: map too large");
        const dim3 gs(col_blocks, (unsigned)gy);
        if (a->kh == 5) { EDE_DISPATCH_DTYPE(a->y_dtype, TY, ((void)launch_pdl(upconv_stream_kernel<5, TY, 8, 256, 1, true>, gs, dim3(256), 0, st, *a, sh, sw, band))); }
        else { EDE_DISPATCH_DTYPE(a->y_dtype, TY, ((void)launch_pdl(upconv_stream_kernel<3, TY, 8, 256, 2, false>, gs, dim3(256), 0, st, *a, sh, sw, band))); }
        EDE_LAUNCH_CHECK("upconv_stream_kernel");
        return EDE_OK;
    }
    if (a->kh == 5) {
        EDE_DISPATCH_DTYPE(a->y_dtype, TY, (upconv_taps_kernel<5, TY><<<(unsigned)g, 256, 0, st>>>(*a, sh, sw)));
    } else {
        EDE_DISPATCH_DTYPE(a->y_dtype, TY, (upconv_taps_kernel<3, TY><<<(unsigned)g, 256, 0, st>>>(*a, sh, sw)));
    }
    EDE_LAUNCH_CHECK("upconv_taps_kernel");
    return EDE_OK;
}

// ---- allocation-free variant of the separable gather: the `rows` workspace lives in a file-scope device array.
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  Every GPU that loads the library holds the
// array until exit.  The array's name is only a host-side stand-in in host code: the device address is looked up once
// per device with cudaGetSymbolAddress (that device current) and handed to the existing kernels through EdeUpConv.rows.
namespace ede {
constexpr long long UPCONV_STATIC_ROWS_FLOATS = 4LL << 20;                 // 16 MiB
static __device__ __align__(256) float g_upconv_static_rows[UPCONV_STATIC_ROWS_FLOATS];
}  // namespace ede

extern "C" int64_t ede_upconv_taps_static_max_bytes(void) { return ede::UPCONV_STATIC_ROWS_FLOATS * (int64_t)sizeof(float); }

extern "C" int ede_upconv_taps_static(const EdeUpConv* a, void* stream) {
    EDE_REQUIRE(a, "ede_upconv_taps_static: null argument");
    EDE_REQUIRE(a->n > 0 && a->h > 0 && a->kh > 0 && a->ow > 0 && a->cgrp > 0, "ede_upconv_taps_static: bad shape");
    const long long need = (long long)a->n * a->h * a->kh * a->ow * a->cgrp;   // floats of rows[n][h][kh][ow][cgrp]
    EDE_REQUIRE(need <= ede::UPCONV_STATIC_ROWS_FLOATS,
                "ede_upconv_taps_static: the rows workspace of this shape needs %lld bytes, the static scratch holds %lld "
                "(use ede_upconv_taps with a caller-provided workspace)",
                need * (long long)sizeof(float), ede::UPCONV_STATIC_ROWS_FLOATS * (long long)sizeof(float));
    static float* dev_ptr[64] = {};                                        // per device, looked up once
    static std::mutex mu;
    int dev = 0;
    int rc = check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
    if (rc) return rc;
    EDE_REQUIRE(dev >= 0 && dev < 64, "ede_upconv_taps_static: device ordinal %d out of range", dev);
    float* rows = nullptr;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!dev_ptr[dev]) {
            void* p = nullptr;
            rc = check_cuda(cudaGetSymbolAddress(&p, ede::g_upconv_static_rows), "cudaGetSymbolAddress(g_upconv_static_rows)");
            if (rc) return rc;
            EDE_REQUIRE(p != nullptr, "ede_upconv_taps_static: cudaGetSymbolAddress returned a null address");
            dev_ptr[dev] = static_cast<float*>(p);
        }
        rows = dev_ptr[dev];
    }
    // upconv_rows_kernel writes every element of rows[n][h][kh][ow][cgrp] before upconv_cols_kernel reads it (stream
    // order), so nothing of the previous call's contents has to be cleared.
    EdeUpConv b = *a;
    b.rows = rows;
    return ede_upconv_taps(&b, stream);
}

extern "C" int ede_tapsum(const EdeTapSum* a, void* stream) {
    EDE_REQUIRE(a && a->z && a->y && a->scale && a->shift, "ede_tapsum: null argument");
    EDE_REQUIRE(((a->z_cs | a->z_co | a->y_cs | a->y_co) & 7) == 0, "ede_tapsum: channel strides must be multiples of 8");
    EDE_REQUIRE(a->n > 0 && a->n <= 65535 && a->h > 0 && a->w > 0 && a->kh > 0 && a->kw > 0 && a->kh * a->kw <= a->z_cs,
                "ede_tapsum: bad shape");
    EDE_REQUIRE(a->kh * a->kw <= 32, "ede_tapsum: more than 32 taps (a block stages 8 float4 lanes per pixel)");
    const dim3 grid((unsigned)ceil_div(a->w, ede::TS_TW), (unsigned)ceil_div(a->h, ede::TS_TH), (unsigned)a->n);
    const size_t smem = (size_t)(ede::TS_TH + a->kh - 1) * (ede::TS_TW + a->kw - 1) * ((a->kh * a->kw) | 1) * sizeof(float);
    EDE_DISPATCH_DTYPE(a->y_dtype, TY, {
        static bool attr_set[64] = {};
        int dev = 0;
        (void)cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64) dev = 0;
        if (!attr_set[dev]) {
            int rc = check_cuda(cudaFuncSetAttribute(tapsum_kernel<TY>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024),
                                "cudaFuncSetAttribute(tapsum_kernel)");
            if (rc) return rc;
            attr_set[dev] = true;
        }
        (void)launch_pdl(tapsum_kernel<TY>, grid, dim3(256), smem, (cudaStream_t)stream, *a);
    });
    EDE_LAUNCH_CHECK("tapsum_kernel");
    return EDE_OK;
}
