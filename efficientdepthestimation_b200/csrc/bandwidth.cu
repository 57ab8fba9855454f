// bandwidth.cu -- the HBM-bound kernels of the ReSIDE forward path (NHWC, 16-byte vectors).
//
// Every thread owns one 8-channel vector (16 B of bf16 / 32 B of fp32) of one pixel, with the
// channel-group index fastest across a warp so that global accesses are fully coalesced.
#include <stdlib.h>

#include <cooperative_groups.h>

#include "common.cuh"

namespace ede {

static inline unsigned grid_for(long long work, int block) {
    long long g = (work + block - 1) / block;
    if (g < 1) g = 1;
    if (g > 0x7fffffffLL) g = 0x7fffffffLL;
    return (unsigned)g;
}
// grid-stride kernels: a few resident blocks per SM that each walk many items (independent loads of consecutive
// iterations overlap) instead of one short-lived block per 256 items
static inline unsigned grid_persistent(long long work, int block, int blocks_per_sm = 8) {
    static const int cap_env = [] { const char* e = getenv("EDE_GRID_CAP"); return e && *e ? atoi(e) : 0; }();
    const long long cap = 148LL * (cap_env > 0 ? cap_env : blocks_per_sm);
    const long long g = grid_for(work, block);
    return (unsigned)(g < cap ? g : cap);
}

// ------------------------------------------------------------------ depthwise conv + BN + act + SE squeeze
template <typename T>
__global__ void __launch_bounds__(256) dwconv_kernel(const EdeDwConv a) {
    extern __shared__ float pool_s[];                     // [c_p] per-block partial sums (if pool_sum)
    const int cg = a.c_p >> 3;
    const int n = blockIdx.y;
    const long long per_img = (long long)a.oh * a.ow * cg;
    if (a.pool_sum) {
        for (int i = threadIdx.x; i < a.c_p; i += blockDim.x) pool_s[i] = 0.f;
        __syncthreads();
    }
    const T* __restrict__ x = reinterpret_cast<const T*>(a.x);
    T* __restrict__ y = reinterpret_cast<T*>(a.y);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < per_img;
         i += (long long)gridDim.x * blockDim.x) {
        const int g = (int)(i % cg);
        const long long pix = i / cg;
        const int ow = (int)(pix % a.ow), oh = (int)(pix / a.ow);
        const int c = g << 3;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        const int ih0 = oh * a.s - a.pt, iw0 = ow * a.s - a.pl;
        for (int ky = 0; ky < a.k; ++ky) {
            const int ih = ih0 + ky;
            if (ih < 0 || ih >= a.h) continue;
            for (int kx = 0; kx < a.k; ++kx) {
                const int iw = iw0 + kx;
                if (iw < 0 || iw >= a.w_) continue;
                float xv[8], wv[8];
                load8<T>(x + (((long long)n * a.h + ih) * a.w_ + iw) * a.x_cs + a.x_co + c, xv);
                load8<float>(a.w + (long long)(ky * a.k + kx) * a.c_p + c, wv);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[j] = fmaf(xv[j], wv[j], acc[j]);
            }
        }
        float sc[8], sf[8];
        load8<float>(a.scale + c, sc);
        load8<float>(a.shift + c, sf);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = apply_act(fmaf(acc[j], sc[j], sf[j]), a.act);
        store8<T>(y + (((long long)n * a.oh + oh) * a.ow + ow) * a.y_cs + a.y_co + c, acc);
        if (a.pool_sum) {
            // the squeeze must see exactly what the next layer reads: the value rounded to T
            float rv[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) rv[j] = round_as<T>(acc[j]);
#pragma unroll
            for (int j = 0; j < 8; ++j) atomicAdd(&pool_s[c + j], rv[j]);
        }
    }
    if (a.pool_sum) {
        __syncthreads();
        for (int i = threadIdx.x; i < a.c_p; i += blockDim.x) atomicAdd(a.pool_sum + (long long)n * a.c_p + i, pool_s[i]);
    }
}


int dwconv_tma_launch(const EdeDwConv& a, cudaStream_t st, int* strips_out);      // dwconv.cu
int dwconv_tile_launch(const EdeDwConv& a, cudaStream_t st, int* bands_out);      // dwconv_tile.cu

// ------------------------------------------------------------------ squeeze-excite FCs
__global__ void __launch_bounds__(1024) se_fc_kernel(const EdeSeFc a) {
    extern __shared__ float sm[];                          // mean[c_p] | hidden[csq]
    float* mean = sm;
    float* hid = sm + a.c_p;
    const int n = blockIdx.x;
    pdl_trigger();
    // The FC weights are constants of the plan but fall out of the L2 between forwards (a step streams ~7 GB through
    // it): pull them in while the depthwise kernel that feeds this one is still running, before the dependency wait
    {
        const char* w1b = reinterpret_cast<const char*>(a.w1);
        const char* w2b = reinterpret_cast<const char*>(a.w2);
        const int bytes = a.csq * a.c_p * 4;
        for (int o = threadIdx.x * 128; o < bytes; o += blockDim.x * 128) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"(w1b + o));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(w2b + o));
        }
        if (a.w_src) {
            const char* wb = reinterpret_cast<const char*>(a.w_src);
            const int wbytes = a.w_passes * a.w_rows * a.w_cols * 2;
            for (int o = threadIdx.x * 128; o < wbytes; o += blockDim.x * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(wb + o));
        }
    }
    pdl_wait();
    const int parts = a.pool_parts > 1 ? a.pool_parts : 1;
    for (int i = threadIdx.x; i < a.c_p; i += blockDim.x) {
        float sum = 0.f;
        for (int s = 0; s < parts; ++s) {                  // slabs in order: the same rounding every forward
            float* ps = a.pool_sum + ((long long)s * a.n + n) * a.c_p + i;
            sum += *ps;
            if (a.clear_pool) *ps = 0.f;                   // ready for the next forward's squeeze
        }
        mean[i] = i < a.c ? sum * a.inv_hw : 0.f;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int j = warp; j < a.csq; j += nw) {
        float s = 0.f;
        for (int c = lane; c < a.c; c += 32) s = fmaf(a.w1[(long long)j * a.c_p + c], mean[c], s);
        s = warp_sum(s);
        if (lane == 0) hid[j] = apply_act(s + a.b1[j], EDE_ACT_SWISH);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < a.c_p; c += blockDim.x) {
        float s = 0.f;
        if (c < a.c) {
            s = a.b2[c];
            if (a.w2_transposed) {
                for (int j = 0; j < a.csq; ++j) s = fmaf(a.w2[(long long)j * a.c_p + c], hid[j], s);
            } else {
                for (int j = 0; j < a.csq; ++j) s = fmaf(a.w2[(long long)c * a.csq + j], hid[j], s);
            }
        }
        const float g = c < a.c ? apply_act(s, EDE_ACT_SIGMOID) : 0.f;
        a.gate[(long long)n * a.c_p + c] = g;
        mean[c] = g;                                       // kept for the weight fold below (mean is dead by now)
    }
    if (a.w_src) {
        // (x * g) . W == x . (g * W): scale column ci of the consumer's 1x1 weights by this image's gate and re-split
        // into bf16 hi / lo, 8 columns (= 8 consecutive channels, also in the doubled-lane split layout) per step
        __syncthreads();
        const int rows = a.w_rows, cols8 = a.w_cols >> 3;
        const uint4* src = reinterpret_cast<const uint4*>(a.w_src);
        uint4* dst = reinterpret_cast<uint4*>(a.w_dst) + (long long)n * a.w_passes * rows * cols8;
        const long long lo_off = (long long)rows * cols8;
        for (int e = threadIdx.x; e < rows * cols8; e += blockDim.x) {
            const int ci = (e % cols8) << 3;
            const int lane = a.w_split ? ((ci >> 4) << 3) : ci;
            const uint4 h = src[e];
            uint4 l = make_uint4(0, 0, 0, 0);
            if (a.w_passes == 2) l = src[lo_off + e];
            const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, lw[4] = {l.x, l.y, l.z, l.w};
            uint32_t oh[4], ol[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float g0 = lane + 2 * q < a.c_p ? mean[lane + 2 * q] : 0.f;
                const float g1 = lane + 2 * q + 1 < a.c_p ? mean[lane + 2 * q + 1] : 0.f;
                const float v0 = (bf16lo_f(hw[q]) + bf16lo_f(lw[q])) * g0, v1 = (bf16hi_f(hw[q]) + bf16hi_f(lw[q])) * g1;
                const __nv_bfloat162 hh = __floats2bfloat162_rn(v0, v1);
                oh[q] = *reinterpret_cast<const uint32_t*>(&hh);
                const __nv_bfloat162 ll = __floats2bfloat162_rn(v0 - bf16lo_f(oh[q]), v1 - bf16hi_f(oh[q]));
                ol[q] = *reinterpret_cast<const uint32_t*>(&ll);
            }
            dst[e] = make_uint4(oh[0], oh[1], oh[2], oh[3]);
            if (a.w_passes == 2) dst[lo_off + e] = make_uint4(ol[0], ol[1], ol[2], ol[3]);
        }
    }
}

// The same op across a thread-block CLUSTER of B CTAs per image.  One block per image leaves most of the machine idle
// at the batch sizes of the BASELINE configs (8..64 blocks on 148 SMs) while the two FC phases of the wide EfficientNet-B4
// stages are 300k MACs each.  Here every CTA of the cluster reads the squeeze vector, the hidden units are split over the
// CTAs and exchanged through DISTRIBUTED SHARED MEMORY, and the gate channels and the weight fold are split as well;
// cluster.sync() orders "everybody has read the slabs" before rank 0 clears them and publishes the gate slices that the
// fold reads back from global memory.  Same arithmetic per output as se_fc_kernel (same summation orders): bit-identical.
__global__ void __launch_bounds__(1024) se_fc_cluster_kernel(const EdeSeFc a) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const int B = (int)cluster.num_blocks(), rank = (int)cluster.block_rank();
    extern __shared__ float sm[];                          // mean[c_p] | hid[csq] (own units) | hid_all[csq]
    float* mean = sm;
    float* hid = sm + a.c_p;
    float* hid_all = hid + a.csq;
    const int n = blockIdx.x / B;
    pdl_trigger();
    {
        // the FC weights fall out of the L2 between forwards: pull this CTA's share in before the dependency wait
        const char* w1b = reinterpret_cast<const char*>(a.w1);
        const char* w2b = reinterpret_cast<const char*>(a.w2);
        const int bytes = a.csq * a.c_p * 4;
        for (int o = (rank * blockDim.x + threadIdx.x) * 128; o < bytes; o += B * blockDim.x * 128) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"(w1b + o));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(w2b + o));
        }
        if (a.w_src) {
            const char* wb = reinterpret_cast<const char*>(a.w_src);
            const int wbytes = a.w_passes * a.w_rows * a.w_cols * 2;
            for (int o = (rank * blockDim.x + threadIdx.x) * 128; o < wbytes; o += B * blockDim.x * 128)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(wb + o));
        }
    }
    pdl_wait();
    const int parts = a.pool_parts > 1 ? a.pool_parts : 1;
    for (int i = threadIdx.x; i < a.c_p; i += blockDim.x) {
        float sum = 0.f;
        for (int s = 0; s < parts; ++s) sum += a.pool_sum[((long long)s * a.n + n) * a.c_p + i];   // slabs in order
        mean[i] = i < a.c ? sum * a.inv_hw : 0.f;
    }
    cluster.sync();                                        // every CTA of the image has read the slabs
    if (a.clear_pool && rank == 0)
        for (int i = threadIdx.x; i < a.c_p; i += blockDim.x)
            for (int s = 0; s < parts; ++s) a.pool_sum[((long long)s * a.n + n) * a.c_p + i] = 0.f;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int j = rank * nw + warp; j < a.csq; j += B * nw) {           // hidden units of this CTA
        float s = 0.f;
        for (int c = lane; c < a.c; c += 32) s = fmaf(a.w1[(long long)j * a.c_p + c], mean[c], s);
        s = warp_sum(s);
        if (lane == 0) hid[j] = apply_act(s + a.b1[j], EDE_ACT_SWISH);
    }
    cluster.sync();                                        // all hidden units exist (release / acquire at cluster scope)
    for (int j = threadIdx.x; j < a.csq; j += blockDim.x) {
        const int owner = (j % (B * nw)) / nw;
        hid_all[j] = cluster.map_shared_rank(hid, owner)[j];           // distributed shared memory read
    }
    __syncthreads();
    const int cs = ((a.c_p + B - 1) / B + 7) & ~7;         // gate channels per CTA
    const int c_lo = rank * cs, c_hi = min(a.c_p, c_lo + cs);
    for (int c = c_lo + threadIdx.x; c < c_hi; c += blockDim.x) {
        float s = 0.f;
        if (c < a.c) {
            s = a.b2[c];
            if (a.w2_transposed) {
                for (int j = 0; j < a.csq; ++j) s = fmaf(a.w2[(long long)j * a.c_p + c], hid_all[j], s);
            } else {
                for (int j = 0; j < a.csq; ++j) s = fmaf(a.w2[(long long)c * a.csq + j], hid_all[j], s);
            }
        }
        a.gate[(long long)n * a.c_p + c] = c < a.c ? apply_act(s, EDE_ACT_SIGMOID) : 0.f;
    }
    __threadfence();
    cluster.sync();                                        // gate slices published; nobody reads remote `hid` after this
    if (a.w_src) {
        for (int c = threadIdx.x; c < a.c_p; c += blockDim.x) mean[c] = a.gate[(long long)n * a.c_p + c];
        __syncthreads();
        const int rows = a.w_rows, cols8 = a.w_cols >> 3;
        const uint4* src = reinterpret_cast<const uint4*>(a.w_src);
        uint4* dst = reinterpret_cast<uint4*>(a.w_dst) + (long long)n * a.w_passes * rows * cols8;
        const long long lo_off = (long long)rows * cols8;
        for (int e = rank * blockDim.x + threadIdx.x; e < rows * cols8; e += B * blockDim.x) {
            const int ci = (e % cols8) << 3;
            const int lane8 = a.w_split ? ((ci >> 4) << 3) : ci;
            const uint4 h = src[e];
            uint4 l = make_uint4(0, 0, 0, 0);
            if (a.w_passes == 2) l = src[lo_off + e];
            const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, lw[4] = {l.x, l.y, l.z, l.w};
            uint32_t oh[4], ol[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float g0 = lane8 + 2 * q < a.c_p ? mean[lane8 + 2 * q] : 0.f;
                const float g1 = lane8 + 2 * q + 1 < a.c_p ? mean[lane8 + 2 * q + 1] : 0.f;
                const float v0 = (bf16lo_f(hw[q]) + bf16lo_f(lw[q])) * g0, v1 = (bf16hi_f(hw[q]) + bf16hi_f(lw[q])) * g1;
                const __nv_bfloat162 hh = __floats2bfloat162_rn(v0, v1);
                oh[q] = *reinterpret_cast<const uint32_t*>(&hh);
                const __nv_bfloat162 ll = __floats2bfloat162_rn(v0 - bf16lo_f(oh[q]), v1 - bf16hi_f(oh[q]));
                ol[q] = *reinterpret_cast<const uint32_t*>(&ll);
            }
            dst[e] = make_uint4(oh[0], oh[1], oh[2], oh[3]);
            if (a.w_passes == 2) dst[lo_off + e] = make_uint4(ol[0], ol[1], ol[2], ol[3]);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256) se_scale_kernel(const EdeSeScale a) {
    const int cg = a.c_p >> 3;
    const long long total = (long long)a.n * a.hw * cg;
    pdl_trigger();
    pdl_wait();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cg) << 3;
        const long long pix = i / cg;
        const int n = (int)(pix / a.hw);
        float v[8], g[8];
        load8<T>(reinterpret_cast<const T*>(a.x) + pix * a.x_cs + a.x_co + c, v);
        load8<float>(a.gate + (long long)n * a.c_p + c, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] *= g[j];
        store8<T>(reinterpret_cast<T*>(a.y) + pix * a.y_cs + a.y_co + c, v);
    }
}

// ------------------------------------------------------------------ bilinear resize, align_corners=True
template <typename TX, typename TY>
__global__ void __launch_bounds__(256) resize_kernel(const EdeResize a, float scale_h, float scale_w) {
    const int cg = a.c_p >> 3;
    const long long total = (long long)a.n * a.oh * a.ow * cg;
    const TX* __restrict__ x = reinterpret_cast<const TX*>(a.x);
    TY* __restrict__ y = reinterpret_cast<TY*>(a.y);
    pdl_trigger();
    pdl_wait();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        int c, ow, oh, n;
        if (total <= 0x7fffffffLL) {                         // 32-bit divisions (64-bit ones cost ~10x as much)
            unsigned pix = (unsigned)i / (unsigned)cg;
            c = (int)((unsigned)i - pix * (unsigned)cg) << 3;
            const unsigned row = pix / (unsigned)a.ow;
            ow = (int)(pix - row * (unsigned)a.ow);
            n = (int)(row / (unsigned)a.oh);
            oh = (int)(row - (unsigned)n * (unsigned)a.oh);
        } else {
            c = (int)(i % cg) << 3;
            long long pix = i / cg;
            ow = (int)(pix % a.ow);
            pix /= a.ow;
            oh = (int)(pix % a.oh);
            n = (int)(pix / a.oh);
        }
        // ATen upsample_bilinear2d (align_corners): src = scale * dst, lambda in fp32
        const float sy = scale_h * oh, sx = scale_w * ow;
        const int y0 = (int)sy, x0 = (int)sx;
        const int y1 = y0 + (y0 < a.h - 1 ? 1 : 0), x1 = x0 + (x0 < a.w - 1 ? 1 : 0);
        const float ly = sy - y0, lx = sx - x0;
        const float hy = 1.f - ly, hx = 1.f - lx;
        const long long base = (long long)n * a.h;
        float v00[8], v01[8], v10[8], v11[8], o[8];
        load8<TX>(x + ((base + y0) * a.w + x0) * a.x_cs + a.x_co + c, v00);
        load8<TX>(x + ((base + y0) * a.w + x1) * a.x_cs + a.x_co + c, v01);
        load8<TX>(x + ((base + y1) * a.w + x0) * a.x_cs + a.x_co + c, v10);
        load8<TX>(x + ((base + y1) * a.w + x1) * a.x_cs + a.x_co + c, v11);
#pragma unroll
        for (int j = 0; j < 8; ++j)
            o[j] = hy * (hx * v00[j] + lx * v01[j]) + ly * (hx * v10[j] + lx * v11[j]);
        store8<TY>(y + (((long long)n * a.oh + oh) * a.ow + ow) * a.y_cs + a.y_co + c, o);
    }
}

// Row-blocked form: a block owns R adjacent output rows of one image and 256 (output column, 8-lane group) items.  The
// row terms (source rows, weights) are block-uniform, the column terms are computed once per thread for the R rows, the
// only division left is by the lane-group count, and a thread has 4 R loads in flight (the flat form above: three
// divisions and 4 loads per 16..32 output bytes -- 0.26 of the HBM peak on the 57x76 -> 114x152 maps of D and MFF).
template <typename TX, typename TY, int R>
__global__ void __launch_bounds__(256) resize_rows_kernel(const EdeResize a, float scale_h, float scale_w) {
    const int cg = a.c_p >> 3;
    const int bands = (a.oh + R - 1) / R;
    const int n = blockIdx.y / bands, oh0 = (blockIdx.y - n * bands) * R;
    pdl_trigger();
    const int t = blockIdx.x * 256 + threadIdx.x;
    if (t >= a.ow * cg) return;
    const int ow = t / cg, c = (t - ow * cg) << 3;
    // ATen upsample_bilinear2d (align_corners): src = scale * dst, lambda in fp32
    const float sx = scale_w * ow;
    const int x0 = (int)sx, x1 = x0 + (x0 < a.w - 1 ? 1 : 0);
    const float lx = sx - x0, hx = 1.f - lx;
    const TX* __restrict__ xin = reinterpret_cast<const TX*>(a.x) + (long long)n * a.h * a.w * a.x_cs + a.x_co + c;
    TY* __restrict__ yout = reinterpret_cast<TY*>(a.y) + (((long long)n * a.oh + oh0) * a.ow + ow) * a.y_cs + a.y_co + c;
    const int o0 = x0 * a.x_cs, o1 = x1 * a.x_cs, row_stride = a.w * a.x_cs;
    pdl_wait();
    float v00[R][8], v01[R][8], v10[R][8], v11[R][8], ly[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int oh = min(oh0 + r, a.oh - 1);               // a ragged last band recomputes the last row (not stored)
        const float sy = scale_h * oh;
        const int y0 = (int)sy, y1 = y0 + (y0 < a.h - 1 ? 1 : 0);
        ly[r] = sy - y0;
        const TX* r0 = xin + (long long)y0 * row_stride;
        const TX* r1 = xin + (long long)y1 * row_stride;
        load8<TX>(r0 + o0, v00[r]);
        load8<TX>(r0 + o1, v01[r]);
        load8<TX>(r1 + o0, v10[r]);
        load8<TX>(r1 + o1, v11[r]);
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (oh0 + r >= a.oh) break;
        const float hy = 1.f - ly[r];
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
            o[j] = hy * (hx * v00[r][j] + lx * v01[r][j]) + ly[r] * (hx * v10[r][j] + lx * v11[r][j]);
        store8<TY>(yout + (long long)r * a.ow * a.y_cs, o);
    }
}

// fp32 planes ([planes][h][w], i.e. NCHW with any C): the post-forward depth-map resize
__global__ void __launch_bounds__(256) resize_planes_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                            long long planes, int h, int w, int oh, int ow,
                                                            float scale_h, float scale_w) {
    const long long total = planes * oh * ow;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int ox = (int)(i % ow);
        const long long r = i / ow;
        const int oy = (int)(r % oh);
        const long long pl = r / oh;
        y[i] = bilerp_ac_plane(x + pl * h * w, h, w, scale_h, scale_w, oy, ox);
    }
}

// ------------------------------------------------------------------ 3x3 s2 p1 max pooling
template <typename T>
__global__ void __launch_bounds__(256) maxpool_kernel(const EdeMaxPool a) {
    const int cg = a.c_p >> 3;
    const long long total = (long long)a.n * a.oh * a.ow * cg;
    const T* __restrict__ x = reinterpret_cast<const T*>(a.x);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cg) << 3;
        long long pix = i / cg;
        const int ow = (int)(pix % a.ow);
        pix /= a.ow;
        const int oh = (int)(pix % a.oh);
        const int n = (int)(pix / a.oh);
        float m[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) m[j] = -INFINITY;
        for (int ky = 0; ky < 3; ++ky) {
            const int ih = oh * 2 - 1 + ky;
            if (ih < 0 || ih >= a.h) continue;
            for (int kx = 0; kx < 3; ++kx) {
                const int iw = ow * 2 - 1 + kx;
                if (iw < 0 || iw >= a.w) continue;
                float v[8];
                load8<T>(x + (((long long)n * a.h + ih) * a.w + iw) * a.x_cs + a.x_co + c, v);
#pragma unroll
                for (int j = 0; j < 8; ++j) m[j] = fmaxf(m[j], v[j]);
            }
        }
        store8<T>(reinterpret_cast<T*>(a.y) + (((long long)n * a.oh + oh) * a.ow + ow) * a.y_cs + a.y_co + c, m);
    }
}

// ------------------------------------------------------------------ elementwise act(x) (+ x2)
template <typename T>
__global__ void __launch_bounds__(256) eltwise_kernel(const EdeEltwise a) {
    const int cg = a.c_p >> 3;
    const long long total = (long long)a.n_pix * cg;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cg) << 3;
        const long long pix = i / cg;
        float v[8];
        load8<T>(reinterpret_cast<const T*>(a.x) + pix * a.x_cs + a.x_co + c, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = apply_act(v[j], a.act);
        if (a.x2) {
            float w[8];
            load8<T>(reinterpret_cast<const T*>(a.x2) + pix * a.x2_cs + a.x2_co + c, w);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] += w[j];
        }
        store8<T>(reinterpret_cast<T*>(a.y) + pix * a.y_cs + a.y_co + c, v);
    }
}

// ------------------------------------------------------------------ layout changes at the model boundary
template <typename T>
__global__ void __launch_bounds__(256) nchw_to_nhwc_kernel(const EdeLayout a) {
    const int cg = a.c_p >> 3;
    const long long hw = (long long)a.h * a.w;
    const long long total = (long long)a.n * hw * cg;
    const float* __restrict__ src = reinterpret_cast<const float*>(a.src);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        // pixel fastest so that the per-channel plane reads coalesce
        const long long pix = i % ((long long)a.n * hw);
        const int c = (int)(i / ((long long)a.n * hw)) << 3;
        const int n = (int)(pix / hw);
        const long long p = pix - (long long)n * hw;
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = (c + j < a.c) ? src[((long long)n * a.c + c + j) * hw + p] : 0.f;
        store8<T>(reinterpret_cast<T*>(a.dst) + pix * a.cs + a.co + c, v);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) nhwc_to_nchw_kernel(const EdeLayout a) {
    const long long hw = (long long)a.h * a.w;
    const long long total = (long long)a.n * a.c * hw;
    const T* __restrict__ src = reinterpret_cast<const T*>(a.src);
    float* __restrict__ dst = reinterpret_cast<float*>(a.dst);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long p = i % hw;
        const long long r = i / hw;
        const int c = (int)(r % a.c);
        const long long n = r / a.c;
        dst[i] = Elt<T>::ld(src + (n * hw + p) * a.cs + a.co + c);
    }
}

}  // namespace ede

using namespace ede;

extern "C" {

int ede_dwconv2d(const EdeDwConv* a, void* stream) {
    EDE_REQUIRE(a && a->x && a->y && a->w && a->scale && a->shift, "ede_dwconv2d: null argument");
    EDE_REQUIRE(((a->c_p | a->x_cs | a->x_co | a->y_cs | a->y_co) & 7) == 0, "ede_dwconv2d: channel counts must be multiples of 8");
    EDE_REQUIRE(a->n > 0 && a->oh > 0 && a->ow > 0 && a->k > 0 && a->s > 0, "ede_dwconv2d: bad shape");
    static const bool generic_only = getenv("EDE_DW_GENERIC") != nullptr;   // debug: force the per-output kernel
    if (!generic_only && dwconv_tile_launch(*a, (cudaStream_t)stream, nullptr) == 0) {   // small / medium maps: one tile per block
        EDE_LAUNCH_CHECK("dwconv_tile_kernel");
        return EDE_OK;
    }
    if (!generic_only && dwconv_tma_launch(*a, (cudaStream_t)stream, nullptr) == 0) {   // k in {3,5}, s in {1,2}: TMA-staged
        EDE_LAUNCH_CHECK("dwconv_tma_kernel");
        return EDE_OK;
    }
    const long long per_img = (long long)a->oh * a->ow * (a->c_p >> 3);
    long long gx = (per_img + 256 * 4 - 1) / (256 * 4);
    if (gx < 1) gx = 1;
    if (gx > 2048) gx = 2048;
    dim3 grid((unsigned)gx, (unsigned)a->n);
    const size_t smem = a->pool_sum ? (size_t)a->c_p * sizeof(float) : 0;
    EDE_DISPATCH_DTYPE_DW(a->dtype, T, (dwconv_kernel<T><<<grid, 256, smem, (cudaStream_t)stream>>>(*a)));
    EDE_LAUNCH_CHECK("dwconv_kernel");
    return EDE_OK;
}

int ede_dwconv2d_pool_parts(const EdeDwConv* a) {
    if (a == nullptr || a->ow <= 0 || a->c_p <= 0) return 1;
    int strips = 1;
    if (getenv("EDE_DW_GENERIC") != nullptr) return 1;
    if (dwconv_tile_launch(*a, nullptr, &strips) == 0) return strips < 1 ? 1 : strips;
    if (dwconv_tma_launch(*a, nullptr, &strips) != 0) return 1;
    return strips < 1 ? 1 : strips;
}

int ede_se_fc(const EdeSeFc* a, void* stream) {
    EDE_REQUIRE(a && a->pool_sum && a->w1 && a->b1 && a->w2 && a->b2 && a->gate, "ede_se_fc: null argument");
    EDE_REQUIRE(a->n > 0 && a->c > 0 && a->c <= a->c_p && a->csq > 0, "ede_se_fc: bad shape");
    EDE_REQUIRE(a->w_src == nullptr || (a->w_dst && a->w_rows > 0 && a->w_cols > 0 && (a->w_cols & 7) == 0 &&
                                        (a->w_passes == 1 || a->w_passes == 2)),
                "ede_se_fc: bad weight-fold description");
    const size_t smem = (size_t)(a->c_p + a->csq) * sizeof(float);
    EDE_REQUIRE(smem <= 48 * 1024, "ede_se_fc: too many channels (%d)", a->c_p);
    // A cluster of B CTAs per image where one block per image would leave the machine idle and the FCs are not trivial
    // (se_fc_cluster_kernel); B = the largest of 8 / 4 / 2 that keeps n * B within two waves.
    const char* ck_env = getenv("EDE_SE_FC_CLUSTER");      // 0: never; 2 / 4 / 8: force that cluster size (tests)
    const int cluster_knob = ck_env && *ck_env ? atoi(ck_env) : -1;
    int B = 1;
    if (cluster_knob != 0) {
        const long long macs = 2LL * a->csq * a->c_p + (a->w_src ? (long long)a->w_passes * a->w_rows * a->w_cols / 4 : 0);
        for (int b = 8; b >= 2; b >>= 1)
            if ((long long)a->n * b <= 296) { B = b; break; }
        if (cluster_knob > 0) B = cluster_knob;
        else if (macs < 200000) B = 1;                       // three cluster barriers cost more than the FCs of the narrower
                                                             // stages save (measured: all of EfficientNet-B0 at batch 64)
    }
    if (B > 1) {
        const size_t smem_c = (size_t)(a->c_p + 2 * a->csq) * sizeof(float);
        const cudaError_t e = launch_pdl_cluster(se_fc_cluster_kernel, dim3((unsigned)(a->n * B)), dim3(a->c_p >= 384 || a->w_src ? 1024 : 512),
                                                 smem_c, (cudaStream_t)stream, (unsigned)B, *a);
        if (e == cudaSuccess) {
            EDE_LAUNCH_CHECK("se_fc_cluster_kernel");
            return EDE_OK;
        }
        (void)cudaGetLastError();                            // cluster launch refused: the one-block form below
    }
    // latency bound (one block per image, three dependent phases): wide blocks for the wide late stages
    (void)launch_pdl(se_fc_kernel, dim3(a->n), dim3(a->c_p >= 384 || a->w_src ? 1024 : 512), smem,
                     (cudaStream_t)stream, *a);
    EDE_LAUNCH_CHECK("se_fc_kernel");
    return EDE_OK;
}

int ede_se_scale(const EdeSeScale* a, void* stream) {
    EDE_REQUIRE(a && a->x && a->y && a->gate, "ede_se_scale: null argument");
    EDE_REQUIRE(((a->c_p | a->x_cs | a->x_co | a->y_cs | a->y_co) & 7) == 0, "ede_se_scale: channel counts must be multiples of 8");
    const long long total = (long long)a->n * a->hw * (a->c_p >> 3);
    EDE_DISPATCH_DTYPE_DW(a->dtype, T, ((void)launch_pdl(se_scale_kernel<T>, dim3(grid_persistent(total, 256)), dim3(256), 0, (cudaStream_t)stream, *a)));
    EDE_LAUNCH_CHECK("se_scale_kernel");
    return EDE_OK;
}

int ede_resize_bilinear_ac(const EdeResize* a, void* stream) {
    EDE_REQUIRE(a && a->x && a->y, "ede_resize_bilinear_ac: null argument");
    EDE_REQUIRE(((a->c_p | a->x_cs | a->x_co | a->y_cs | a->y_co) & 7) == 0, "ede_resize_bilinear_ac: channel counts must be multiples of 8");
    EDE_REQUIRE(a->n > 0 && a->h > 0 && a->w > 0 && a->oh > 0 && a->ow > 0, "ede_resize_bilinear_ac: bad shape");
    const float sh = a->oh > 1 ? (float)(a->h - 1) / (float)(a->oh - 1) : 0.f;
    const float sw = a->ow > 1 ? (float)(a->w - 1) / (float)(a->ow - 1) : 0.f;
    const long long total = (long long)a->n * a->oh * a->ow * (a->c_p >> 3);
    cudaStream_t st = (cudaStream_t)stream;
    static const char* flat_knob = getenv("EDE_RESIZE_FLAT");             // 1: the flat grid-stride kernel everywhere
    constexpr int R = 2;
    const long long rows = (long long)a->n * ((a->oh + R - 1) / R);
    const long long items = (long long)a->ow * (a->c_p >> 3);
    if (rows <= 65535 && items >= 128 && !(flat_knob && flat_knob[0] == '1')) {
        const dim3 grid((unsigned)((items + 255) / 256), (unsigned)rows);
        EDE_DISPATCH_DTYPE(a->x_dtype, TX,
                           EDE_DISPATCH_DTYPE(a->y_dtype, TY, ((void)launch_pdl(resize_rows_kernel<TX, TY, R>, grid, dim3(256), 0, st, *a, sh, sw))));
        EDE_LAUNCH_CHECK("resize_rows_kernel");
        return EDE_OK;
    }
    const unsigned g = grid_persistent(total, 256);
    EDE_DISPATCH_DTYPE(a->x_dtype, TX,
                       EDE_DISPATCH_DTYPE(a->y_dtype, TY, ((void)launch_pdl(resize_kernel<TX, TY>, dim3(g), dim3(256), 0, st, *a, sh, sw))));
    EDE_LAUNCH_CHECK("resize_kernel");
    return EDE_OK;
}

int ede_resize_planes_f32(const float* x, float* y, int64_t planes, int32_t h, int32_t w, int32_t oh, int32_t ow,
                          void* stream) {
    EDE_REQUIRE(x && y && planes > 0 && h > 0 && w > 0 && oh > 0 && ow > 0, "ede_resize_planes_f32: bad arguments");
    EDE_REQUIRE(h < (1 << 22) && w < (1 << 22) && (long long)h * w < 0x7fffffffLL, "ede_resize_planes_f32: plane too large");
    const float sh = oh > 1 ? (float)(h - 1) / (float)(oh - 1) : 0.f;
    const float sw = ow > 1 ? (float)(w - 1) / (float)(ow - 1) : 0.f;
    resize_planes_kernel<<<grid_for(planes * oh * ow, 256), 256, 0, (cudaStream_t)stream>>>(x, y, planes, h, w, oh, ow,
                                                                                            sh, sw);
    EDE_LAUNCH_CHECK("resize_planes_kernel");
    return EDE_OK;
}

int ede_maxpool3x3s2(const EdeMaxPool* a, void* stream) {
    EDE_REQUIRE(a && a->x && a->y, "ede_maxpool3x3s2: null argument");
    EDE_REQUIRE(((a->c_p | a->x_cs | a->x_co | a->y_cs | a->y_co) & 7) == 0, "ede_maxpool3x3s2: channel counts must be multiples of 8");
    const long long total = (long long)a->n * a->oh * a->ow * (a->c_p >> 3);
    EDE_DISPATCH_DTYPE(a->dtype, T, (maxpool_kernel<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(*a)));
    EDE_LAUNCH_CHECK("maxpool_kernel");
    return EDE_OK;
}

int ede_eltwise(const EdeEltwise* a, void* stream) {
    EDE_REQUIRE(a && a->x && a->y, "ede_eltwise: null argument");
    EDE_REQUIRE(((a->c_p | a->x_cs | a->x_co | a->y_cs | a->y_co | a->x2_cs | a->x2_co) & 7) == 0, "ede_eltwise: channel counts must be multiples of 8");
    const long long total = (long long)a->n_pix * (a->c_p >> 3);
    EDE_DISPATCH_DTYPE(a->dtype, T, (eltwise_kernel<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(*a)));
    EDE_LAUNCH_CHECK("eltwise_kernel");
    return EDE_OK;
}

int ede_nchw_to_nhwc(const EdeLayout* a, void* stream) {
    EDE_REQUIRE(a && a->src && a->dst, "ede_nchw_to_nhwc: null argument");
    EDE_REQUIRE(((a->c_p | a->cs | a->co) & 7) == 0 && a->c <= a->c_p, "ede_nchw_to_nhwc: bad channel layout");
    const long long total = (long long)a->n * a->h * a->w * (a->c_p >> 3);
    EDE_DISPATCH_DTYPE(a->dtype, T, (nchw_to_nhwc_kernel<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(*a)));
    EDE_LAUNCH_CHECK("nchw_to_nhwc_kernel");
    return EDE_OK;
}

int ede_nhwc_to_nchw(const EdeLayout* a, void* stream) {
    EDE_REQUIRE(a && a->src && a->dst, "ede_nhwc_to_nchw: null argument");
    const long long total = (long long)a->n * a->c * a->h * a->w;
    EDE_DISPATCH_DTYPE(a->dtype, T, (nhwc_to_nchw_kernel<T><<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(*a)));
    EDE_LAUNCH_CHECK("nhwc_to_nchw_kernel");
    return EDE_OK;
}

}  // extern "C"
