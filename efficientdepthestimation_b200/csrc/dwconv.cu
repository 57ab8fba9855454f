// dwconv.cu -- depthwise k x k convolution + BN + activation + squeeze-excite pooling, TMA-staged.
// (the k = 3 / 5, stride 1 / 2 depthwise stage of every EfficientNet MBConv: SURVEY.md Appendix A.3)
//
// HBM-bound: every input element is fetched from global memory once per block and every output written once.
//   * A block owns (image n, 64-channel chunk, `cols` adjacent output columns) and walks DOWN the output rows.
//   * One producer warp streams input rows through a ring of smem row slots with TMA (tensor map over the NHWC
//     tensor, box = 64 channels x (cols*S + K - S) pixels x 1 row).  Zero padding, ragged right/bottom edges and
//     channel tails are TMA out-of-bounds fill: the compute loop has no bounds checks on its inputs.
//   * Consumer warp w owns output column w of the tile; lane l owns two channels.  The K x K window and the
//     K x K weights of those two channels live in registers; per output row a thread reads only the S new
//     input rows (K pixels each) from the ring.  Window slots rotate by compile-time indices because the walk
//     is unrolled K rows at a time.
//   * The squeeze (per-(n, c) sum of the STORED activations) is a per-thread running sum, one smem reduction
//     over the block's columns and one atomicAdd per channel per block.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "tma.cuh"

namespace ede {

__device__ __forceinline__ float2 bf16pair_to_f2(uint32_t u) {
    return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
}
__device__ __forceinline__ uint32_t f2_to_bf16pair(float2 v) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// two consecutive channels (2*lane, 2*lane+1) of one pixel: smem read (64-channel slice at `px`), global store
template <typename T> struct Pair;
template <> struct Pair<float> {
    static constexpr int kPixelBytes = 256;
    static __device__ __forceinline__ float2 lds(const uint8_t* px, int lane) {
        return *reinterpret_cast<const float2*>(px + lane * 8);
    }
    static __device__ __forceinline__ float2 st(float* px, int c, float2 v) {
        *reinterpret_cast<float2*>(px + c) = v;
        return v;
    }
};
template <> struct Pair<__nv_bfloat16> {
    static constexpr int kPixelBytes = 128;
    static __device__ __forceinline__ float2 lds(const uint8_t* px, int lane) {
        return bf16pair_to_f2(*reinterpret_cast<const uint32_t*>(px + lane * 4));
    }
    static __device__ __forceinline__ float2 st(__nv_bfloat16* px, int c, float2 v) {   // returns the stored value
        const uint32_t u = f2_to_bf16pair(v);
        *reinterpret_cast<uint32_t*>(px + c) = u;
        return bf16pair_to_f2(u);
    }
};
template <> struct Pair<bf16x2_t> {
    // 8-lane group g at byte 32 g: [8 x hi][8 x lo]; channels (c, c+1), c even -> one u32 in each half
    static constexpr int kPixelBytes = 256;
    static __device__ __forceinline__ float2 lds(const uint8_t* px, int lane) {
        const uint32_t* g = reinterpret_cast<const uint32_t*>(px + (lane >> 2) * 32 + (lane & 3) * 4);
        const float2 h = bf16pair_to_f2(g[0]), l = bf16pair_to_f2(g[4]);
        return make_float2(h.x + l.x, h.y + l.y);
    }
    static __device__ __forceinline__ float2 st(bf16x2_t* px, int c, float2 v) {
        uint32_t* g = reinterpret_cast<uint32_t*>(px + (c & ~7)) + ((c & 7) >> 1);
        const uint32_t uh = f2_to_bf16pair(v);
        const float2 h = bf16pair_to_f2(uh);
        const uint32_t ul = f2_to_bf16pair(make_float2(v.x - h.x, v.y - h.y));
        const float2 l = bf16pair_to_f2(ul);
        g[0] = uh;
        g[4] = ul;
        return make_float2(h.x + l.x, h.y + l.y);
    }
};

struct alignas(64) DwParams {
    CUtensorMap tm_x;      // (64-channel lanes, W, H, N), element = one lane of the dtype
    EdeDwConv a;
    int cols;              // output columns (= consumer warps) per block
    int row_bytes;         // bytes of one ring slot: (cols*S + K - S) pixels x 64 lanes
};

template <int K> struct DwDepth { static constexpr int value = K >= 5 ? 16 : 8; };   // ring slots (power of 2)
constexpr int DW_MAX_COLS = 16;
// consumer warps per block: the 5x5 window + weights need ~170 registers per thread -> 12 warps (384 threads)
template <int K> struct DwCols { static constexpr int value = K >= 5 ? 11 : DW_MAX_COLS; };

template <typename T, int K, int S>
__global__ void __launch_bounds__(32 * (DwCols<K>::value + 1), 1) dwconv_tma_kernel(const __grid_constant__ DwParams p) {
    constexpr int DEPTH = DwDepth<K>::value;
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) unsigned long long bars[2 * DEPTH];
    __shared__ float pool_s[DW_MAX_COLS][64];
    const EdeDwConv& a = p.a;
    const uint32_t ring = (smem_u32(smem_raw) + 127u) & ~127u;
    const uint8_t* const ring_g = smem_raw + (ring - smem_u32(smem_raw));
    const uint32_t bar0 = smem_u32(bars);
    auto full = [&](int s) { return bar0 + 8u * s; };
    auto empty = [&](int s) { return bar0 + 8u * (DEPTH + s); };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ow0 = blockIdx.x * p.cols;
    const int c0 = blockIdx.y * 64;
    const int n = blockIdx.z;
    const int n_act = min(p.cols, a.ow - ow0);            // consumer warps that own a real output column
    const int rows_in = (a.oh - 1) * S + K;               // input rows (incl. padding rows) the walk touches

    if (threadIdx.x == 0) {
        for (int s = 0; s < DEPTH; ++s) {
            mbar_init(full(s), 1);
            mbar_init(empty(s), n_act);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    float2 psum = make_float2(0.f, 0.f);
    if (warp == p.cols) {
        // ===================== producer: one TMA row per ring slot
        for (int ir = 0; ir < rows_in; ++ir) {
            const int slot = ir & (DEPTH - 1);
            mbar_wait(empty(slot), ((ir / DEPTH) & 1) ^ 1);
            if (elect_one()) {
                mbar_expect_tx(full(slot), (uint32_t)p.row_bytes);
                tma_load_4d(ring + slot * p.row_bytes, &p.tm_x, full(slot), c0, ow0 * S - a.pl, ir - a.pt, n);
            }
            __syncwarp();
        }
    } else if (warp < n_act) {
        // ===================== consumers
        const int c = c0 + 2 * lane;
        const bool c_ok = c < a.c_p;
        const int cw = c_ok ? c : 0;                       // idle lanes compute on channel 0 and never store
        float2 w[K * K];
#pragma unroll
        for (int t = 0; t < K * K; ++t) w[t] = *reinterpret_cast<const float2*>(a.w + (long long)t * a.c_p + cw);
        const float2 sc = *reinterpret_cast<const float2*>(a.scale + cw);
        const float2 sf = *reinterpret_cast<const float2*>(a.shift + cw);
        const int ow = ow0 + warp;
        T* const ycol = reinterpret_cast<T*>(a.y) + ((long long)n * a.oh * a.ow + ow) * a.y_cs + a.y_co;
        const uint8_t* const colbase = ring_g + (size_t)(warp * S) * Pair<T>::kPixelBytes;
        float2 xw[K][K];                                   // [row slot][kx]
        auto load_row = [&](int ir, float2 (&row)[K]) {
            const int slot = ir & (DEPTH - 1);
            mbar_wait(full(slot), (ir / DEPTH) & 1);
            const uint8_t* px = colbase + (size_t)slot * p.row_bytes;
#pragma unroll
            for (int kx = 0; kx < K; ++kx) row[kx] = Pair<T>::lds(px + kx * Pair<T>::kPixelBytes, lane);
            // the row now lives in registers of every lane: hand the slot back to the producer at once
            // (the arrive has release semantics, so the reads above are ordered before the TMA refill)
            __syncwarp();
            if (lane == 0) mbar_arrive(empty(slot));
        };
#pragma unroll
        for (int ky = 0; ky < K - S; ++ky) load_row(ky, xw[ky]);
        for (int oh0 = 0; oh0 < a.oh; oh0 += K) {
#pragma unroll
            for (int u = 0; u < K; ++u) {
                const int oh = oh0 + u;
                if (oh < a.oh) {                           // block-uniform
                    // window row ky of output row oh lives in slot (ky + S*u) % K; the S new rows arrive now
#pragma unroll
                    for (int j = 0; j < S; ++j) load_row(oh * S + K - S + j, xw[(K - S + j + S * u) % K]);
                    float2 acc0 = make_float2(0.f, 0.f), acc1 = make_float2(0.f, 0.f);
#pragma unroll
                    for (int ky = 0; ky < K; ++ky)
#pragma unroll
                        for (int kx = 0; kx < K; ++kx) {
                            const float2 xv = xw[(ky + S * u) % K][kx], wv = w[ky * K + kx];
                            if ((ky * K + kx) & 1) { acc1.x = fmaf(xv.x, wv.x, acc1.x); acc1.y = fmaf(xv.y, wv.y, acc1.y); }
                            else { acc0.x = fmaf(xv.x, wv.x, acc0.x); acc0.y = fmaf(xv.y, wv.y, acc0.y); }
                        }
                    float2 v = make_float2(apply_act(fmaf(acc0.x + acc1.x, sc.x, sf.x), a.act),
                                           apply_act(fmaf(acc0.y + acc1.y, sc.y, sf.y), a.act));
                    if (c_ok) {
                        v = Pair<T>::st(ycol + (long long)oh * a.ow * a.y_cs, c, v);   // the squeeze sums what is stored
                        psum.x += v.x;
                        psum.y += v.y;
                    }
                }
            }
        }
    }
    if (a.pool_sum) {
        if (warp < p.cols) {
            pool_s[warp][2 * lane] = psum.x;
            pool_s[warp][2 * lane + 1] = psum.y;
        }
        __syncthreads();
        if (threadIdx.x < 64 && c0 + threadIdx.x < a.c_p) {
            float s = 0.f;
            for (int r = 0; r < n_act; ++r) s += pool_s[r][threadIdx.x];
            atomicAdd(a.pool_sum + (long long)n * a.c_p + c0 + threadIdx.x, s);
        }
    }
}

// columns per block: as many consumer warps as registers allow, sized to waste as little of the last tile as possible
static int pick_cols(int ow, int max_cols) {
    int best = 1;
    double best_cost = 1e30;
    for (int c = 1; c <= max_cols; ++c) {
        const int tiles = ceil_div(ow, c);
        const double cost = (double)tiles * (c + 1.5);       // +1.5: producer warp + halo columns per tile
        if (cost < best_cost - 1e-9 || (cost < best_cost + 1e-9 && c > best)) { best_cost = cost; best = c; }
    }
    return best;
}

template <typename T, int K, int S>
static int dwconv_tma_launch_ks(const EdeDwConv& a, cudaStream_t st) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return 1;
    constexpr int DEPTH = DwDepth<K>::value;
    DwParams p;
    memset(&p, 0, sizeof(p));
    p.a = a;
    p.cols = pick_cols(a.ow, DwCols<K>::value);
    const int esz = a.dtype == EDE_BF16 ? 2 : 4;
    const int box_w = p.cols * S + K - S;
    p.row_bytes = box_w * 64 * esz;
    cuuint64_t dims[4] = {(cuuint64_t)a.c_p, (cuuint64_t)a.w_, (cuuint64_t)a.h, (cuuint64_t)a.n};
    const cuuint64_t pix = (cuuint64_t)a.x_cs * esz;
    cuuint64_t strides[3] = {pix, pix * a.w_, pix * a.w_ * a.h};
    cuuint32_t box[4] = {64, (cuuint32_t)box_w, 1, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    void* base = (void*)(reinterpret_cast<const uint8_t*>(a.x) + (size_t)a.x_co * esz);
    CUresult r = enc(&p.tm_x, esz == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, base, dims,
                     strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return 1;
    const int smem = DEPTH * p.row_bytes + 128;
    static bool attr_set[64] = {};
    int dev = 0;
    (void)cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!attr_set[dev]) {
        if (cudaFuncSetAttribute(dwconv_tma_kernel<T, K, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) !=
            cudaSuccess) {
            (void)cudaGetLastError();
            return 1;
        }
        attr_set[dev] = true;
    }
    dim3 grid((unsigned)ceil_div(a.ow, p.cols), (unsigned)ceil_div(a.c_p, 64), (unsigned)a.n);
    if (grid.y > 65535u || grid.z > 65535u || smem > 200 * 1024) return 1;
    dwconv_tma_kernel<T, K, S><<<grid, 32 * (p.cols + 1), smem, st>>>(p);
    return 0;
}

// returns 0 when the TMA kernel was launched, 1 when this (k, s) / device has no TMA path (caller falls back)
int dwconv_tma_launch(const EdeDwConv& a, cudaStream_t st) {
    if ((reinterpret_cast<uintptr_t>(a.x) & 15) != 0) return 1;
    const int ks = a.k * 10 + a.s;
    int rc = 1;
    EDE_DISPATCH_DTYPE(a.dtype, T, {
        if (ks == 31) rc = dwconv_tma_launch_ks<T, 3, 1>(a, st);
        else if (ks == 32) rc = dwconv_tma_launch_ks<T, 3, 2>(a, st);
        else if (ks == 51) rc = dwconv_tma_launch_ks<T, 5, 1>(a, st);
        else if (ks == 52) rc = dwconv_tma_launch_ks<T, 5, 2>(a, st);
    });
    return rc;
}

}  // namespace ede
