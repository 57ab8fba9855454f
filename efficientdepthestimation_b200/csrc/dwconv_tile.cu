// dwconv_tile.cu -- depthwise k x k convolution + BN + activation + squeeze sums for SMALL AND MEDIUM feature maps
// (EfficientNet stages at 57x76 and below: 11 of the 16 depthwise convs of EfficientNet-B0).
//
// The row-streaming kernel (dwconv.cu) keeps its window in registers and is the right shape for the wide early maps, but
// on a 14x19 or 7x9 map one block per (channel chunk, column strip) walking image after image leaves an SM with five
// active warps and a latency-bound row pipeline (0.9-1.8 TB/s measured).  Here a block owns a whole BAND of output rows
// of one image and CH channel lanes:
//   1. ONE TMA box brings the band's input tile (rows x cols x CH lanes, padding and ragged edges = out-of-bounds zero
//      fill) into shared memory;
//   2. split-bf16 lanes are expanded to fp32 IN PLACE once per element (a group of 8 lanes is 32 bytes in either form),
//      so the tap loop below is one LDS.64 + one packed FFMA2 per tap and channel pair instead of two loads and six
//      integer / add instructions per tap;
//   3. every half-/full warp produces two adjacent output columns per step from a K x (K+S) window read row by row
//      (weights of the thread's channel pair live in registers), applies BN + activation, stores, and keeps the squeeze
//      sum of what it stored;
//   4. the block's squeeze sums are reduced in a fixed order and STORED into the band's slab of pool_sum (no atomics:
//      bit-reproducible; ede_se_fc adds the slabs in order).
// Several blocks are resident per SM (<= 72 KB of shared memory, <= 128 registers), so one block's load overlaps the
// others' arithmetic.  HBM-bound integer/fp32 work: no tensor cores.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <type_traits>

#include "common.cuh"
#include "tma.cuh"

namespace ede {

constexpr int DT_THREADS = 256;
constexpr int DT_SMEM_TARGET = 72 * 1024;

struct alignas(64) DwTileParams {
    CUtensorMap tm_x;      // (channel lanes, W, H, N), element = one 4-byte lane
    EdeDwConv a;
    int band_rows;         // output rows per block
    int box_w, box_h;      // input tile (pixels) one block loads
    int pool_write;        // 1: store the squeeze sums into slab blockIdx.x; 0: atomics on slab 0
};

template <typename T> __device__ __forceinline__ void expand_group(uint8_t* g);
template <> __device__ __forceinline__ void expand_group<float>(uint8_t*) {}
template <> __device__ __forceinline__ void expand_group<f32in_t>(uint8_t*) {}     // fp32 lanes already
template <> __device__ __forceinline__ void expand_group<bf16x2_t>(uint8_t* g) {
    const uint4 hi = *reinterpret_cast<const uint4*>(g), lo = *reinterpret_cast<const uint4*>(g + 16);
    float4 a, b;
    a.x = bf16lo_f(hi.x) + bf16lo_f(lo.x); a.y = bf16hi_f(hi.x) + bf16hi_f(lo.x);
    a.z = bf16lo_f(hi.y) + bf16lo_f(lo.y); a.w = bf16hi_f(hi.y) + bf16hi_f(lo.y);
    b.x = bf16lo_f(hi.z) + bf16lo_f(lo.z); b.y = bf16hi_f(hi.z) + bf16hi_f(lo.z);
    b.z = bf16lo_f(hi.w) + bf16lo_f(lo.w); b.w = bf16hi_f(hi.w) + bf16hi_f(lo.w);
    *reinterpret_cast<float4*>(g) = a;
    *reinterpret_cast<float4*>(g + 16) = b;
}

// store two consecutive channels (c even) of one pixel row `px`; returns the value as the next layer will read it
template <typename T> __device__ __forceinline__ float2 store_pair(T* px, int c, float2 v);
template <> __device__ __forceinline__ float2 store_pair<float>(float* px, int c, float2 v) {
    *reinterpret_cast<float2*>(px + c) = v;
    return v;
}
template <> __device__ __forceinline__ float2 store_pair<bf16x2_t>(bf16x2_t* px, int c, float2 v);
template <> __device__ __forceinline__ float2 store_pair<f32in_t>(f32in_t* px, int c, float2 v) {
    return store_pair<bf16x2_t>(static_cast<bf16x2_t*>(px), c, v);
}
template <> __device__ __forceinline__ float2 store_pair<bf16x2_t>(bf16x2_t* px, int c, float2 v) {
    uint32_t* g = reinterpret_cast<uint32_t*>(px + (c & ~7)) + ((c & 7) >> 1);
    const __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
    const uint32_t uh = *reinterpret_cast<const uint32_t*>(&h);
    const float hx = bf16lo_f(uh), hy = bf16hi_f(uh);
    const __nv_bfloat162 l = __floats2bfloat162_rn(v.x - hx, v.y - hy);
    const uint32_t ul = *reinterpret_cast<const uint32_t*>(&l);
    g[0] = uh;
    g[4] = ul;
    return make_float2(hx + bf16lo_f(ul), hy + bf16hi_f(ul));
}

template <int S> struct DwTileRows { static constexpr int value = S == 1 ? 3 : 2; };   // output rows per item

template <typename T, int K, int S, int CH>
__global__ void __launch_bounds__(DT_THREADS, 2) dwconv_tile_kernel(const __grid_constant__ DwTileParams p) {
    constexpr int SUB = 64 / CH;                           // pixel groups per warp (CH / 2 lanes each)
    constexpr int NWARP = DT_THREADS / 32;
    constexpr int PB = CH * 4;                             // bytes of one staged pixel
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) unsigned long long bar;
    __shared__ float pool_s[NWARP * SUB * CH];
    const EdeDwConv& a = p.a;
    uint8_t* const tile = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int band = blockIdx.x, c0 = blockIdx.y * CH, n = blockIdx.z;
    const int oh0 = band * p.band_rows;
    const int rows = min(p.band_rows, a.oh - oh0);
    const int sub = lane / (CH / 2), cl = lane % (CH / 2);
    const int c = c0 + 2 * cl;
    const bool c_ok = c < a.c_p;
    const int cw = c_ok ? c : 0;

    if (threadIdx.x == 0) {
        mbar_init(smem_u32(&bar), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    pdl_trigger();
    // weights / folded BN of this thread's channel pair: constants of the plan, safe to read before the dependency wait
    float2 w[K * K];
#pragma unroll
    for (int t = 0; t < K * K; ++t) w[t] = *reinterpret_cast<const float2*>(a.w + (long long)t * a.c_p + cw);
    const float2 sc = *reinterpret_cast<const float2*>(a.scale + cw);
    const float2 sf = *reinterpret_cast<const float2*>(a.shift + cw);
    __syncthreads();
    pdl_wait();
    if (threadIdx.x == 0) {
        mbar_expect_tx(smem_u32(&bar), (uint32_t)(p.box_w * p.box_h * PB));
        tma_load_4d(smem_u32(tile), &p.tm_x, smem_u32(&bar), c0, -a.pl, oh0 * S - a.pt, n);
    }
    mbar_wait(smem_u32(&bar), 0);
    // split bf16 -> fp32 in place, once per element
    if (std::is_same<T, bf16x2_t>::value) {
        const int groups = p.box_h * p.box_w * (CH / 8);
        for (int g = threadIdx.x; g < groups; g += DT_THREADS) expand_group<T>(tile + (size_t)g * 32);
        __syncthreads();
    }
    // an item = NR output rows x 2 output columns of the thread's channel pair: its (NR-1)*S+K input rows are read once,
    // (K+S) LDS.64 per row, and feed up to NR*2*K packed FMAs -- 2-3x less shared-memory traffic than one row per item
    constexpr int NR = DwTileRows<S>::value;
    constexpr int RIN = (NR - 1) * S + K;
    const bool swish = a.act == EDE_ACT_SWISH;
    const int cpairs = (a.ow + 1) >> 1;
    const int items = ((rows + NR - 1) / NR) * cpairs;
    float2 psum = make_float2(0.f, 0.f);
    for (int it = warp * SUB + sub; it < items; it += NWARP * SUB) {
        const int ib = it / cpairs;
        const int oy = ib * NR, ox = (it - ib * cpairs) * 2;
        const uint8_t* base = tile + ((size_t)(oy * S) * p.box_w + ox * S) * PB + cl * 8;
        float2 acc[NR][2];
#pragma unroll
        for (int j = 0; j < NR; ++j) acc[j][0] = acc[j][1] = make_float2(0.f, 0.f);
#pragma unroll
        for (int r = 0; r < RIN; ++r) {
            float2 xr[K + S];
#pragma unroll
            for (int kx = 0; kx < K + S; ++kx) xr[kx] = *reinterpret_cast<const float2*>(base + kx * PB);
#pragma unroll
            for (int j = 0; j < NR; ++j) {
                const int ky = r - j * S;                  // compile-time after unrolling
                if (ky >= 0 && ky < K) {
#pragma unroll
                    for (int kx = 0; kx < K; ++kx) {
                        acc[j][0] = ffma2(xr[kx], w[ky * K + kx], acc[j][0]);
                        acc[j][1] = ffma2(xr[kx + S], w[ky * K + kx], acc[j][1]);
                    }
                }
            }
            base += (size_t)p.box_w * PB;
        }
        if (c_ok) {
#pragma unroll
            for (int j = 0; j < NR; ++j) {
                if (oy + j >= rows) break;
                T* yrow = reinterpret_cast<T*>(a.y) + (((long long)n * a.oh + oh0 + oy + j) * a.ow + ox) * a.y_cs + a.y_co;
                float2 v0 = make_float2(fmaf(acc[j][0].x, sc.x, sf.x), fmaf(acc[j][0].y, sc.y, sf.y));
                float2 v1 = make_float2(fmaf(acc[j][1].x, sc.x, sf.x), fmaf(acc[j][1].y, sc.y, sf.y));
                if (swish) {
                    v0 = make_float2(swish_fast(v0.x), swish_fast(v0.y));
                    v1 = make_float2(swish_fast(v1.x), swish_fast(v1.y));
                } else {
                    v0 = make_float2(apply_act(v0.x, a.act), apply_act(v0.y, a.act));
                    v1 = make_float2(apply_act(v1.x, a.act), apply_act(v1.y, a.act));
                }
                v0 = store_pair<T>(yrow, c, v0);
                psum.x += v0.x; psum.y += v0.y;
                if (ox + 1 < a.ow) {
                    v1 = store_pair<T>(yrow + a.y_cs, c, v1);
                    psum.x += v1.x; psum.y += v1.y;
                }
            }
        }
    }
    if (a.pool_sum) {                                      // fixed-order block reduction, then one store per channel
        pool_s[(warp * SUB + sub) * CH + 2 * cl] = psum.x;
        pool_s[(warp * SUB + sub) * CH + 2 * cl + 1] = psum.y;
        __syncthreads();
        for (int ch = threadIdx.x; ch < CH; ch += DT_THREADS) {
            if (c0 + ch < a.c_p) {
                float sum = 0.f;
#pragma unroll
                for (int r = 0; r < NWARP * SUB; ++r) sum += pool_s[r * CH + ch];
                if (p.pool_write) a.pool_sum[((long long)band * a.n + n) * a.c_p + c0 + ch] = sum;
                else atomicAdd(a.pool_sum + (long long)n * a.c_p + c0 + ch, sum);
            }
        }
    }
}

// Tile geometry for a shape, or false when the row-streaming kernel is the better fit (wide maps, where a 72 KB tile
// would hold too few rows and re-read its K - S halo rows too often).
struct DwTileCfg { int ch, band_rows, n_bands, box_w, box_h, smem; };

static int env_int_or(const char* name, int dflt) {
    const char* e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
}

static bool dwconv_tile_config(const EdeDwConv& a, DwTileCfg& c) {
    static const char* env = getenv("EDE_DW_TILE");        // "0": never, "1": whenever the shape fits (tests / tuning)
    if (env && env[0] == '0') return false;
    if (a.dtype == EDE_BF16) return false;                 // 2-byte lanes cannot be expanded in place
    if (!((a.k == 3 || a.k == 5) && (a.s == 1 || a.s == 2))) return false;
    const int ow2 = round_up(a.ow, 2);
    const int box_w = (ow2 - 1) * a.s + a.k;
    if (box_w > 256 || a.n > 65535) return false;
    const bool force = env && env[0] == '1';
    double best_cost = 1e30;
    bool found = false;
    // 64 / 32 lanes: a (half-)warp's LDS.64 covers whole 128-byte wavefronts; 16-lane chunks would put the four pixel
    // groups of a warp on the same banks (measured: 4x the shared-memory wavefronts), so they are not offered
    const int chs[2] = {64, 32};
    for (int i = 0; i < 2; ++i) {
        const int ch = chs[i];
        const int row_bytes = box_w * ch * 4;
        const int box_h_max = std::min(256, (DT_SMEM_TARGET - 256) / row_bytes);
        const int nr = a.s == 1 ? 3 : 2;                   // DwTileRows: bands are whole items
        if (box_h_max < a.k) continue;
        const int max_rows = ((box_h_max - a.k) / a.s + 1) / nr * nr;
        if (max_rows < nr) continue;
        const int n_bands = ceil_div(a.oh, max_rows);
        const int band = round_up(ceil_div(a.oh, n_bands), nr);
        const int box_h = (band - 1) * a.s + a.k;
        // cost ~ bytes fetched (halo rows of neighbouring bands re-read, idle channel lanes of the last chunk included;
        // padding rows / columns are TMA zero fill and cost nothing) + a fixed cost per block
        const double chunks = ceil_div(a.c_p, ch);
        const double fetched = (double)(a.h + (n_bands - 1) * (a.k - a.s)) * a.w_ * chunks * ch * 4;
        const double cost = fetched + (double)n_bands * chunks * 6000.0;
        if (cost < best_cost) {
            best_cost = cost;
            c.ch = ch; c.band_rows = band; c.n_bands = n_bands; c.box_w = box_w; c.box_h = box_h;
            c.smem = box_h * row_bytes + 256;
            found = true;
        }
    }
    if (!found) return false;
    if (force) return true;
    // the row-streaming kernel re-uses its register window down the rows (5x less shared-memory traffic per output): it
    // wins on the wide maps; this kernel wins where one map is too small to fill a row pipeline (<= 28x38 outputs)
    static const int max_px = env_int_or("EDE_DW_TILE_MAX_PX", 1100);
    if (a.oh * a.ow > max_px) return false;
    const double ideal = (double)a.h * a.w_ * a.c_p * 4;
    const double fetched = (double)(a.h + (c.n_bands - 1) * (a.k - a.s)) * a.w_ * ceil_div(a.c_p, c.ch) * c.ch * 4;
    return fetched <= 1.5 * ideal + 64.0 * 1024;
}

template <typename T, int K, int S, int CH>
static int dwconv_tile_launch_ch(const EdeDwConv& a, const DwTileCfg& c, cudaStream_t st) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return 1;
    DwTileParams p;
    memset(&p, 0, sizeof(p));
    p.a = a;
    p.band_rows = c.band_rows; p.box_w = c.box_w; p.box_h = c.box_h;
    p.pool_write = a.pool_sum != nullptr && a.pool_parts >= c.n_bands && (c.n_bands == 1 || a.pool_parts > 1);
    cuuint64_t dims[4] = {(cuuint64_t)a.c_p, (cuuint64_t)a.w_, (cuuint64_t)a.h, (cuuint64_t)a.n};
    const cuuint64_t pix = (cuuint64_t)a.x_cs * 4;
    cuuint64_t strides[3] = {pix, pix * a.w_, pix * a.w_ * a.h};
    cuuint32_t box[4] = {(cuuint32_t)CH, (cuuint32_t)c.box_w, (cuuint32_t)c.box_h, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    void* base = (void*)(reinterpret_cast<const uint8_t*>(a.x) + (size_t)a.x_co * 4);
    if (enc(&p.tm_x, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return 1;
    static bool attr_set[64] = {};
    int dev = 0;
    (void)cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!attr_set[dev]) {
        if (cudaFuncSetAttribute(dwconv_tile_kernel<T, K, S, CH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 DT_SMEM_TARGET + 1024) != cudaSuccess) {
            (void)cudaGetLastError();
            return 1;
        }
        attr_set[dev] = true;
    }
    dim3 grid((unsigned)c.n_bands, (unsigned)ceil_div(a.c_p, CH), (unsigned)a.n);
    if (launch_pdl(dwconv_tile_kernel<T, K, S, CH>, grid, dim3(DT_THREADS), (size_t)c.smem, st, p) != cudaSuccess) {
        (void)cudaGetLastError();
        return 1;
    }
    return 0;
}

template <typename T, int K, int S>
static int dwconv_tile_launch_ks(const EdeDwConv& a, const DwTileCfg& c, cudaStream_t st) {
    if (c.ch == 64) return dwconv_tile_launch_ch<T, K, S, 64>(a, c, st);
    if (c.ch == 32) return dwconv_tile_launch_ch<T, K, S, 32>(a, c, st);
    return dwconv_tile_launch_ch<T, K, S, 16>(a, c, st);
}

// returns 0 when the tile kernel took the op (or, with `bands_out`, when it WOULD: host query, nothing launched)
int dwconv_tile_launch(const EdeDwConv& a, cudaStream_t st, int* bands_out) {
    DwTileCfg c;
    if (!dwconv_tile_config(a, c)) return 1;
    if (bands_out) { *bands_out = c.n_bands; return 0; }
    if ((reinterpret_cast<uintptr_t>(a.x) & 15) != 0 || ((a.x_cs * 4) & 15) != 0) return 1;
    const int ks = a.k * 10 + a.s;
    int rc = 1;
    if (a.dtype == EDE_BF16X2) {
        using T = bf16x2_t;
        if (ks == 31) rc = dwconv_tile_launch_ks<T, 3, 1>(a, c, st);
        else if (ks == 32) rc = dwconv_tile_launch_ks<T, 3, 2>(a, c, st);
        else if (ks == 51) rc = dwconv_tile_launch_ks<T, 5, 1>(a, c, st);
        else if (ks == 52) rc = dwconv_tile_launch_ks<T, 5, 2>(a, c, st);
    } else if (a.dtype == EDE_F32_TO_BF16X2) {
        using T = f32in_t;
        if (ks == 31) rc = dwconv_tile_launch_ks<T, 3, 1>(a, c, st);
        else if (ks == 32) rc = dwconv_tile_launch_ks<T, 3, 2>(a, c, st);
        else if (ks == 51) rc = dwconv_tile_launch_ks<T, 5, 1>(a, c, st);
        else if (ks == 52) rc = dwconv_tile_launch_ks<T, 5, 2>(a, c, st);
    } else {
        using T = float;
        if (ks == 31) rc = dwconv_tile_launch_ks<T, 3, 1>(a, c, st);
        else if (ks == 32) rc = dwconv_tile_launch_ks<T, 3, 2>(a, c, st);
        else if (ks == 51) rc = dwconv_tile_launch_ks<T, 5, 1>(a, c, st);
        else if (ks == 52) rc = dwconv_tile_launch_ks<T, 5, 2>(a, c, st);
    }
    return rc;
}

}  // namespace ede
