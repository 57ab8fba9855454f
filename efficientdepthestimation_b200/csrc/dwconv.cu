// dwconv.cu -- depthwise k x k convolution + BN + activation + squeeze-excite pooling, TMA-staged.
// (the k = 3 / 5, stride 1 / 2 depthwise stage of every EfficientNet MBConv: SURVEY.md Appendix A.3)
//
// HBM-bound: every input element is fetched from global memory once per block and every output written once.
//   * A block owns (image n, 64-channel chunk, `cols` adjacent output columns) and walks DOWN the output rows.
//   * One producer warp streams input rows through a ring of smem row slots with TMA (tensor map over the NHWC
//     tensor, box = 64 channels x (cols*S + K - S) pixels x 1 row).  Zero padding, ragged right/bottom edges and
//     channel tails are TMA out-of-bounds fill: the compute loop has no bounds checks on its inputs.
//   * Consumer warp w owns NC adjacent output columns of the tile (NC = 2 where the registers allow: the walk is
//     instruction-issue bound, and two columns share the row wait, the addressing and K - S of their K input
//     pixels); lane l owns two channels.  The K x (K + (NC-1) S) window and the K x K weights of those two
//     channels live in registers; per output row a thread reads only the S new input rows from the ring.
//     Window slots rotate by compile-time indices because the walk is unrolled K rows at a time.
//   * The squeeze (per-(n, c) sum of the STORED activations) is a per-thread running sum, one smem reduction
//     over the block's columns and one store per channel per block into the block's slab of pool_sum (atomicAdd
//     only in the single-slab form).
#include <stdlib.h>
#include <string.h>

#include <algorithm>

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

template <> struct Pair<f32in_t> {       // fp32 lanes in, split-bf16 lanes out
    static constexpr int kPixelBytes = 256;
    static __device__ __forceinline__ float2 lds(const uint8_t* px, int lane) { return Pair<float>::lds(px, lane); }
    static __device__ __forceinline__ float2 st(f32in_t* px, int c, float2 v) {
        return Pair<bf16x2_t>::st(static_cast<bf16x2_t*>(px), c, v);
    }
};

struct alignas(64) DwParams {
    CUtensorMap tm_x;      // (channel lanes, W, H, N), element = one lane of the dtype
    EdeDwConv a;
    int cols;              // consumer warps per block (each owns 64 / CH adjacent output columns)
    int row_bytes;         // bytes of one ring slot: (cols * (64/CH) * S + K - S) pixels x CH lanes
    int tx_bytes;          // bytes one TMA row box delivers (row_bytes before the 128-byte rounding)
    int zsplit;            // grid.z: a block walks (image, row band) units blockIdx.z, blockIdx.z + zsplit, ...
    int bands, band_oh;    // row bands per image and output rows per band: a walk can start at any output row (it
                           // re-reads K - S input rows), so small batches of large maps still fill the machine
    int pool_write;        // 1: every column strip stores its squeeze sums into its own slab of pool_sum
};

template <int K> struct DwDepth { static constexpr int value = K >= 5 ? 16 : 8; };   // ring slots (power of 2)
constexpr int DW_MAX_COLS = 16;
// consumer warps per block: the 5x5 window + weights need ~130 registers per thread -> 12 warps (384 threads)
template <int K> struct DwCols { static constexpr int value = K >= 5 ? 11 : DW_MAX_COLS; };

// CH = channel lanes per block: 64 (lane = channel pair, warp = one output column) or 32 (half-warp = channel
// pairs, the two halves of a warp own two adjacent output columns) -- the latter when c_p % 64 is in (0, 32],
// e.g. the 32 / 96 / 144 / 480 / 672-channel stages of EfficientNet-B0, so that no lane idles.
template <typename T, int K, int S, int CH, int NC>
__global__ void __launch_bounds__(32 * (DwCols<K>::value + 1), 1) dwconv_tma_kernel(const __grid_constant__ DwParams p) {
    constexpr int DEPTH = DwDepth<K>::value;
    constexpr int SUB = 64 / CH;                           // column groups per warp (each NC output columns)
    constexpr int NI = (NC - 1) * S + K;                   // input pixels of a row one thread reads
    constexpr int PB = Pair<T>::kPixelBytes * CH / 64;     // bytes of one staged pixel
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) unsigned long long bars[2 * DEPTH];
    __shared__ float pool_s[DW_MAX_COLS * 64];             // [column of the tile][CH]
    const EdeDwConv& a = p.a;
    const uint32_t ring = (smem_u32(smem_raw) + 127u) & ~127u;
    const uint8_t* const ring_g = smem_raw + (ring - smem_u32(smem_raw));
    const uint32_t bar0 = smem_u32(bars);
    auto full = [&](int s) { return bar0 + 8u * s; };
    auto empty = [&](int s) { return bar0 + 8u * (DEPTH + s); };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ow0 = blockIdx.x * p.cols * SUB * NC;
    const int c0 = blockIdx.y * CH;
    const int n_act = min(p.cols, (a.ow - ow0 + SUB * NC - 1) / (SUB * NC));   // consumer warps that own a real output column
    const int n_units = a.n * p.bands;                    // (image, row band) walks

    if (threadIdx.x == 0) {
        for (int s = 0; s < DEPTH; ++s) {
            mbar_init(full(s), 1);
            mbar_init(empty(s), n_act);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    pdl_trigger();
    __syncthreads();
    pdl_wait();

    if (warp == p.cols) {
        // ===================== producer: one TMA row per ring slot, images back to back
        if (lane == 0) {
            int g = 0;                                     // rows issued so far (ring position)
            for (int v = blockIdx.z; v < n_units; v += p.zsplit) {
                const int n = v / p.bands, oh_first = (v - n * p.bands) * p.band_oh;
                const int rows_out = min(p.band_oh, a.oh - oh_first);
                const int ir0 = oh_first * S, rows_in = (rows_out - 1) * S + K;   // input rows (incl. padding rows) of the walk
                for (int ir = ir0; ir < ir0 + rows_in; ++ir, ++g) {
                    const int slot = g & (DEPTH - 1);
                    mbar_wait(empty(slot), ((g / DEPTH) & 1) ^ 1);
                    mbar_expect_tx(full(slot), (uint32_t)p.tx_bytes);
                    tma_load_4d(ring + slot * p.row_bytes, &p.tm_x, full(slot), c0, ow0 * S - a.pl, ir - a.pt, n);
                }
            }
        }
    } else if (warp < n_act) {
        // ===================== consumers
        const int sub = lane / (CH / 2), cl = lane % (CH / 2);     // column within the warp, channel pair
        const int c = c0 + 2 * cl;
        const int col = (warp * SUB + sub) * NC, ow = ow0 + col;   // first of this thread's NC output columns
        const bool ok = c < a.c_p && ow < a.ow;
        const bool swish = a.act == EDE_ACT_SWISH;
        const int cw = c < a.c_p ? c : 0;                  // idle lanes compute on channel 0 and never store
        float2 w[K * K];
#pragma unroll
        for (int t = 0; t < K * K; ++t) w[t] = *reinterpret_cast<const float2*>(a.w + (long long)t * a.c_p + cw);
        const float2 sc = *reinterpret_cast<const float2*>(a.scale + cw);
        const float2 sf = *reinterpret_cast<const float2*>(a.shift + cw);
        const uint8_t* const colbase = ring_g + (size_t)(col * S) * PB;
        const int n_cons = n_act * 32, ctid = threadIdx.x;     // consumer warps are warps 0 .. n_act-1
        int slot = 0;                                      // ring slot of the next row this warp loads ...
        uint32_t phase = 0;                                // ... and its barrier phase
        const long long y_row = (long long)a.ow * a.y_cs;  // elements between two output rows
        for (int v = blockIdx.z; v < n_units; v += p.zsplit) {
            const int n = v / p.bands, band = v - n * p.bands, oh_first = band * p.band_oh;
            const int rows_out = min(p.band_oh, a.oh - oh_first);
            T* yrow = reinterpret_cast<T*>(a.y) + (((long long)n * a.oh + oh_first) * a.ow + (ok ? ow : 0)) * a.y_cs + a.y_co;
            float2 psum = make_float2(0.f, 0.f);
            float2 xw[K][NI];                              // [row slot][input column]
            auto load_row = [&](float2 (&row)[NI]) {
                mbar_wait(full(slot), phase);
                const uint8_t* px = colbase + slot * p.row_bytes;
#pragma unroll
                for (int kx = 0; kx < NI; ++kx) row[kx] = Pair<T>::lds(px + kx * PB, cl);
                // the row now lives in registers of every lane: hand the slot back to the producer at once
                // (the arrive has release semantics, so the reads above are ordered before the TMA refill)
                __syncwarp();
                if (lane == 0) mbar_arrive(empty(slot));
                if (++slot == DEPTH) { slot = 0; phase ^= 1u; }
            };
#pragma unroll
            for (int ky = 0; ky < K - S; ++ky) load_row(xw[ky]);
            for (int oh0 = 0; oh0 < rows_out; oh0 += K) {  // output rows relative to the band
#pragma unroll
                for (int u = 0; u < K; ++u) {
                    const int oh = oh0 + u;
                    if (oh < rows_out) {                   // block-uniform
                        // window row ky of output row oh lives in slot (ky + S*u) % K; the S new rows arrive now
#pragma unroll
                        for (int j = 0; j < S; ++j) load_row(xw[(K - S + j + S * u) % K]);
#pragma unroll
                        for (int j = 0; j < NC; ++j) {
                            float2 acc0 = make_float2(0.f, 0.f), acc1 = make_float2(0.f, 0.f);
#pragma unroll
                            for (int ky = 0; ky < K; ++ky)
#pragma unroll
                                for (int kx = 0; kx < K; ++kx) {
                                    const float2 xv = xw[(ky + S * u) % K][kx + j * S], wv = w[ky * K + kx];
                                    // the channel pair is one packed fp32 FMA (FFMA2): same values, half the issue slots
                                    if ((ky * K + kx) & 1) acc1 = ffma2(xv, wv, acc1);
                                    else acc0 = ffma2(xv, wv, acc0);
                                }
                            float2 v = make_float2(fmaf(acc0.x + acc1.x, sc.x, sf.x), fmaf(acc0.y + acc1.y, sc.y, sf.y));
                            if (swish) v = make_float2(swish_fast(v.x), swish_fast(v.y));
                            else v = make_float2(apply_act(v.x, a.act), apply_act(v.y, a.act));
                            if (ok && ow + j < a.ow) {
                                v = Pair<T>::st(yrow + j * a.y_cs, c, v);       // the squeeze sums what is stored
                                psum.x += v.x;
                                psum.y += v.y;
                            }
                        }
                        yrow += y_row;
                    }
                }
            }
            if (a.pool_sum) {                              // per-image squeeze: reduce over the tile's columns
                pool_s[(col / NC) * CH + 2 * cl] = psum.x;
                pool_s[(col / NC) * CH + 2 * cl + 1] = psum.y;
                asm volatile("bar.sync 1, %0;" ::"r"(n_cons) : "memory");
                for (int ch = ctid; ch < CH; ch += n_cons) {
                    if (c0 + ch < a.c_p) {
                        float sum = 0.f;
                        for (int r = 0; r < n_act * SUB; ++r) sum += pool_s[r * CH + ch];
                        // one slab per column strip: plain stores, summed in slab order by se_fc (reproducible);
                        // more strips than slabs (very wide maps): atomics on slab 0
                        if (p.pool_write) a.pool_sum[((long long)(blockIdx.x * p.bands + band) * a.n + n) * a.c_p + c0 + ch] = sum;
                        else atomicAdd(a.pool_sum + (long long)n * a.c_p + c0 + ch, sum);
                    }
                }
                asm volatile("bar.sync 1, %0;" ::"r"(n_cons) : "memory");
            }
        }
    }
}

// Row bands and image groups.  One block per SM is resident (registers).  A block's fixed cost (weights, barrier init,
// ring fill) is a few rows' worth of work, so small maps let one block walk several images, and large maps at a small
// batch are cut into row bands (each band re-reads K - S input rows): pick (bands, z) that minimises
// waves x (walks per block x (rows per walk + 3) + 6).  Pure host arithmetic on 148 SMs: ede_dwconv2d_pool_parts must
// give the same answer without a device (the squeeze slabs are per column strip AND band).
static void pick_bands_z(const EdeDwConv& a, int units, int* bands_out, int* z_out) {
    const char* b_env = getenv("EDE_DW_BANDS");               // debug / tests: force the number of row bands
    static const char* z_env = getenv("EDE_DW_ZSPLIT");       // debug / tests: force the number of walk groups
    const int sms = 148;
    struct Memo { int n, oh, units, b, z; };
    thread_local Memo memo[16];
    thread_local int memo_n = 0;
    if (!(b_env && *b_env) && !(z_env && *z_env))
        for (int i = 0; i < memo_n; ++i)
            if (memo[i].n == a.n && memo[i].oh == a.oh && memo[i].units == units) { *bands_out = memo[i].b; *z_out = memo[i].z; return; }
    int best_b = 1, best_z = a.n;
    double best_cost = 1e30;
    // bands only where whole images cannot fill the machine (measured: with >= one walk per SM the band split of the
    // 114x152x32 map of ENB0-HU batch 64 cost 8 %: more halo rows, same occupancy)
    const int max_b = b_env && *b_env ? std::max(1, std::min(a.oh, atoi(b_env)))
                                      : (units * a.n >= sms ? 1 : std::max(1, a.oh / 12));
    for (int b = (b_env && *b_env) ? max_b : 1; b <= max_b; ++b) {
        const int band_oh = ceil_div(a.oh, b);
        if (ceil_div(a.oh, band_oh) != b) continue;           // every band owns at least one row
        const int walks = a.n * b;
        for (int z = 1; z <= walks; ++z) {
            const double waves = (double)ceil_div(units * z, sms);
            const double cost = waves * ((double)ceil_div(walks, z) * (band_oh + 3.0) + 6.0);
            if (cost < best_cost - 1e-9) { best_cost = cost; best_b = b; best_z = z; }
        }
    }
    if (z_env && *z_env) best_z = std::max(1, std::min(a.n * best_b, atoi(z_env)));
    else if (!(b_env && *b_env)) { memo[memo_n % 16] = Memo{a.n, a.oh, units, best_b, best_z}; if (memo_n < 16) ++memo_n; }
    *bands_out = best_b;
    *z_out = best_z;
}

// consumer warps per block: as many as registers allow, sized to waste as little of the last tile as possible
static int pick_cols(int ow_units, int max_cols) {
    int best = 1;
    double best_cost = 1e30;
    for (int c = 1; c <= max_cols; ++c) {
        const int tiles = ceil_div(ow_units, c);
        const double cost = (double)tiles * (c + 1.5);       // +1.5: producer warp + halo columns per tile
        if (cost < best_cost - 1e-9 || (cost < best_cost + 1e-9 && c > best)) { best_cost = cost; best = c; }
    }
    return best;
}

template <typename T, int K, int S, int CH, int NC>
static int dwconv_tma_launch_ch(const EdeDwConv& a, cudaStream_t st, int* strips_out) {
    constexpr int DEPTH = DwDepth<K>::value;
    constexpr int SUB = 64 / CH;
    DwParams p;
    memset(&p, 0, sizeof(p));
    p.a = a;
    p.cols = pick_cols(ceil_div(a.ow, SUB * NC), DwCols<K>::value);
    const int units = ceil_div(a.ow, p.cols * SUB * NC) * ceil_div(a.c_p, CH);
    int bands = 1, best_z = a.n;
    pick_bands_z(a, units, &bands, &best_z);
    if (strips_out) {                                     // host query: squeeze slabs = column strips (grid.x) x row bands
        if (p.cols * SUB * NC * S + K - S > 256) return 1;
        *strips_out = ceil_div(a.ow, p.cols * SUB * NC) * bands;
        return 0;
    }
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return 1;
    const int esz = a.dtype == EDE_BF16 ? 2 : 4;
    const int box_w = p.cols * SUB * NC * S + K - S;
    if (box_w > 256) return 1;
    p.tx_bytes = box_w * CH * esz;
    p.row_bytes = round_up(box_w * CH * esz, 128);          // TMA shared-memory destinations are 128-byte aligned
    cuuint64_t dims[4] = {(cuuint64_t)a.c_p, (cuuint64_t)a.w_, (cuuint64_t)a.h, (cuuint64_t)a.n};
    const cuuint64_t pix = (cuuint64_t)a.x_cs * esz;
    cuuint64_t strides[3] = {pix, pix * a.w_, pix * a.w_ * a.h};
    cuuint32_t box[4] = {(cuuint32_t)CH, (cuuint32_t)box_w, 1, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    void* base = (void*)(reinterpret_cast<const uint8_t*>(a.x) + (size_t)a.x_co * esz);
    CUresult r = enc(&p.tm_x, esz == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, base, dims,
                     strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return 1;
    const int smem = DEPTH * p.row_bytes + 128;
    static bool attr_set[64] = {};
    static int sm_count[64] = {};
    int dev = 0;
    (void)cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!attr_set[dev]) {
        if (cudaFuncSetAttribute(dwconv_tma_kernel<T, K, S, CH, NC>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 200 * 1024) != cudaSuccess ||
            cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
            (void)cudaGetLastError();
            return 1;
        }
        attr_set[dev] = true;
    }
    p.zsplit = best_z;
    p.bands = bands;
    p.band_oh = ceil_div(a.oh, bands);
    dim3 grid((unsigned)ceil_div(a.ow, p.cols * SUB * NC), (unsigned)ceil_div(a.c_p, CH), (unsigned)p.zsplit);
    if (grid.y > 65535u || grid.z > 65535u || smem > 200 * 1024) return 1;
    p.pool_write = a.pool_sum != nullptr && a.pool_parts > 1 && (int)grid.x * bands <= a.pool_parts;
    if (launch_pdl(dwconv_tma_kernel<T, K, S, CH, NC>, grid, dim3(32 * (p.cols + 1)), (size_t)smem, st, p) != cudaSuccess) {
        (void)cudaGetLastError();
        return 1;
    }
    return 0;
}

template <typename T, int K, int S>
static int dwconv_tma_launch_ks(const EdeDwConv& a, cudaStream_t st, int* so) {
    // 32-lane chunks when they cut the idle lanes of the last chunk by a quarter or more (c_p = 32, 96, ...)
    // two output columns per thread except where the window would not fit the registers (5x5 stride 2) or the map
    // is so narrow that half of the column pairs would be padding
    constexpr int NC = (K >= 5 && S == 2) ? 1 : 2;
    static const char* nc_env = getenv("EDE_DW_NC");          // debug / tests: force one column per thread
    const bool one = NC == 1 || (nc_env && atoi(nc_env) == 1);
    const bool ch32 = 4 * round_up(a.c_p, 32) <= 3 * round_up(a.c_p, 64);
    if (one) return ch32 ? dwconv_tma_launch_ch<T, K, S, 32, 1>(a, st, so) : dwconv_tma_launch_ch<T, K, S, 64, 1>(a, st, so);
    // 3x3 stride 1: four columns per thread on the widest maps (the window is small enough for the registers)
    constexpr int NC4 = K == 3 ? 4 : NC;
    const bool four = K == 3 && S == 1 && a.ow >= 128 && !(nc_env && atoi(nc_env) == 2);   // measured: pays on 114x152 only
    if (four) return ch32 ? dwconv_tma_launch_ch<T, K, S, 32, NC4>(a, st, so) : dwconv_tma_launch_ch<T, K, S, 64, NC4>(a, st, so);
    return ch32 ? dwconv_tma_launch_ch<T, K, S, 32, NC>(a, st, so) : dwconv_tma_launch_ch<T, K, S, 64, NC>(a, st, so);
}

// returns 0 when the TMA kernel was launched, 1 when this (k, s) / device has no TMA path (caller falls back);
// with `strips_out` nothing is launched: it receives the number of column strips (grid.x) the kernel would use
int dwconv_tma_launch(const EdeDwConv& a, cudaStream_t st, int* strips_out) {
    if (!strips_out && (reinterpret_cast<uintptr_t>(a.x) & 15) != 0) return 1;
    const int ks = a.k * 10 + a.s;
    int rc = 1;
    EDE_DISPATCH_DTYPE_DW(a.dtype, T, {
        if (ks == 31) rc = dwconv_tma_launch_ks<T, 3, 1>(a, st, strips_out);
        else if (ks == 32) rc = dwconv_tma_launch_ks<T, 3, 2>(a, st, strips_out);
        else if (ks == 51) rc = dwconv_tma_launch_ks<T, 5, 1>(a, st, strips_out);
        else if (ks == 52) rc = dwconv_tma_launch_ks<T, 5, 2>(a, st, strips_out);
    });
    return rc;
}

}  // namespace ede
