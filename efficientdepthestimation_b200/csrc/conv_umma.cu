// conv_umma.cu -- implicit-GEMM convolution on the 5th-gen tensor cores (tcgen05 / TMEM / TMA).
//
// GEMM view of a convolution over NHWC bf16 activations:
//     D[pixel, cout] = sum_{tap, cin} A[pixel (+tap shift), cin] * W[tap][cout][cin]
// UMMA M = 128 output pixels (TMEM lanes), UMMA N = n_pad output channels (TMEM columns,
// fp32), K = 16 bf16 channels per tcgen05.mma, accumulated over channel chunks and taps.
//
// Two ways of feeding the A operand, both through TMA tiled tensor maps with zero OOB fill
// (so zero padding, ragged edges and channel tails cost nothing and are never materialised):
//   a_mode 1 ("per-tap"):  every (tap, channel-chunk) step TMA-loads a box of 128 output
//       pixels (bw x bh, element strides = conv stride) at the tap-shifted coordinate.
//       Works for every stride/padding; A is re-read kh*kw times from L2.
//   a_mode 0 ("resident"): stride-1 k x k convs.  One TMA box loads the (bh+kh-1) x (bw+kw-1)
//       input halo tile of a channel chunk ONCE into smem as a flat array of pixel rows
//       (pitch = bw+kw-1); tap (r,s) is then just the UMMA smem descriptor start address
//       advanced by (r*pitch+s) rows, because the K-major swizzled canonical layout is a flat
//       array of rows.  Output rows that fall into the halo columns are computed and dropped.
//       A is read once per channel chunk instead of kh*kw times.
//
// Warp roles (192 threads): warp 0 = TMA producer (one lane), warp 1 = TMEM alloc + MMA issuer
// (one lane), warps 2..5 = epilogue (tcgen05.ld -> folded BN scale/shift, residual, activation
// -> 16-byte NHWC stores into a channel slice of the destination).
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include <stddef.h>

#include <algorithm>
#include <cmath>
#include <unordered_map>

#include "common.cuh"
#include "tma.cuh"

namespace ede {

// ------------------------------------------------------------------ tcgen05 PTX wrappers (TMA / mbarrier ones: tma.cuh)
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// descriptors passed as (lo, hi) words: the hi word and the constant lo bits are loop invariant, only the
// 14-bit start-address field changes per MMA, so the issuing thread spends ~6 instructions per tcgen05.mma
__device__ __forceinline__ void tc_mma_bf16_w(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi,
                                              uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t"
        "}" ::"r"(d_tmem),
        "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate));
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------ kernel parameters
constexpr int MAX_STAGES = 12;
constexpr int MAX_ACC_STAGES = 4;           // TMEM accumulator stages of conv_npack_kernel (2 everywhere else)
constexpr int EPI_WARPS = 8;                 // two warps per TMEM lane quarter, interleaved over 16-column chunks
constexpr int EPI_THREADS = 32 * EPI_WARPS;
constexpr int NUM_THREADS = 64 + EPI_THREADS;   // warp 0 TMA, warp 1 MMA, warps 2.. epilogue

struct alignas(64) UmmaParams {
    CUtensorMap tm_a;     // activations: (C, W, H, N) bf16
    CUtensorMap tm_b;     // weights: (Cin_p, Cout_p, taps) bf16
    int oh, ow, tiles_w, tiles_h, n_img, n_tiles, total_tiles;
    int a_mode, ck_bytes, k_bytes, n_chunks, kh, kw, sh, sw, pt, pl, w_passes;
    int bw, bh, pitch, box_bytes;
    int m_tiles, n_pad, acc_cols, swap;      // swap: weights are the M-side operand, 256 (240 packed) pixels on N
    int pack_g, kw_steps, n_cols;            // pack_g taps share one MMA (M rows = [tap j][128/pack_g channels]); kw_steps = ceil(kw/pack_g); n_cols = UMMA N of the swapped form
    int a_stages, b_stages, a_stage_bytes, b_stage_bytes, b_tx_bytes;
    int b_group;          // taps per weight stage (one TMA box, one wait, one commit); > 1 only swapped + resident
    // ---- N-packed pixels-on-M form (conv_npack_kernel)
    int npack, slot;      // npack = G > 0: the G taps of a kernel row sit on the N axis (column j*slot + c), smem pixel
                          // pitch 32 = one epilogue warp per window row; the epilogue adds the slots lane-shifted
    int b_resident;       // 1: the whole weight set sits in the b_stages (loaded once, never released)
    int acc_stages;       // accumulator stages in TMEM (acc_stages * acc_cols <= 512)
    int tiles_per_cta;    // a CTA walks that many CONSECUTIVE tiles (coordinates advance without divisions)
    int w_img_taps;       // > 0: the weight tensor map holds one set of that many taps per image (SE gate folded in)
    int layout_type, sbo_bytes, tmem_cols, flat;
    uint32_t idesc;
    void* y;
    const void* res;
    const float* scale;
    const float* shift;
    int y_cs, y_co, res_cs, res_co, cout_p, act, act_c_end, y_dtype;
    unsigned long long* prof;   // debug: 8 u64 per CTA (EDE_UMMA_PROF=1), else nullptr
};

__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ void epi_bar_sync() {      // named barrier 1 over the epilogue threads
    asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory");
}

// constant part of a K-major swizzled UMMA shared-memory descriptor (start address field = 0)
__device__ __forceinline__ uint64_t smem_desc_const(const UmmaParams& p) {
    uint64_t d = (uint64_t)1 << 16;                          // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)((uint32_t)p.sbo_bytes >> 4) << 32;       // stride between 8-row groups
    d |= (uint64_t)1 << 46;                                  // descriptor version (Blackwell)
    d |= (uint64_t)p.layout_type << 61;
    return d;
}

struct TileCoord {
    int ntile, img, oh0, ow0;
};

__device__ __forceinline__ TileCoord decode_tile(const UmmaParams& p, int tile) {
    TileCoord t;
    const int per_ntile = p.n_img * p.tiles_h * p.tiles_w;
    t.ntile = tile / per_ntile;
    int r = tile - t.ntile * per_ntile;
    const int per_img = p.tiles_h * p.tiles_w;
    t.img = r / per_img;
    r -= t.img * per_img;
    const int th = r / p.tiles_w, tw = r - th * p.tiles_w;
    const int tile_w_pix = p.a_mode == 0 ? p.bw : p.bw * p.m_tiles;   // output columns covered by one tile
    t.ow0 = tw * tile_w_pix;
    t.oh0 = th * p.bh;
    return t;
}

// the tile after `t` in raster order (conv_npack_kernel: consecutive tiles per CTA, one n-tile, resident pixel tile)
__device__ __forceinline__ void next_tile_contig(const UmmaParams& p, TileCoord& t) {
    t.ow0 += p.bw;
    if (t.ow0 >= p.tiles_w * p.bw) {
        t.ow0 = 0;
        t.oh0 += p.bh;
        if (t.oh0 >= p.tiles_h * p.bh) { t.oh0 = 0; ++t.img; }
    }
}

// one 8-channel group of the epilogue: folded BN, residual, activation, typed store.  ACT and RES are compile
// time (the tile loop dispatches once per tile), so the hot loop carries no per-element branches.
template <int ACT>
__device__ __forceinline__ float act_ct(float v) {
    if (ACT == EDE_ACT_RELU) return fmaxf(v, 0.f);
    if (ACT == EDE_ACT_SWISH) return swish_fast(v);
    return v;
}
#ifndef EDE_UNIFORM_ISSUER
#define EDE_UNIFORM_ISSUER 1       // MMA issuer: whole warp walks the loop, one elected lane issues (else lane 0 alone)
#endif
#ifndef EDE_UNIFORM_PRODUCER
#define EDE_UNIFORM_PRODUCER 0     // same choice for the TMA producer warp
#endif
constexpr int ACT_RUNTIME = 99;              // kernel variant that reads p.act at run time (sigmoid, anything else)

template <typename TY, int ACT, bool RES>
__device__ __forceinline__ void epilogue_store8(const UmmaParams& p, TY* yrow, const TY* rrow, int c, int c_local,
                                                const uint32_t* acc, const float* s_scale, const float* s_shift) {
    float v[8];
    const float4 s0 = *reinterpret_cast<const float4*>(s_scale + c_local);
    const float4 s1 = *reinterpret_cast<const float4*>(s_scale + c_local + 4);
    const float4 f0 = *reinterpret_cast<const float4*>(s_shift + c_local);
    const float4 f1 = *reinterpret_cast<const float4*>(s_shift + c_local + 4);
    v[0] = fmaf(__uint_as_float(acc[0]), s0.x, f0.x); v[1] = fmaf(__uint_as_float(acc[1]), s0.y, f0.y);
    v[2] = fmaf(__uint_as_float(acc[2]), s0.z, f0.z); v[3] = fmaf(__uint_as_float(acc[3]), s0.w, f0.w);
    v[4] = fmaf(__uint_as_float(acc[4]), s1.x, f1.x); v[5] = fmaf(__uint_as_float(acc[5]), s1.y, f1.y);
    v[6] = fmaf(__uint_as_float(acc[6]), s1.z, f1.z); v[7] = fmaf(__uint_as_float(acc[7]), s1.w, f1.w);
    if (RES) {
        float r[8];
        load8<TY>(rrow + c, r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] += r[j];
    }
    if (ACT != EDE_ACT_NONE && c < p.act_c_end) {          // act_c_end is a multiple of 8: whole groups are (in)active
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = ACT == ACT_RUNTIME ? apply_act(v[j], p.act) : act_ct<ACT>(v[j]);
    }
    store8<TY>(yrow + c, v);
}

// all 16-column chunks of one accumulator row that this warp drains (chunks half, half + 2, ...): the next
// tcgen05.ld is in flight while the current chunk is converted and stored
template <typename TY, int ACT, bool RES>
__device__ __forceinline__ void epilogue_row(const UmmaParams& p, uint32_t t_row, int nch, int half, bool valid,
                                             long long pix, int cbase, const float* s_scale, const float* s_shift) {
    TY* const yrow = reinterpret_cast<TY*>(p.y) + pix * p.y_cs + p.y_co;
    const TY* const rrow = RES ? reinterpret_cast<const TY*>(p.res) + pix * p.res_cs + p.res_co : nullptr;
    uint32_t acc0[16], acc1[16];
    int ci = half;
    if (ci < nch) tc_ld16(t_row + (uint32_t)(ci << 4), acc0);
    while (ci < nch) {
        tc_ld_wait();
        const int cn = ci + EPI_WARPS / 4;
        if (cn < nch) tc_ld16(t_row + (uint32_t)(cn << 4), acc1);
        if (valid) {
            const int c0l = ci << 4, c = cbase + c0l;
            if (c < p.cout_p) epilogue_store8<TY, ACT, RES>(p, yrow, rrow, c, c0l, acc0, s_scale, s_shift);
            if (c + 8 < p.cout_p) epilogue_store8<TY, ACT, RES>(p, yrow, rrow, c + 8, c0l + 8, acc0 + 8, s_scale, s_shift);
        }
        if (cn >= nch) break;
        tc_ld_wait();
        ci = cn + EPI_WARPS / 4;
        if (ci < nch) tc_ld16(t_row + (uint32_t)(ci << 4), acc0);
        if (valid) {
            const int c0l = cn << 4, c = cbase + c0l;
            if (c < p.cout_p) epilogue_store8<TY, ACT, RES>(p, yrow, rrow, c, c0l, acc1, s_scale, s_shift);
            if (c + 8 < p.cout_p) epilogue_store8<TY, ACT, RES>(p, yrow, rrow, c + 8, c0l + 8, acc1 + 8, s_scale, s_shift);
        }
    }
}

// Epilogue of the operand-swapped mode: TMEM lane = [tap slot j][output channel], column = pixel of the window.
// The four warps of a half-set (one per TMEM lane quarter) stage 16 window positions at a time in a shared
// [16][128] fp32 tile -- the warp that holds tap slot j reads its columns j further right, so row k of the tile
// holds, for every j, the partial sum that belongs to output position c0 + k -- and after one named barrier all
// 128 threads finish 4-channel units of one pixel: sum over the G tap slots, folded BN, residual, activation,
// 8/16-byte stores that are contiguous along the channels of a pixel.  Two tiles alternate (one barrier/chunk).
// A thread always finishes the same 4 channels (scale/shift live in registers) and finds its pixel through a
// 256-entry window-position table built once per kernel: the finishing loop carries no divisions.
constexpr int STG_PITCH = 132;               // floats per staged window position (128 lanes + 4: no bank conflicts)
constexpr int STG_FLOATS = 16 * STG_PITCH;

template <typename TY, int ACT, bool RES, int G>
__device__ __forceinline__ void epilogue_swapped(const UmmaParams& p, const TileCoord& t, uint32_t t_base, int quarter,
                                                 int half, int lane, int tid_h, float* s_stage, const float* s_scale,
                                                 const float* s_shift, const uint16_t* s_lut, long long* ph) {
    constexpr int CG = 128 / G;                              // channels per tap slot (128 / 64 / 32)
    constexpr int UPP = CG / 4;                              // 4-channel units per staged position: 32 / 16 / 8
    constexpr int UPT = 16 * UPP / 128;                      // units per thread and chunk: 4 / 2 / 1
    constexpr int PSTEP = 128 / UPP;                         // positions between two units of a thread: 4 / 8 / 16
    const int j = (quarter * 32) / CG;                       // tap slot of this warp's 32 lanes
    const bool live = quarter * 32 - j * CG < p.cout_p;      // do these lanes hold real output channels?
    const int r = quarter * 32 + lane;
    const int c = (tid_h & (UPP - 1)) << 2;                  // this thread's four channels, the same for every unit
    const int pj0 = tid_h / UPP;
    const bool c_live = c < p.cout_p;
    const float4 sc = *reinterpret_cast<const float4*>(s_scale + c), sf = *reinterpret_cast<const float4*>(s_shift + c);
    const bool act_on = ACT != EDE_ACT_NONE && c < p.act_c_end;        // act_c_end is a multiple of 8
    const float* const src0 = s_stage + pj0 * STG_PITCH + c;
    const long long img_row0 = (long long)t.img * p.oh;
    const int step = 16 * (EPI_WARPS / 4);
    // where the units of chunk c0 go; the residual is fetched one chunk ahead so that its latency hides behind the
    // staging barrier and the previous chunk's stores
    long long pix[UPT];
    uint32_t okm = 0;
    float rs[UPT][4];
    auto locate = [&](int c0) {
        okm = 0;
#pragma unroll
        for (int k = 0; k < UPT; ++k) {
            const uint32_t e = c_live && c0 < p.n_cols ? s_lut[c0 + pj0 + k * PSTEP] : 0xffffu;   // position -> tile (row, col)
            const int oh = t.oh0 + (int)(e >> 8), ow = t.ow0 + (int)(e & 255u);
            const bool in = e != 0xffffu && oh < p.oh && ow < p.ow;
            okm |= in ? 1u << k : 0u;
            pix[k] = (img_row0 + oh) * p.ow + ow;
            if (RES && in) load4<TY>(reinterpret_cast<const TY*>(p.res) + pix[k] * p.res_cs + p.res_co, c, rs[k]);
        }
    };
    locate(half * 16);
    uint32_t acc[16];
    if (live) {                                              // chunk 0 -> staging tile 0
        tc_ld16(t_base + (uint32_t)(half * 16 + j), acc);
        tc_ld_wait();
#pragma unroll
        for (int k = 0; k < 16; ++k) s_stage[k * STG_PITCH + r] = __uint_as_float(acc[k]);
    }
    int it = 0;
    for (int c0 = half * 16; c0 < p.n_cols; c0 += step, ++it) {
        const float* const buf = src0 + (it & 1) * STG_FLOATS;
        const bool next = live && c0 + step < p.n_cols;
        if (next) tc_ld16(t_base + (uint32_t)(c0 + step + j), acc);      // in flight across the barrier and phase 2
        const long long c_a = ph ? clock64() : 0;
        if (half == 0) asm volatile("bar.sync 2, 128;" ::: "memory");
        else asm volatile("bar.sync 3, 128;" ::: "memory");
        const long long c_b = ph ? clock64() : 0;
        {
#pragma unroll
            for (int k = 0; k < UPT; ++k) {
                if (!(okm >> k & 1u)) continue;
                const float* src = buf + k * PSTEP * STG_PITCH;
                float4 a4 = *reinterpret_cast<const float4*>(src);
#pragma unroll
                for (int jj = 1; jj < G; ++jj) {
                    const float4 b4 = *reinterpret_cast<const float4*>(src + jj * CG);
                    a4.x += b4.x; a4.y += b4.y; a4.z += b4.z; a4.w += b4.w;
                }
                float v[4] = {fmaf(a4.x, sc.x, sf.x), fmaf(a4.y, sc.y, sf.y), fmaf(a4.z, sc.z, sf.z), fmaf(a4.w, sc.w, sf.w)};
                if (RES) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) v[q] += rs[k][q];
                }
                if (act_on) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) v[q] = ACT == ACT_RUNTIME ? apply_act(v[q], p.act) : act_ct<ACT>(v[q]);
                }
                store4<TY>(reinterpret_cast<TY*>(p.y) + pix[k] * p.y_cs + p.y_co, c, v);
            }
            locate(c0 + step);
        }
        const long long c_c = ph ? clock64() : 0;
        if (ph) { ph[0] += c_b - c_a; ph[1] += c_c - c_b; }
        if (next) {                                          // chunk it+1 -> the other staging tile (its readers of
            tc_ld_wait();                                    // chunk it-1 all passed this iteration's barrier)
            float* const nb = s_stage + ((it + 1) & 1) * STG_FLOATS;
#pragma unroll
            for (int k = 0; k < 16; ++k) nb[k * STG_PITCH + r] = __uint_as_float(acc[k]);
        }
        if (ph) ph[2] += clock64() - c_c;
    }
}

// Persistent, warp-specialised implicit-GEMM convolution.  One CTA per SM walks the tile list
// (tile = n-tile x image x output patch) with a static stride; TMEM holds TWO accumulator stages so the
// epilogue of tile i overlaps the MMAs of tile i+1; the operand rings run on across tile boundaries.
template <typename TY, int ACT, bool RES>
__global__ void __launch_bounds__(NUM_THREADS, 1) conv_umma_kernel(const __grid_constant__ UmmaParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = smem_base;
    const uint32_t b_base = a_base + (uint32_t)p.a_stages * p.a_stage_bytes;
    const uint32_t bar_base = b_base + (uint32_t)p.b_stages * p.b_stage_bytes;
    // barriers: a_full[8] a_empty[8] b_full[8] b_empty[8] acc_full[2] acc_empty[2], TMEM base slot, scale/shift
    auto a_full = [&](int s) { return bar_base + 8u * s; };
    auto a_empty = [&](int s) { return bar_base + 8u * (MAX_STAGES + s); };
    auto b_full = [&](int s) { return bar_base + 8u * (2 * MAX_STAGES + s); };
    auto b_empty = [&](int s) { return bar_base + 8u * (3 * MAX_STAGES + s); };
    auto acc_full = [&](int s) { return bar_base + 8u * (4 * MAX_STAGES + s); };
    auto acc_empty = [&](int s) { return bar_base + 8u * (4 * MAX_STAGES + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (4 * MAX_STAGES + 4);
    float* const s_scale = reinterpret_cast<float*>(smem_raw + (bar_base - smem_u32(smem_raw)) + 512);
    float* const s_shift = s_scale + 256;
    // swapped mode (n_pad = 128 uses half of the scale area): window position -> (tile row << 8 | tile column)
    uint16_t* const s_lut = reinterpret_cast<uint16_t*>(s_scale + 128);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sp_taps = p.kh * p.kw_steps;        // spatial steps (kw_steps = kw unless taps are packed on M)
    const int taps = sp_taps * p.w_passes;        // weight tiles per channel chunk ([w_hi taps ; w_lo taps])
    unsigned long long prof_t0 = 0;
    if (p.prof && threadIdx.x == 0) prof_t0 = gtime();
    pdl_trigger();

    if (threadIdx.x == 0) {
        for (int s = 0; s < MAX_STAGES; ++s) {
            mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1);
            mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1);
        }
        mbar_init(acc_full(0), 1); mbar_init(acc_full(1), 1);
        mbar_init(acc_empty(0), EPI_WARPS); mbar_init(acc_empty(1), EPI_WARPS);   // one arrival per epilogue warp
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                     "r"((uint32_t)p.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (p.swap) {
        for (int f = threadIdx.x; f < 256; f += NUM_THREADS) {
            int dy, dx;
            bool ok = true;
            if (p.a_mode == 0) { dy = f / p.pitch; dx = f - dy * p.pitch; ok = dx < p.bw && dy < p.bh; }
            else { const int g = f & 127; dy = g / p.bw; dx = (f >> 7) * p.bw + g - dy * p.bw; }
            s_lut[f] = ok ? (uint16_t)((dy << 8) | dx) : (uint16_t)0xffffu;
        }
    }
    if (p.a_mode == 0) {
        // The shifted UMMA windows of the resident mode read past the TMA box (into the tail of the stage, never
        // written by TMA).  Those rows only meet invalid output positions or, with packed taps, the zero weight
        // rows of taps >= kw -- but 0 * stale NaN is NaN, so the tails are zeroed once and stay zero.
        const int tail0 = p.box_bytes & ~15, tail_n = (p.a_stage_bytes - tail0) >> 4;
        for (int st = 0; st < p.a_stages; ++st) {
            uint8_t* base = smem_raw + (a_base - smem_u32(smem_raw)) + (size_t)st * p.a_stage_bytes + tail0;
            for (int i = threadIdx.x; i < tail_n; i += NUM_THREADS) reinterpret_cast<uint4*>(base)[i] = make_uint4(0, 0, 0, 0);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    pdl_wait();            // the prologue above overlapped the previous kernel's tail; its outputs are visible now
    uint32_t tmem_addr;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_addr) : "r"(tmem_slot) : "memory");

    if (warp == 0) {
        // ===================== TMA producer: ONE thread walks the loop (no warp-collective ops on the hot path; a
        // warp-uniform loop with an elected issuing lane measured 7 % slower on the 5x5 decoder convs)
#if EDE_UNIFORM_PRODUCER
        {
            const bool leader = elect_one();
#else
        if (lane == 0) {
            const bool leader = true;
#endif
            int a_st = 0, a_ph = 0, b_st = 0, b_ph = 0;
            const int ck = p.ck_bytes >> 1;
            const int a_mode = p.a_mode, n_chunks = p.n_chunks, m_tiles = p.m_tiles, kws = p.kw_steps;
            const int a_stages = p.a_stages, b_stages = p.b_stages;
            const uint32_t a_stage_bytes = (uint32_t)p.a_stage_bytes, b_stage_bytes = (uint32_t)p.b_stage_bytes;
            const int b_group = p.b_group;
            const uint32_t box_bytes = (uint32_t)p.box_bytes, b_tx = (uint32_t)(p.b_tx_bytes * b_group);
            const uint32_t a_tx = (uint32_t)(m_tiles * 128 * p.ck_bytes), m_bytes = (uint32_t)(128 * p.ck_bytes);
            const bool prof = p.prof != nullptr;
            long long w_empty = 0;
            for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
                const TileCoord t = decode_tile(p, tile);
                const int wrow = t.ntile * p.n_pad;
                for (int chunk = 0; chunk < n_chunks; ++chunk) {
                    const int cch = chunk * ck;
                    if (a_mode == 0) {
                        mbar_wait(a_empty(a_st), a_ph ^ 1);
                        if (leader) {
                            mbar_expect_tx(a_full(a_st), box_bytes);
                            tma_load_4d(a_base + a_st * a_stage_bytes, &p.tm_a, a_full(a_st), cch, t.ow0 - p.pl,
                                        t.oh0 - p.pt, t.img);
                        }
                        if (++a_st == a_stages) { a_st = 0; a_ph ^= 1; }
                    }
                    int ky = 0, kx = 0;
                    for (int tap = 0; tap < taps; tap += b_group) {
                        if (a_mode != 0) {
                            if (tap == sp_taps) { ky = 0; kx = 0; }
                            mbar_wait(a_empty(a_st), a_ph ^ 1);
                            if (leader) {
                                mbar_expect_tx(a_full(a_st), a_tx);
                                for (int m = 0; m < m_tiles; ++m)
                                    tma_load_4d(a_base + a_st * a_stage_bytes + m * m_bytes, &p.tm_a, a_full(a_st), cch,
                                                (t.ow0 + m * p.bw) * p.sw - p.pl + kx, t.oh0 * p.sh - p.pt + ky, t.img);
                            }
                            if (++a_st == a_stages) { a_st = 0; a_ph ^= 1; }
                            if (++kx == kws) { kx = 0; ++ky; }
                        }
                        const long long c0 = prof ? clock64() : 0;
                        mbar_wait(b_empty(b_st), b_ph ^ 1);
                        if (prof) w_empty += clock64() - c0;
                        if (leader) {
                            mbar_expect_tx(b_full(b_st), b_tx);
                            tma_load_3d(b_base + b_st * b_stage_bytes, &p.tm_b, b_full(b_st), cch, wrow, tap + t.img * p.w_img_taps);
                        }
                        if (++b_st == b_stages) { b_st = 0; b_ph ^= 1; }
#if EDE_UNIFORM_PRODUCER
                        __syncwarp();
#endif
                    }
                }
            }
            if (prof && lane == 0) p.prof[(size_t)blockIdx.x * 16 + 5] = (unsigned long long)w_empty;
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: ONE thread; per (tap, chunk) step it waits for the operands, issues the
        // K steps that hold real channels and commits the stage back -- kept to a few dozen instructions because
        // narrow layers have only 1..2 MMAs (128..256 tensor-pipe cycles) per step
#if EDE_UNIFORM_ISSUER
        {
            const bool leader = elect_one();
#else
        if (lane == 0) {
            const bool leader = true;
#endif
            const int a_mode = p.a_mode, m_tiles = p.m_tiles, n_pad = p.n_pad, n_chunks = p.n_chunks, swap = p.swap;
            const int a_stages = p.a_stages, b_stages = p.b_stages, kws = p.kw_steps;
            const uint32_t ck16 = (uint32_t)p.ck_bytes >> 4;            // (bytes >> 4) units of the address field
            const uint32_t a_stage16 = (uint32_t)p.a_stage_bytes >> 4, b_stage16 = (uint32_t)p.b_stage_bytes >> 4;
            const uint32_t m_step16 = 128u * ck16;                       // 128 pixel rows
            const int ksteps = p.ck_bytes >> 5;                          // 32 bytes of K (16 bf16) per tcgen05.mma
            const uint32_t idesc = p.idesc;
            const uint64_t dconst = smem_desc_const(p);
            const uint32_t desc_hi = (uint32_t)(dconst >> 32);
            const uint32_t a_base16 = (uint32_t)dconst + (a_base >> 4), b_base16 = (uint32_t)dconst + (b_base >> 4);
            const uint32_t kx_step16 = a_mode == 0 ? (uint32_t)p.pack_g * ck16 : 0u;       // next tap of a kernel row
            const uint32_t ky_step16 = a_mode == 0 ? (uint32_t)(p.pitch - kws * p.pack_g) * ck16 : 0u;   // row wrap
            const int k_bytes = p.k_bytes, ck_bytes = p.ck_bytes;
            int a_st = 0, a_ph = 0, b_st = 0, b_ph = 0, acc_st = 0, acc_ph = 0;
            uint32_t a_cur = a_base16, b_cur = b_base16;                 // descriptor low words of the current stages
            unsigned long long t_start = 0;
            const bool prof = p.prof != nullptr;
            long long w_ops = 0, w_acc = 0;
            const long long c_start = prof ? clock64() : 0;
            if (prof) t_start = gtime();
            if (a_mode == 0) {
                // ---- resident pixel tile (either orientation): a weight stage holds b_group taps, so one wait and one
                // commit cover b_group * ks MMAs (the per-step bookkeeping of this thread, not the tensor pipe, was
                // the limit of the 5x5 decoder convs: tools/micro/mma_contend.cu, profiles/r01_issue_overhead.md)
                const int T = p.b_group;
                const uint32_t tap16 = (uint32_t)p.b_tx_bytes >> 4;
                for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
                    const long long ca = prof ? clock64() : 0;
                    mbar_wait(acc_empty(acc_st), acc_ph ^ 1);
                    if (prof) w_acc += clock64() - ca;
                    tc_fence_after();
                    const uint32_t d_base = tmem_addr + (uint32_t)(acc_st * p.acc_cols);
                    uint32_t accum = 0;
                    for (int chunk = 0; chunk < n_chunks; ++chunk) {
                        const int ks = min(ksteps, (k_bytes - chunk * ck_bytes + 31) >> 5);
                        mbar_wait(a_full(a_st), a_ph);
                        uint32_t shift16 = 0;
                        int kx = 0, tap_in_pass = 0;
                        for (int tap0 = 0; tap0 < taps; tap0 += T) {
                            const long long c0 = prof ? clock64() : 0;
                            mbar_wait(b_full(b_st), b_ph);
                            if (prof) w_ops += clock64() - c0;
                            tc_fence_after();
                            const int nt = min(T, taps - tap0);
                            uint32_t bt = b_cur;
                            for (int t = 0; t < nt; ++t) {
                                const uint32_t a0 = a_cur + shift16;
                                if (leader) {
                                    if (swap) {
                                        // D[cout lanes][pixel columns] += W[128 x 16] . X[n_cols x 16]^T
                                        if (ks == 4) {
                                            tc_mma_bf16_w(d_base, bt, a0, desc_hi, idesc, accum);
                                            tc_mma_bf16_w(d_base, bt + 2u, a0 + 2u, desc_hi, idesc, 1u);
                                            tc_mma_bf16_w(d_base, bt + 4u, a0 + 4u, desc_hi, idesc, 1u);
                                            tc_mma_bf16_w(d_base, bt + 6u, a0 + 6u, desc_hi, idesc, 1u);
                                        } else {
                                            tc_mma_bf16_w(d_base, bt, a0, desc_hi, idesc, accum);
                                            if (ks > 1) tc_mma_bf16_w(d_base, bt + 2u, a0 + 2u, desc_hi, idesc, 1u);
                                            if (ks > 2) tc_mma_bf16_w(d_base, bt + 4u, a0 + 4u, desc_hi, idesc, 1u);
                                        }
                                    } else {
                                        uint32_t am = a0, dm = d_base;
                                        for (int m = 0; m < m_tiles; ++m) {
                                            if (ks == 4) {
                                                tc_mma_bf16_w(dm, am, bt, desc_hi, idesc, accum);
                                                tc_mma_bf16_w(dm, am + 2u, bt + 2u, desc_hi, idesc, 1u);
                                                tc_mma_bf16_w(dm, am + 4u, bt + 4u, desc_hi, idesc, 1u);
                                                tc_mma_bf16_w(dm, am + 6u, bt + 6u, desc_hi, idesc, 1u);
                                            } else {
                                                tc_mma_bf16_w(dm, am, bt, desc_hi, idesc, accum);
                                                if (ks > 1) tc_mma_bf16_w(dm, am + 2u, bt + 2u, desc_hi, idesc, 1u);
                                                if (ks > 2) tc_mma_bf16_w(dm, am + 4u, bt + 4u, desc_hi, idesc, 1u);
                                            }
                                            am += m_step16;
                                            dm += (uint32_t)n_pad;
                                        }
                                    }
                                }
                                accum = 1;
                                bt += tap16;
                                shift16 += kx_step16;
                                if (++kx == kws) { kx = 0; shift16 += ky_step16; }
                                if (++tap_in_pass == sp_taps) { tap_in_pass = 0; shift16 = 0; kx = 0; }
                            }
                            if (leader) tc_commit(b_empty(b_st));
#if EDE_UNIFORM_ISSUER
                            __syncwarp();
#endif
                            b_cur += b_stage16;
                            if (++b_st == b_stages) { b_st = 0; b_ph ^= 1; b_cur = b_base16; }
                        }
                        if (leader) tc_commit(a_empty(a_st));
                        a_cur += a_stage16;
                        if (++a_st == a_stages) { a_st = 0; a_ph ^= 1; a_cur = a_base16; }
                    }
                    if (leader) tc_commit(acc_full(acc_st));
                    if (++acc_st == 2) { acc_st = 0; acc_ph ^= 1; }
                }
            } else {
            const int pass_group = p.b_group;
            const uint32_t pass_tap16 = (uint32_t)p.b_tx_bytes >> 4;
            for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
                const long long ca = prof ? clock64() : 0;
                mbar_wait(acc_empty(acc_st), acc_ph ^ 1);                // epilogue has drained this accumulator stage
                if (prof) w_acc += clock64() - ca;
                tc_fence_after();
                const uint32_t d_base = tmem_addr + (uint32_t)(acc_st * p.acc_cols);
                uint32_t accum = 0;
                for (int chunk = 0; chunk < n_chunks; ++chunk) {
                    // K steps that hold real channels: the tail of the last chunk is TMA zero fill, never multiplied
                    const int ks = min(ksteps, (k_bytes - chunk * ck_bytes + 31) >> 5);
                    if (a_mode == 0) mbar_wait(a_full(a_st), a_ph);
                    uint32_t shift16 = 0;
                    int kx = 0, tap_in_pass = 0;
                    // a weight stage holds pass_group tiles that share ONE pixel box: 1, or 2 = [w_hi ; w_lo] of a 1x1 conv
                    for (int tap = 0; tap < taps; tap += pass_group) {
                        const long long c0 = prof ? clock64() : 0;
                        if (a_mode != 0) mbar_wait(a_full(a_st), a_ph);
                        mbar_wait(b_full(b_st), b_ph);
                        if (prof) w_ops += clock64() - c0;
                        tc_fence_after();
                        const uint32_t a0 = a_cur + shift16;
                        if (leader) {
                            uint32_t acc_flag = accum, b0 = b_cur;
                            for (int t = 0; t < pass_group; ++t, b0 += pass_tap16, acc_flag = 1u) {
                                if (swap) {
                                    // D[cout lanes][pixel columns] += W[128 x 16] . X[n_cols x 16]^T
                                    tc_mma_bf16_w(d_base, b0, a0, desc_hi, idesc, acc_flag);
                                    if (ks > 1) tc_mma_bf16_w(d_base, b0 + 2u, a0 + 2u, desc_hi, idesc, 1u);
                                    if (ks > 2) tc_mma_bf16_w(d_base, b0 + 4u, a0 + 4u, desc_hi, idesc, 1u);
                                    if (ks > 3) tc_mma_bf16_w(d_base, b0 + 6u, a0 + 6u, desc_hi, idesc, 1u);
                                } else {
                                    uint32_t am = a0, dm = d_base;
                                    for (int m = 0; m < m_tiles; ++m) {
                                        tc_mma_bf16_w(dm, am, b0, desc_hi, idesc, acc_flag);
                                        if (ks > 1) tc_mma_bf16_w(dm, am + 2u, b0 + 2u, desc_hi, idesc, 1u);
                                        if (ks > 2) tc_mma_bf16_w(dm, am + 4u, b0 + 4u, desc_hi, idesc, 1u);
                                        if (ks > 3) tc_mma_bf16_w(dm, am + 6u, b0 + 6u, desc_hi, idesc, 1u);
                                        am += m_step16;
                                        dm += (uint32_t)n_pad;
                                    }
                                }
                            }
                            tc_commit(b_empty(b_st));
                            if (a_mode != 0) tc_commit(a_empty(a_st));
                        }
#if EDE_UNIFORM_ISSUER
                        __syncwarp();
#endif
                        accum = 1;
                        b_cur += b_stage16;
                        if (++b_st == b_stages) { b_st = 0; b_ph ^= 1; b_cur = b_base16; }
                        if (a_mode != 0) {
                            a_cur += a_stage16;
                            if (++a_st == a_stages) { a_st = 0; a_ph ^= 1; a_cur = a_base16; }
                        } else {
                            shift16 += kx_step16;                        // next tap slot of this kernel row ...
                            if (++kx == kws) { kx = 0; shift16 += ky_step16; }           // ... or the next row
                            if (++tap_in_pass == sp_taps) { tap_in_pass = 0; shift16 = 0; kx = 0; }   // w_lo pass
                        }
                    }
                    if (a_mode == 0) {
                        if (leader) tc_commit(a_empty(a_st));
                        a_cur += a_stage16;
                        if (++a_st == a_stages) { a_st = 0; a_ph ^= 1; a_cur = a_base16; }
                    }
                }
                if (leader) tc_commit(acc_full(acc_st));
                if (++acc_st == 2) { acc_st = 0; acc_ph ^= 1; }
            }
            }
            if (p.prof && lane == 0) {
                unsigned long long* q = p.prof + (size_t)blockIdx.x * 16;
                q[1] = t_start; q[2] = gtime(); q[4] = (unsigned long long)w_ops; q[8] = (unsigned long long)w_acc;
                q[12] = (unsigned long long)(clock64() - c_start);
            }
        }
    } else {
        // ===================== epilogue (warps 2..): TMEM -> BN/residual/activation -> NHWC stores ========
        const int quarter = warp & 3;                 // TMEM lanes [32q, 32q+32) are visible to warp%4 == q
        const int half = (warp - 2) >> 2;             // which interleaved set of 16-column chunks this warp drains
        const int r = quarter * 32 + lane;
        const int et = threadIdx.x - 64;              // 0..EPI_THREADS-1
        int acc_st = 0, acc_ph = 0, n_done = 0;
        long long w_full = 0, ph_clk[3] = {0, 0, 0};
        for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
            const TileCoord t = decode_tile(p, tile);
            if (!p.swap || n_done == 0) {             // the swapped form has one n-tile: loaded once
                epi_bar_sync();                       // everyone is done with the previous tile's scale/shift
                for (int e = et; e < p.n_pad; e += EPI_THREADS) {
                    const int c = t.ntile * p.n_pad + e;
                    const bool in = c < p.cout_p;
                    s_scale[e] = in ? __ldg(p.scale + c) : 0.f;
                    s_shift[e] = in ? __ldg(p.shift + c) : 0.f;
                }
                epi_bar_sync();
            }
            const long long c0 = p.prof ? clock64() : 0;
            mbar_wait(acc_full(acc_st), acc_ph);
            if (p.prof) w_full += clock64() - c0;
            tc_fence_after();
            const uint32_t t_base = tmem_addr + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc_st * p.acc_cols);
            if (p.swap) {
                // accumulator lane = [tap slot][output channel], column = pixel of the window
                float* const s_stage = s_shift + 256 + half * (2 * STG_FLOATS);
                const int tid_h = (warp - 2 - 4 * half) * 32 + lane;
                const uint32_t t_col = t_base;
                long long* const phc = p.prof ? ph_clk : nullptr;
                if (p.pack_g == 4)
                    epilogue_swapped<TY, ACT, RES, 4>(p, t, t_col, quarter, half, lane, tid_h, s_stage, s_scale, s_shift, s_lut, phc);
                else if (p.pack_g == 2)
                    epilogue_swapped<TY, ACT, RES, 2>(p, t, t_col, quarter, half, lane, tid_h, s_stage, s_scale, s_shift, s_lut, phc);
                else
                    epilogue_swapped<TY, ACT, RES, 1>(p, t, t_col, quarter, half, lane, tid_h, s_stage, s_scale, s_shift, s_lut, phc);
            } else {
                const int nch = p.n_pad >> 4;         // 16-column chunks of one accumulator row (n_pad % 16 == 0)
                for (int m = 0; m < p.m_tiles; ++m) {
                    // which output pixel does accumulator row (m, r) hold?
                    int oh, ow;
                    bool valid;
                    if (p.a_mode == 0) {
                        const int f = m * 128 + r;
                        const int fr = f / p.pitch, fc = f - fr * p.pitch;
                        oh = t.oh0 + fr; ow = t.ow0 + fc;
                        valid = fc < p.bw && fr < p.bh;
                    } else {
                        const int fr = r / p.bw, fc = r - fr * p.bw;
                        oh = t.oh0 + fr; ow = t.ow0 + m * p.bw + fc;
                        valid = true;
                    }
                    valid = valid && oh < p.oh && ow < p.ow;
                    const long long pix = ((long long)t.img * p.oh + oh) * p.ow + ow;
                    const uint32_t t_row = t_base + (uint32_t)(m * p.n_pad);
                    const int cbase = t.ntile * p.n_pad;
                    epilogue_row<TY, ACT, RES>(p, t_row, nch, half, valid, pix, cbase, s_scale, s_shift);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty(acc_st));            // this warp has drained its TMEM lanes
            if (++acc_st == 2) { acc_st = 0; acc_ph ^= 1; }
            ++n_done;
        }
        if (p.prof && et == 0) {
            unsigned long long* q = p.prof + (size_t)blockIdx.x * 16;
            q[3] = gtime(); q[6] = (unsigned long long)w_full; q[7] = (unsigned long long)n_done;
            q[9] = (unsigned long long)ph_clk[0]; q[10] = (unsigned long long)ph_clk[1]; q[11] = (unsigned long long)ph_clk[2];
        }
    }

    tc_fence_before();
    __syncthreads();
    if (p.prof && threadIdx.x == 0) p.prof[(size_t)blockIdx.x * 16] = prof_t0;
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_addr),
                     "r"((uint32_t)p.tmem_cols)
                     : "memory");
    }
}


// =====================================================================================================================
// conv_npack_kernel -- narrow stride-1 k x k convolutions (Cout <= 80, 2 <= kw <= 5) in the "N-packed pixels-on-M" form.
//
// GEMM orientation: M = 128 window positions of the resident halo tile, N = G * slot columns where column j*slot + c
// holds kernel-row tap slot j of output channel c (slot = cout_p16, G = min(kw, 256 / slot) taps of a kernel row per
// MMA), K = input channels.  One tcgen05.mma costs max(97, N/2) cycles whatever N <= 192 is, so a kernel row costs
// ceil(kw / G) MMAs per 128 positions instead of kw (3x3 32->16: 6 instead of 18 per m-tile).  The partial sum of tap
// slot j for output position q sits in accumulator row q + j: the smem pixel pitch is fixed at 32 so that one epilogue
// warp holds exactly one window row (lanes >= bw = 33 - kw are halo columns) and the slot sum is a warp shuffle -- no
// shared-memory staging, no barrier.  The whole weight set is loaded ONCE per CTA and stays resident (the picker only
// offers the form when it fits); a CTA walks CONSECUTIVE tiles, so coordinates advance without divisions and halo
// rows of neighbouring tiles meet in L2.  Roles as in conv_umma_kernel: warp 0 = TMA producer (pixel tiles only after
// the first tile), warp 1 = MMA issuer, warps 2-9 = epilogue, two per TMEM lane quarter, which alternate over the
// (m-tile, 16-channel chunk) units of a tile.  Up to four TMEM accumulator stages.
// Measured (ENB0-HU batch 64, tools/cta_timeline.py): these warps run alone on their schedulers, so every dependent
// instruction costs its full latency -- the epilogue keeps G compile-time (a run-time G put every SHFL -> FADD pair
// behind its own uniform branch: one 2200-cycle dependent chain per unit) and carries no divisions.
template <typename TY, int ACT, bool RES, int G>
__device__ __forceinline__ void npack_tile_epilogue(const UmmaParams& p, uint32_t t_base, int half, int quarter, int lane,
                                                    const TileCoord& t, const float* s_scale, const float* s_shift) {
    const int nch = p.slot >> 4;
    const bool col_ok = lane < p.bw && t.ow0 + lane < p.ow;
    int oh = t.oh0 + quarter;
    long long pix = ((long long)t.img * p.oh + oh) * p.ow + t.ow0 + lane;
    const long long pix_step = 4ll * p.ow;
    uint32_t t_row = t_base;
    int par = 0;                                             // (m * nch + ci) % 2: which warp of the pair owns the unit
    for (int m = 0; m < p.m_tiles; ++m, t_row += (uint32_t)p.n_pad, pix += pix_step, oh += 4) {
        const bool valid = col_ok && oh < p.oh;
        TY* const yrow = reinterpret_cast<TY*>(p.y) + pix * p.y_cs + p.y_co;
        const TY* const rrow = RES ? reinterpret_cast<const TY*>(p.res) + pix * p.res_cs + p.res_co : nullptr;
        for (int ci = 0; ci < nch; ++ci, par ^= 1) {
            if (par != half) continue;
            uint32_t acc[G][16];
            const uint32_t col = t_row + (uint32_t)(ci << 4);
#pragma unroll
            for (int j = 0; j < G; ++j) tc_ld16(col + (uint32_t)(j * p.slot), acc[j]);      // all slots in flight, ONE wait
            tc_ld_wait();
            float v[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                float x = __uint_as_float(acc[0][k]);
#pragma unroll
                for (int j = 1; j < G; ++j) x += __shfl_down_sync(0xffffffffu, __uint_as_float(acc[j][k]), j);
                v[k] = x;
            }
            if (!valid) continue;
            const int c = ci << 4;
            uint32_t u[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) u[k] = __float_as_uint(v[k]);
            if (c < p.cout_p) epilogue_store8<TY, ACT, RES>(p, yrow, rrow, c, c, u, s_scale, s_shift);
            if (c + 8 < p.cout_p) epilogue_store8<TY, ACT, RES>(p, yrow, rrow, c + 8, c + 8, u + 8, s_scale, s_shift);
        }
    }
}

template <typename TY, int ACT, bool RES>
__global__ void __launch_bounds__(NUM_THREADS, 1) conv_npack_kernel(const __grid_constant__ UmmaParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = smem_base;
    const uint32_t b_base = a_base + (uint32_t)p.a_stages * p.a_stage_bytes;
    const uint32_t bar_base = b_base + (uint32_t)p.b_stages * p.b_stage_bytes;
    // barriers: a_full[12] a_empty[12] b_full[12] acc_full[4] acc_empty[4], TMEM base slot; scale/shift at +512
    auto a_full = [&](int s) { return bar_base + 8u * s; };
    auto a_empty = [&](int s) { return bar_base + 8u * (MAX_STAGES + s); };
    auto b_full = [&](int s) { return bar_base + 8u * (2 * MAX_STAGES + s); };
    auto acc_full = [&](int s) { return bar_base + 8u * (3 * MAX_STAGES + s); };
    auto acc_empty = [&](int s) { return bar_base + 8u * (3 * MAX_STAGES + MAX_ACC_STAGES + s); };
    const uint32_t tmem_slot = bar_base + 8u * (3 * MAX_STAGES + 2 * MAX_ACC_STAGES);
    float* const s_scale = reinterpret_cast<float*>(smem_raw + (bar_base - smem_u32(smem_raw)) + 512);
    float* const s_shift = s_scale + 256;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sp_taps = p.kh * p.kw_steps;        // (kernel row, tap group) steps of one weight pass
    const int taps = sp_taps * p.w_passes;
    pdl_trigger();

    if (threadIdx.x == 0) {
        for (int s = 0; s < MAX_STAGES; ++s) { mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1); mbar_init(b_full(s), 1); }
        for (int s = 0; s < MAX_ACC_STAGES; ++s) { mbar_init(acc_full(s), 1); mbar_init(acc_empty(s), EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                     "r"((uint32_t)p.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    {
        // the shifted UMMA windows read past the TMA box into the tail of the stage (never written by TMA): those rows
        // only meet halo positions or zero weight rows, but 0 * stale NaN is NaN -- zeroed once
        const int tail0 = p.box_bytes & ~15, tail_n = (p.a_stage_bytes - tail0) >> 4;
        for (int st = 0; st < p.a_stages; ++st) {
            uint8_t* base = smem_raw + (a_base - smem_u32(smem_raw)) + (size_t)st * p.a_stage_bytes + tail0;
            for (int i = threadIdx.x; i < tail_n; i += NUM_THREADS) reinterpret_cast<uint4*>(base)[i] = make_uint4(0, 0, 0, 0);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    for (int e = threadIdx.x; e < 256; e += NUM_THREADS) {   // folded-BN vectors: one n-tile, loaded once (weights, not
        const bool in = e < p.cout_p;                        // activations: no dependency on the previous kernel)
        s_scale[e] = in ? __ldg(p.scale + e) : 0.f;
        s_shift[e] = in ? __ldg(p.shift + e) : 0.f;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_addr;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_addr) : "r"(tmem_slot) : "memory");
    const int tile_first = (int)blockIdx.x * p.tiles_per_cta;
    const int tile_end = min(p.total_tiles, tile_first + p.tiles_per_cta);
    const int n_chunks = p.n_chunks;

    if (warp == 0) {
        // ===================== TMA producer (one thread): the weight set once, then one halo tile per (tile, chunk)
        if (lane == 0 && tile_first < tile_end) {
            const int ck = p.ck_bytes >> 1;
            const uint32_t b_tx = (uint32_t)(p.b_tx_bytes * p.b_group);
            for (int chunk = 0; chunk < n_chunks; ++chunk) {                  // weights do not depend on the previous kernel
                mbar_expect_tx(b_full(chunk), b_tx);
                tma_load_3d(b_base + chunk * (uint32_t)p.b_stage_bytes, &p.tm_b, b_full(chunk), chunk * ck, 0, 0);
            }
            pdl_wait();                                                       // activations of the previous kernel
            int a_st = 0, a_ph = 0;
            TileCoord t = decode_tile(p, tile_first);
            for (int tile = tile_first; tile < tile_end; ++tile) {
                for (int chunk = 0; chunk < n_chunks; ++chunk) {
                    mbar_wait(a_empty(a_st), a_ph ^ 1);
                    mbar_expect_tx(a_full(a_st), (uint32_t)p.box_bytes);
                    tma_load_4d(a_base + a_st * (uint32_t)p.a_stage_bytes, &p.tm_a, a_full(a_st), chunk * ck, t.ow0 - p.pl,
                                t.oh0 - p.pt, t.img);
                    if (++a_st == p.a_stages) { a_st = 0; a_ph ^= 1; }
                }
                next_tile_contig(p, t);
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: the whole warp walks the loop, one elected lane issues
        const bool leader = elect_one();
        const int m_tiles = p.m_tiles, n_pad = p.n_pad, kws = p.kw_steps, a_stages = p.a_stages, acc_stages = p.acc_stages;
        const uint32_t ck16 = (uint32_t)p.ck_bytes >> 4;
        const uint32_t a_stage16 = (uint32_t)p.a_stage_bytes >> 4, b_stage16 = (uint32_t)p.b_stage_bytes >> 4;
        const uint32_t m_step16 = 128u * ck16;                       // 128 window positions = 4 rows of pitch 32
        const int ksteps = p.ck_bytes >> 5;
        const uint32_t idesc = p.idesc;
        const uint64_t dconst = smem_desc_const(p);
        const uint32_t desc_hi = (uint32_t)(dconst >> 32);
        const uint32_t a_base16 = (uint32_t)dconst + (a_base >> 4), b_base16 = (uint32_t)dconst + (b_base >> 4);
        const uint32_t kx_step16 = (uint32_t)p.pack_g * ck16;                          // next tap group of a kernel row
        const uint32_t ky_step16 = (uint32_t)(p.pitch - kws * p.pack_g) * ck16;        // wrap to the next kernel row
        const uint32_t tap16 = (uint32_t)p.b_tx_bytes >> 4;
        int a_st = 0, a_ph = 0, acc_st = 0, acc_ph = 0;
        uint32_t a_cur = a_base16;
        if (tile_first < tile_end)
            for (int chunk = 0; chunk < n_chunks; ++chunk) mbar_wait(b_full(chunk), 0);   // resident from here on
        tc_fence_after();
        for (int tile = tile_first; tile < tile_end; ++tile) {
            mbar_wait(acc_empty(acc_st), acc_ph ^ 1);
            tc_fence_after();
            const uint32_t d_base = tmem_addr + (uint32_t)(acc_st * p.acc_cols);
            uint32_t accum = 0;
            for (int chunk = 0; chunk < n_chunks; ++chunk) {
                const int ks = min(ksteps, (p.k_bytes - chunk * p.ck_bytes + 31) >> 5);   // K steps that hold real channels
                mbar_wait(a_full(a_st), a_ph);
                tc_fence_after();
                uint32_t bt = b_base16 + (uint32_t)chunk * b_stage16, shift16 = 0;
                int kx = 0, tap_in_pass = 0;
                for (int tap = 0; tap < taps; ++tap) {
                    if (leader) {
                        uint32_t am = a_cur + shift16, dm = d_base;
                        for (int m = 0; m < m_tiles; ++m) {
                            tc_mma_bf16_w(dm, am, bt, desc_hi, idesc, accum);
                            if (ks > 1) tc_mma_bf16_w(dm, am + 2u, bt + 2u, desc_hi, idesc, 1u);
                            if (ks > 2) tc_mma_bf16_w(dm, am + 4u, bt + 4u, desc_hi, idesc, 1u);
                            if (ks > 3) tc_mma_bf16_w(dm, am + 6u, bt + 6u, desc_hi, idesc, 1u);
                            am += m_step16;
                            dm += (uint32_t)n_pad;
                        }
                    }
                    accum = 1;
                    bt += tap16;
                    shift16 += kx_step16;
                    if (++kx == kws) { kx = 0; shift16 += ky_step16; }
                    if (++tap_in_pass == sp_taps) { tap_in_pass = 0; shift16 = 0; kx = 0; }   // the w_lo pass restarts
                }
                if (leader) tc_commit(a_empty(a_st));
                __syncwarp();
                a_cur += a_stage16;
                if (++a_st == a_stages) { a_st = 0; a_ph ^= 1; a_cur = a_base16; }
            }
            if (leader) tc_commit(acc_full(acc_st));
            __syncwarp();
            if (++acc_st == acc_stages) { acc_st = 0; acc_ph ^= 1; }
        }
    } else {
        // ===================== epilogue (warps 2..9): TMEM -> slot sums -> BN / residual / activation -> NHWC stores
        const int quarter = warp & 3;                 // TMEM lanes [32q, 32q+32) are visible to warp%4 == q
        const int half = (warp - 2) >> 2;
        const int acc_stages = p.acc_stages;
        int acc_st = 0, acc_ph = 0;
        pdl_wait();                                   // residual reads and the stores below touch the arena
        if (tile_first < tile_end) {
            TileCoord t = decode_tile(p, tile_first);
            for (int tile = tile_first; tile < tile_end; ++tile) {
                mbar_wait(acc_full(acc_st), acc_ph);
                tc_fence_after();
                const uint32_t t_base = tmem_addr + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc_st * p.acc_cols);
                if (p.npack == 5) npack_tile_epilogue<TY, ACT, RES, 5>(p, t_base, half, quarter, lane, t, s_scale, s_shift);
                else if (p.npack == 3) npack_tile_epilogue<TY, ACT, RES, 3>(p, t_base, half, quarter, lane, t, s_scale, s_shift);
                else if (p.npack == 4) npack_tile_epilogue<TY, ACT, RES, 4>(p, t_base, half, quarter, lane, t, s_scale, s_shift);
                else npack_tile_epilogue<TY, ACT, RES, 2>(p, t_base, half, quarter, lane, t, s_scale, s_shift);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(acc_empty(acc_st));
                if (++acc_st == acc_stages) { acc_st = 0; acc_ph ^= 1; }
                next_tile_contig(p, t);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_addr),
                     "r"((uint32_t)p.tmem_cols)
                     : "memory");
    }
}

// ------------------------------------------------------------------ host side
static int env_int(const char* name, int dflt) {
    const char* s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}

constexpr int SMEM_BUDGET = 189 * 1024;      // operand rings; barriers + alignment slack on top
constexpr int SMEM_EXTRA_NPACK = 1024 + 512 + 2048;                 // conv_npack_kernel stages nothing in its epilogue
constexpr int SMEM_BUDGET_NPACK = 227 * 1024 - SMEM_EXTRA_NPACK - 1024;
constexpr int SMEM_EXTRA = 1024 + 512 + 2048 + 4 * 16 * 132 * 4;   // alignment slack + barriers + scale/shift slice + 2 half-sets x 2 staging tiles (swapped epilogue)

static bool umma_eligible(const EdeConv& a) {
    if (a.force_direct || a.w_umma == nullptr) return false;
    if (a.x_dtype != EDE_BF16 || a.x_layout != EDE_NHWC || a.relu_in) return false;
    if (a.sh != a.sw || (a.sh != 1 && a.sh != 2)) return false;
    if (a.kh * a.kw > 64 || a.kh > 8 || a.kw > 8) return false;
    if (a.w_passes != 1 && a.w_passes != 2) return false;
    if ((a.x_cs | a.x_co | a.y_cs | a.y_co | a.cin_p | a.cout_p) & 7) return false;
    if (a.res && ((a.res_cs | a.res_co) & 7)) return false;
    if (a.w_umma_cin_p & 7) return false;
    if (a.w_pack_g > 1 && (a.w_pack_g & EDE_PACK_N)) {      // taps on the N axis: G slots of cout_p16 rows
        const int G = a.w_pack_g & 15;
        if (G < 2 || G > 5 || a.w_umma_cout_p != G * round_up(a.cout_p, 16) || a.w_umma_cout_p > 256) return false;
    } else if (a.w_pack_g > 1 && (a.w_umma_cout_p != 128 || (a.w_pack_g != 2 && a.w_pack_g != 4))) return false;
    return true;
}

// operand ring depth: the persistent kernel owns a whole SM, so use the full smem budget
static int fit_stages(int fixed_bytes, int stage_bytes, int /*want*/) {
    int s = (SMEM_BUDGET - fixed_bytes) / stage_bytes;
    if (s > MAX_STAGES) s = MAX_STAGES;
    return s;
}

static int pow2_cols(int c) {
    int v = 32;
    while (v < c) v <<= 1;
    return v;
}

// ---- tile configuration by an explicit time model (numbers measured on B200, tools/micro/mma_rate.cu):
// one tcgen05.mma (M=128, K=16, operands in smem) takes max(N/2, base) cycles with base = 97 / 154 / 214 for
// the 128B / 64B / 32B swizzle (the A-operand fetch dominates below N ~ 192), independent of the start
// address alignment.  Operand bytes per tile are charged at the share of L2->SM bandwidth one SM gets.
static double mma_cycles(int swizzle_bytes, int n_pad) {
    const double base = swizzle_bytes == 128 ? 97.0 : swizzle_bytes == 64 ? 154.0 : 214.0;
    return std::max(base, n_pad * 0.5);
}
constexpr double L2_BYTES_PER_CLK_PER_SM = 27.0;     // TMA box bytes (zero fill included) one SM ingests per clock:
                                                     // 26..28 measured on the decoder convs (ncu, xbar2l1tex tma_ld)
constexpr double EPI_CYCLES_PER_ROW_COL = 0.012;     // epilogue cost per accumulator element (4 warps)
constexpr int NUM_SMS_MODEL = 148;

// 32-channel chunks (64-byte swizzle) only when they cut the padded K -- i.e. the TMA bytes -- by at least `gain`:
// measured, K = 24 / 32 run 1.6x faster with them, K = 64 / 80 slower (the tensor core reads narrow rows slower).
static bool narrow_pays(int cin_p, double gain) {
    const int k64 = round_up(cin_p, 64), k32 = round_up(cin_p, 32);
    return k32 < k64 && (double)(k64 - k32) >= gain * k64;
}

// Taps per weight stage of the resident modes.  Few large stages pipeline worse than many small ones (the issuer
// starts a stage only when all of it has landed, the producer refills it only when all of it is consumed), but a
// stage must carry enough MMAs to hide the issuing thread's per-stage bookkeeping (`target_clk`, measured): group
// just enough taps for that, judged on the chunk with the fewest K steps (the trimmed last one).
static int group_want(double mma_clk_per_kstep, int cin_p, int ck, int force_group, double target_clk,
                      double pair_below_clk) {
    if (force_group > 0) return force_group;
    const int tail = cin_p % ck ? ceil_div(cin_p % ck, 16) : ck / 16;        // K steps of the last chunk
    const double stage_clk = std::max(1.0, tail * mma_clk_per_kstep);
    const int t = std::max(1, std::min(6, (int)std::ceil(target_clk / stage_clk - 0.2)));
    return t == 1 && stage_clk < pair_below_clk ? 2 : t;
}

// TMEM accumulator stages: two, so that the epilogue of tile i overlaps the MMAs of tile i+1.  The N-packed form has
// tiles of 1..3 m-tiles whose issue -> MMA -> wake-up -> epilogue -> release round trip (5-8k cycles, tools/cta_timeline.py)
// is several times the MMA time of a tile: it takes as many stages as fit (up to 4).
static int umma_acc_stages(const EdeUmmaCfg& c) {
    if (c.swap || !(c.pack_g & EDE_PACK_N)) return 2;
    static const int forced = env_int("EDE_UMMA_ACC_STAGES", 0);
    const int fit = std::min(MAX_ACC_STAGES, 512 / std::max(1, c.m_tiles * c.n_pad));
    return std::max(2, forced > 0 ? std::min(forced, fit) : fit);
}

static void umma_config(const EdeConv& a, EdeUmmaCfg& c) {
    memset(&c, 0, sizeof(c));
    if (!umma_eligible(a)) return;
    const int force_mode = env_int("EDE_UMMA_MODE", -1);
    const int force_ck = env_int("EDE_UMMA_CK", 0);
    const int force_npad = env_int("EDE_UMMA_NPAD", 0);
    const int force_group = env_int("EDE_UMMA_GROUP", 0);
    const int n16 = round_up(a.cout_p, 16);
    // 1x1 convs fold N*H*W into one pixel axis -- unless every image has its own weights (tiles must not span images)
    const bool is_1x1 = a.kh == 1 && a.kw == 1 && a.sh == 1 && a.pt == 0 && a.pl == 0 && a.oh == a.h && a.ow == a.w &&
                        a.w_per_image <= 0;
    const int OH = is_1x1 ? 1 : a.oh;
    const long long OWl = is_1x1 ? (long long)a.n * a.h * a.w : a.ow;
    if (OWl > 0x7fffffffLL) return;
    const int OW = (int)OWl;
    const int n_img = is_1x1 ? 1 : a.n;
    const int taps_w = a.kh * a.kw * a.w_passes;

    EdeUmmaCfg best{};
    double best_time = 1e300;
    const bool unpacked_ok = a.w_pack_g <= 1;            // -1: caller asks which packing to prepack for
    // Weight boxes of the pixels-on-M candidates are charged at 40 B/clk, not 27: every CTA streams the SAME weights (L2
    // hits, no DRAM behind them) and the wide convs of ResNet-50's decoder were measured at ~50 B/clk of TMA ingest with
    // N = 256 tiles the 27 B/clk model had rejected (D.up2 / D.up3 of RN50-HU: 0.71 -> 0.65 / 0.63 ms; sum of the plan's
    // ops 11.85 -> 11.59 ms, ENB4-HU 4.56 -> 4.50, ENB0-HU unchanged).  EDE_UMMA_WBPC overrides.
    const double w_bpc = (double)env_int("EDE_UMMA_WBPC", 40);
    // contention factors on top of the isolated MMA rates (TMA fills and the MMA operand reads share the
    // 128 B/clk of shared-memory bandwidth; measured on the decoder convs with tools/cta_timeline.py)
    const double F_NOSWAP = 1.5, F_SWAP = 1.3;
    // Channel chunk: the 128-byte swizzle (64 bf16 channels per smem row) or the 64-byte one (32 channels).
    // Narrower swizzles feed the tensor core at 63 % / 45 % of the rate (profiles/r01_mma_rate.md) and the MMA
    // issuer skips the K steps of the last chunk that only hold TMA zero fill, so padding K to 64 costs no
    // tensor-core time -- but TMA ingests the zero fill at the same ~27 B/clk as real bytes, so layers whose K
    // is far from a multiple of 64 (24, 32, 96 ...) run faster with 32-channel chunks.  The time model decides.
    const int cks[3] = {64, 32, 16};
    const int ks_total = ceil_div(a.cin_p, 16);           // K = 16 steps that hold real channels
    for (int ci = 0; ci < 3; ++ci) {
        const int ck = cks[ci];
        if (force_ck ? force_ck != ck : (ck == 16 || (ck == 32 && !narrow_pays(a.cin_p, 0.25)))) continue;
        const int ck_bytes = ck * 2;
        const int n_chunks = ceil_div(a.cin_p, ck);
        for (int n_tiles = ceil_div(n16, 256); n_tiles <= ceil_div(n16, 16) && n_tiles <= 64; ++n_tiles) {
            const int n_pad = round_up(ceil_div(n16, n_tiles), 16);
            if (n_pad > 256) continue;
            if (force_npad && n_pad != force_npad) continue;
            if (n_tiles > ceil_div(n16, 256) && n_pad < 96) break;      // finer N splits never pay
            const int b_bytes = round_up(n_pad * ck_bytes, 1024);
            const int max_t = 256 / n_pad;                // two accumulator stages share the 512 TMEM columns
            if (max_t < 1) continue;
            const double t_mma = mma_cycles(ck_bytes, n_pad) * F_NOSWAP;
            auto consider = [&](const EdeUmmaCfg& cand, double a_bytes_tile, double b_bytes_tile) {
                const long long tiles = (long long)cand.tiles_w * cand.tiles_h * n_img * n_tiles;
                const double rounds = (double)ceil_div((int)std::min<long long>(tiles, 1 << 30), NUM_SMS_MODEL);
                const double mma = (double)taps_w * cand.m_tiles * ks_total * t_mma;
                const double mem = a_bytes_tile / L2_BYTES_PER_CLK_PER_SM + b_bytes_tile / w_bpc;
                const double epi = cand.m_tiles * 128.0 * n_pad * EPI_CYCLES_PER_ROW_COL * 4.0 + 600.0;
                // operand ring too shallow to cover the L2 latency (~1500 clk) -> exposed waits
                const double per_stage = mma / std::max(1, n_chunks * taps_w);
                const double lat = std::max(0.0, 1500.0 - cand.b_stages * per_stage) * 0.15 * n_chunks * taps_w /
                                   std::max(1, cand.b_stages);
                const double t = rounds * (std::max(std::max(mma, mem), epi) + lat + 300.0) + 4000.0;
                if (t < best_time) {
                    best_time = t;
                    best = cand;
                    best.ck = ck; best.swizzle = ck_bytes; best.n_pad = n_pad; best.n_tiles = n_tiles;
                    if (best.b_stage_bytes == 0) best.b_stage_bytes = b_bytes;
                    const double kfrac = (double)a.cin_p / (n_chunks * ck);
                    const double nfrac = (double)a.cout_p / ((double)n_tiles * n_pad);
                    best.useful_frac = cand.useful_frac * kfrac * nfrac;
                }
            };
            // ---- per-tap candidates: every (tap, chunk) TMA-loads its own 128-pixel boxes
            if (force_mode != 0 && unpacked_ok) {
                for (int bw = 128; bw >= 1; bw >>= 1) {
                    const int bh = 128 / bw;
                    if (bw * a.sw > 256 || bh * a.sh > 256) continue;
                    if (is_1x1 && bh != 1) continue;
                    if (!is_1x1 && bw > 2 * OW && bw > 1) continue;
                    for (int T = std::min(4, max_t); T >= 1; --T) {
                        const int a_bytes = T * 128 * ck_bytes;
                        const int stages = fit_stages(0, a_bytes + b_bytes, 4);
                        if (stages < 3) continue;
                        EdeUmmaCfg cand{};
                        cand.a_mode = 1; cand.bh = bh; cand.bw = bw; cand.pitch = bw; cand.box_rows = bh;
                        cand.m_tiles = T; cand.a_stages = cand.b_stages = stages; cand.a_stage_bytes = a_bytes;
                        cand.tiles_w = ceil_div(OW, bw * T); cand.tiles_h = ceil_div(OH, bh);
                        cand.useful_frac = (double)OH * OW / ((double)cand.tiles_w * cand.tiles_h * T * 128);
                        consider(cand, (double)n_chunks * taps_w / a.w_passes * a_bytes * (a.w_passes),
                                 (double)n_chunks * taps_w * n_pad * ck_bytes);
                    }
                }
            }
            // ---- resident candidates (stride 1, k x k with k > 1): halo tile loaded once per chunk
            if (a.sh == 1 && a.kh * a.kw >= 9 && force_mode != 1 && !is_1x1 && unpacked_ok) {
                for (int div = 1; div <= 8; ++div) {
                    const int bw = ceil_div(OW, div);
                    const int P = bw + a.kw - 1;
                    if (P > 256) continue;
                    for (int bh = 1; bh <= 64 && bh <= OH + 8; ++bh) {
                        const int box_rows = bh + a.kh - 1;
                        if (box_rows > 256) break;
                        const int m_need = (bh - 1) * P + bw;
                        const int T = ceil_div(m_need, 128);
                        if (T > max_t) break;
                        const int rows_alloc = std::max(box_rows * P, T * 128 + (a.kh - 1) * P + a.kw - 1);
                        const int a_bytes = round_up(rows_alloc * ck_bytes, 1024);
                        if (fit_stages(2 * a_bytes, b_bytes, 4) < 3) continue;
                        // taps per weight stage (see the swapped candidates): three stages, or two that each hold
                        // at least ~1500 cycles of MMAs (the load of one hides behind the other)
                        int grp = 1;
                        const int tap_bytes = n_pad * ck_bytes;
                        const int want = group_want(T * mma_cycles(ck_bytes, n_pad), a.cin_p, ck, force_group, 776.0, 1000.0);   // m-tile loop: more bookkeeping per tap, pairs pay up to ~1000 clk
                        for (int t = tap_bytes % 1024 ? 1 : std::min(taps_w, want); t > 1; --t) {
                            const int st = fit_stages(2 * a_bytes, t * tap_bytes, 4);
                            if (st >= 3 || (st == 2 && t * T * ceil_div(ck, 16) * t_mma >= 1500.0)) { grp = t; break; }
                        }
                        const int b_grp_bytes = grp > 1 ? grp * tap_bytes : b_bytes;
                        const int b_stages = fit_stages(2 * a_bytes, b_grp_bytes, 4);
                        EdeUmmaCfg cand{};
                        cand.a_mode = 0; cand.bh = bh; cand.bw = bw; cand.pitch = P; cand.box_rows = box_rows;
                        cand.m_tiles = T; cand.a_stages = 2; cand.b_stages = b_stages; cand.a_stage_bytes = a_bytes;
                        cand.b_stage_bytes = b_grp_bytes;
                        cand.tiles_w = ceil_div(OW, bw); cand.tiles_h = ceil_div(OH, bh);
                        cand.useful_frac = (double)OH * OW / ((double)cand.tiles_w * cand.tiles_h * T * 128);
                        consider(cand, (double)n_chunks * box_rows * P * ck_bytes,
                                 (double)n_chunks * taps_w * n_pad * ck_bytes);
                    }
                }
            }
        }
    }
    // ---- operand-swapped candidates (Cout <= 128): weights on the M side (128 rows), pixels on N.  One MMA costs
    // 128 cycles for 256 pixels instead of 97 for 128.  With Cout <= 64 / 32 the 128 M rows hold G = 2 / 4 taps
    // (kx, kx+1, ...) of the kernel row at once -- row [j][c] = W[c][.][ky][gx*G + j] -- so a kernel row costs
    // ceil(kw / G) MMAs instead of kw; the epilogue adds the G partial sums, each read j columns further right.
    // Packed tiles use N = 240 so that those shifted reads stay inside the 256-column accumulator stage.
    const int force_swap = env_int("EDE_UMMA_SWAP", -1);
    const int force_pack = env_int("EDE_UMMA_PACK", -1);
    if (n16 <= 128 && force_swap != 0) {
        if (force_swap == 1) best_time = 1e300;
        const int gs[3] = {4, 2, 1};
        for (int gi = 0; gi < 3; ++gi) {
            const int G = gs[gi];
            if (G > 1 && (a.cout_p > 128 / G || a.kw < 2 || a.sh != 1 || is_1x1)) continue;
            if (a.w_pack_g >= 0 && std::max(1, a.w_pack_g) != G) continue;      // weights are already packed for one G (or on N)
            if (force_pack >= 0 && a.w_pack_g < 0 && std::max(1, force_pack) != G) continue;
            const int kws = ceil_div(a.kw, G);
            const int steps = a.kh * kws * a.w_passes;
            const int n_cols = G > 1 ? 240 : 256;
          for (int cki = 0; cki < 2; ++cki) {
            const int ck = cks[cki];
            if (force_ck ? force_ck != ck : (ck == 32 && !narrow_pays(a.cin_p, 0.3))) continue;
            const int ck_bytes = ck * 2, n_chunks = ceil_div(a.cin_p, ck);
            // unpacked: the TMA box stops at the real output channels; the MMA still reads 128 rows, the rows past
            // the box are stale smem that lands in accumulator lanes nobody reads (each lane is independent)
            const int w_rows_box = G > 1 ? 128 : std::min(128, round_up(a.cout_p, 8));
            const int b_bytes = round_up(w_rows_box * ck_bytes, 1024);
            const double t_mma = n_cols * 0.5 * F_SWAP * (ck == 64 ? 1.0 : ck == 32 ? 1.3 : 1.8);
            auto consider = [&](const EdeUmmaCfg& cand, double a_bytes_tile) {
                const long long tiles = (long long)cand.tiles_w * cand.tiles_h * n_img;
                const double rounds = (double)ceil_div((int)std::min<long long>(tiles, 1 << 30), NUM_SMS_MODEL);
                const double mma = (double)steps * ks_total * t_mma;
                const double w_rows = G > 1 ? 128.0 : (double)std::min(128, round_up(a.cout_p, 8));
                const double mem = (a_bytes_tile + (double)steps * n_chunks * w_rows * ck_bytes) / L2_BYTES_PER_CLK_PER_SM;
                const double epi = 16.0 * 350.0 + 600.0;      // 2 x 8 staged chunks, each a ~500-cycle dependent chain (measured)
                const double per_stage = mma / std::max(1, n_chunks * steps);
                const double lat = std::max(0.0, 1500.0 - cand.b_stages * per_stage) * 0.15 * n_chunks * steps /
                                   std::max(1, cand.b_stages);
                const double t = rounds * (std::max(std::max(mma, mem), epi) + lat + 300.0) + 4000.0;
                if (t < best_time) {
                    best_time = t;
                    best = cand;
                    best.swap = 1;
                    best.pack_g = G;
                    best.ck = ck; best.swizzle = ck_bytes; best.n_pad = 128; best.n_tiles = 1;
                    if (best.b_stage_bytes == 0) best.b_stage_bytes = b_bytes;
                    best.useful_frac = cand.useful_frac * ((double)a.cin_p / (ks_total * 16)) *
                                       ((double)a.cout_p * a.kw / (128.0 * kws));
                }
            };
            if (force_mode != 0 && G == 1) {
                for (int bw = 128; bw >= 1; bw >>= 1) {
                    const int bh = 128 / bw;
                    if (bw * a.sw > 256 || bh * a.sh > 256) continue;
                    if (is_1x1 && bh != 1) continue;
                    if (!is_1x1 && bw > 2 * OW && bw > 1) continue;
                    const int a_bytes = 2 * 128 * ck_bytes;
                    const int stages = fit_stages(0, a_bytes + b_bytes, 4);
                    if (stages < 3) continue;
                    EdeUmmaCfg cand{};
                    cand.a_mode = 1; cand.bh = bh; cand.bw = bw; cand.pitch = bw; cand.box_rows = bh;
                    cand.m_tiles = 2; cand.a_stages = cand.b_stages = stages; cand.a_stage_bytes = a_bytes;
                    cand.tiles_w = ceil_div(OW, bw * 2); cand.tiles_h = ceil_div(OH, bh);
                    cand.useful_frac = (double)OH * OW / ((double)cand.tiles_w * cand.tiles_h * 256);
                    consider(cand, (double)n_chunks * steps * a_bytes);
                }
            }
            if (a.sh == 1 && a.kh * a.kw >= 9 && force_mode != 1 && !is_1x1) {
                for (int div = 1; div <= 8; ++div) {
                    const int bw = ceil_div(OW, div);
                    const int P = bw + a.kw - 1;
                    if (P > 256) continue;
                    for (int bh = 1; bh <= 64 && bh <= OH + 8; ++bh) {
                        const int box_rows = bh + a.kh - 1;
                        if (box_rows > 256) break;
                        const int m_need = (bh - 1) * P + bw;
                        if (m_need + G - 1 > n_cols) break;
                        const int rows_alloc = std::max(box_rows * P, n_cols + (a.kh - 1) * P + kws * G);
                        const int a_bytes = round_up(rows_alloc * ck_bytes, 1024);
                        // taps per weight stage: as many as still leave three stages (one wait + one commit per
                        // stage; the issuing thread's bookkeeping per stage costs about as much as four MMAs)
                        int grp = 1;
                        const int want = group_want(mma_cycles(ck_bytes, n_cols), a.cin_p, ck, force_group, 500.0, 0.0);
                        for (int t = (w_rows_box * ck_bytes) % 1024 ? 1 : std::min(steps, want); t > 1; --t)
                            if (fit_stages(2 * a_bytes, round_up(t * w_rows_box * ck_bytes, 1024), 4) >= 3) { grp = t; break; }
                        const int b_grp_bytes = round_up(grp * w_rows_box * ck_bytes, 1024);
                        const int b_stages = fit_stages(2 * a_bytes, b_grp_bytes, 4);
                        if (b_stages < 3) continue;
                        EdeUmmaCfg cand{};
                        cand.a_mode = 0; cand.bh = bh; cand.bw = bw; cand.pitch = P; cand.box_rows = box_rows;
                        cand.m_tiles = 2; cand.a_stages = 2; cand.b_stages = b_stages; cand.a_stage_bytes = a_bytes;
                        cand.b_stage_bytes = b_grp_bytes;
                        cand.tiles_w = ceil_div(OW, bw); cand.tiles_h = ceil_div(OH, bh);
                        cand.useful_frac = (double)OH * OW / ((double)cand.tiles_w * cand.tiles_h * n_cols);
                        consider(cand, (double)n_chunks * box_rows * P * ck_bytes);
                    }
                }
            }
          }
        }
    }
    // ---- N-packed pixels-on-M candidates (narrow stride-1 k x k convs): the G taps of a kernel row sit side by side on
    // the UMMA N axis (column j*slot + c, slot = cout_p16), so a kernel row costs ceil(kw/G) MMAs of max(97, N/2) cycles
    // per 128 window positions instead of kw.  The smem pixel pitch is 32 -- one warp of the epilogue per window row -- so
    // the slot sums are warp shuffles (epilogue_npack).  Offered when the whole weight set stays resident in shared
    // memory: with 112..120 useful pixels per m-tile, re-streaming the weights per tile would bind on TMA ingest.
    const int force_npack = env_int("EDE_UMMA_NPACK", -1);
    if (a.sh == 1 && a.kh * a.kw >= 9 && a.kw >= 2 && a.kw <= 5 && !is_1x1 && force_npack != 0 && force_mode != 1 &&
        (a.w_pack_g < 0 || (a.w_pack_g > 1 && (a.w_pack_g & EDE_PACK_N))) && a.w_per_image <= 0) {
        if (force_npack == 1) best_time = 1e300;
        const int slot = n16;
        int G = std::min(a.kw, 256 / std::max(16, slot));
        if (a.w_pack_g > 1) G = a.w_pack_g & 15;
        const int n_pad = G * slot;
        const int kws = ceil_div(a.kw, std::max(1, G));
        const int steps = a.kh * kws * a.w_passes;
        const int bw = 32 - (a.kw - 1);
        for (int cki = 0; cki < 2 && G >= 2 && n_pad <= 256; ++cki) {
            const int ck = cks[cki];
            if (force_ck && force_ck != ck) continue;
            if (!force_ck && ck == 32 && round_up(a.cin_p, 32) >= round_up(a.cin_p, 64)) continue;
            const int ck_bytes = ck * 2, n_chunks = ceil_div(a.cin_p, ck);
            const int w_chunk_bytes = steps * n_pad * ck_bytes;              // one stage = all taps of a channel chunk
            if (n_chunks > MAX_STAGES || steps > 256 || n_chunks * w_chunk_bytes > 160 * 1024) continue;
            const double t_mma = mma_cycles(ck_bytes, n_pad) * 1.15;
            const int force_T = env_int("EDE_UMMA_NPACK_T", 0);
            for (int T = 1; T <= 256 / n_pad && T <= 8; ++T) {
                if (force_T && T != force_T && force_T <= 256 / n_pad) continue;
                const int bh = 4 * T, box_rows = bh + a.kh - 1;
                const int rows_alloc = std::max(box_rows * 32, T * 128 + (a.kh - 1) * 32 + kws * G);
                const int a_bytes = round_up(rows_alloc * ck_bytes, 1024);
                const int a_stages = std::min(6, (SMEM_BUDGET_NPACK - n_chunks * w_chunk_bytes) / a_bytes);
                if (a_stages < 2) continue;
                EdeUmmaCfg cand{};
                cand.a_mode = 0; cand.bh = bh; cand.bw = bw; cand.pitch = 32; cand.box_rows = box_rows;
                cand.m_tiles = T; cand.a_stages = a_stages; cand.b_stages = n_chunks; cand.a_stage_bytes = a_bytes;
                cand.b_stage_bytes = w_chunk_bytes;
                cand.tiles_w = ceil_div(OW, bw); cand.tiles_h = ceil_div(OH, bh);
                cand.useful_frac = (double)OH * OW / ((double)cand.tiles_w * cand.tiles_h * T * 128);
                const long long tiles = (long long)cand.tiles_w * cand.tiles_h * n_img;
                const double rounds = (double)ceil_div((int)std::min<long long>(tiles, 1 << 30), NUM_SMS_MODEL);
                const double mma = (double)steps * T * ks_total * t_mma;
                const double mem = (double)n_chunks * box_rows * 32 * ck_bytes / L2_BYTES_PER_CLK_PER_SM;
                // per (m-tile, 16-channel chunk) unit: G tcgen05.ld + 16 (G-1) shuffles + BN / activation / stores; two
                // warps per lane quarter share the units
                const double epi = std::ceil(T * (slot / 16) / 2.0) * (260.0 + 90.0 * G) + 300.0;
                // every role pays ~1500 cycles of per-tile bookkeeping (waits, fences, commits run at full latency: one
                // warp per scheduler), and a pixel tile lands ~3000 cycles after its stage was released: with few stages
                // the difference is exposed (measured: 5x5 24->32 runs 20 % faster with 32-channel chunks = 6 stages of
                // 16 KB than with 64-channel chunks = 2 stages of 33 KB although its MMAs are 1.6x slower)
                double t_tile = std::max(std::max(std::max(mma, mem), epi), 1500.0);
                if ((a_stages - 1) * t_tile < 3000.0) t_tile += (3000.0 - (a_stages - 1) * t_tile) / a_stages;
                const double t = rounds * (t_tile + 400.0) + 4000.0;
                if (t < best_time) {
                    best_time = t;
                    best = cand;
                    best.swap = 0;
                    best.pack_g = EDE_PACK_N | G;
                    best.ck = ck; best.swizzle = ck_bytes; best.n_pad = n_pad; best.n_tiles = 1;
                    best.useful_frac = cand.useful_frac * ((double)a.cin_p / (ks_total * 16)) *
                                       ((double)a.cout_p * a.kw / ((double)n_pad * kws));
                }
            }
        }
    }
    if (best.m_tiles == 0) return;
    // 1x1 convs with hi + lo weights (per-tap mode): both passes read the SAME pixel box, so a weight stage can hold both
    // tiles and the box is loaded once per chunk instead of once per pass -- the deep-K convs are bound by TMA ingest
    // (project 2304 -> 192 @7x9 moved 2 MB per 128-pixel tile: 0.039 -> 0.025 ms).  Applied to the tile the picker
    // chose, and only where the doubled weight stage still leaves three ring stages (letting the picker re-tile
    // for it made the shallow-K convs of ResNet-50 slower: more n-tiles).
    if (best.a_mode == 1 && !best.swap && a.kh * a.kw == 1 && a.w_passes == 2 && best.ck >= 32 &&
        env_int("EDE_UMMA_SHARE_PASSES", 1) != 0) {
        const int b2 = 2 * best.n_pad * best.ck * 2;
        const int st = fit_stages(0, best.a_stage_bytes + b2, 4);
        if (st >= 3) { best.b_stage_bytes = b2; best.a_stages = best.b_stages = st; }
    }
    c = best;
    c.eligible = 1;
    if (c.pack_g < 1) c.pack_g = 1;
    c.tmem_cols = c.swap ? 512 : pow2_cols(umma_acc_stages(c) * c.m_tiles * c.n_pad);
    c.smem_bytes = c.a_stages * c.a_stage_bytes + c.b_stages * c.b_stage_bytes +
                   ((c.pack_g & EDE_PACK_N) ? SMEM_EXTRA_NPACK : SMEM_EXTRA);
    c.flat = is_1x1;
}

int conv_direct_launch(const EdeConv& a, cudaStream_t st);

static unsigned long long* g_prof_last = nullptr;
static size_t g_prof_ctas = 0;

template <typename TY, int ACT, bool RES>
static int launch_variant(const UmmaParams& p, int n_cta, int smem, int dev, cudaStream_t st) {
    static bool attr_set[64] = {};
    if (!attr_set[dev]) {
        int rc = check_cuda(cudaFuncSetAttribute(conv_umma_kernel<TY, ACT, RES>,
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024),
                            "cudaFuncSetAttribute(conv_umma_kernel)");
        if (rc) return rc;
        attr_set[dev] = true;
    }
    return check_cuda(launch_pdl(conv_umma_kernel<TY, ACT, RES>, dim3(n_cta), dim3(NUM_THREADS), (size_t)smem, st, p),
                      "launch(conv_umma_kernel)");
}

template <typename TY, int ACT, bool RES>
static int launch_variant_npack(const UmmaParams& p, int n_cta, int smem, int dev, cudaStream_t st) {
    static bool attr_set[64] = {};
    if (!attr_set[dev]) {
        int rc = check_cuda(cudaFuncSetAttribute(conv_npack_kernel<TY, ACT, RES>,
                                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024),
                            "cudaFuncSetAttribute(conv_npack_kernel)");
        if (rc) return rc;
        attr_set[dev] = true;
    }
    return check_cuda(launch_pdl(conv_npack_kernel<TY, ACT, RES>, dim3(n_cta), dim3(NUM_THREADS), (size_t)smem, st, p),
                      "launch(conv_npack_kernel)");
}

// Everything a launch needs, derived once per distinct EdeConv (ede_conv2d keeps these in a small cache keyed by
// the struct's bytes: a forward plan replays the same structs, so per-forward host work is one hash + one launch).
struct ConvLaunch {
    UmmaParams p;
    int n_cta, smem;
};

static int conv_umma_build(const EdeConv& a, const EdeUmmaCfg& c, ConvLaunch& L) {
    UmmaParams& p = L.p;
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        set_error("ede_conv2d: cuTensorMapEncodeTiled driver entry point unavailable");
        return EDE_E_DRIVER;
    }
    memset(&p, 0, sizeof(p));
    const CUtensorMapSwizzle sw = c.swizzle == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                  : c.swizzle == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                    : CU_TENSOR_MAP_SWIZZLE_32B;
    // ---- activations
    {
        const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(a.x) + a.x_co;
        cuuint64_t dims[4], strides[3];
        cuuint32_t box[4], es[4];
        const cuuint64_t pix = (cuuint64_t)a.x_cs * 2;
        if (c.flat) {
            const cuuint64_t npx = (cuuint64_t)a.n * a.h * a.w;
            dims[0] = a.cin_p; dims[1] = npx; dims[2] = 1; dims[3] = 1;
            strides[0] = pix; strides[1] = pix * npx; strides[2] = pix * npx;
        } else {
            dims[0] = a.cin_p; dims[1] = a.w; dims[2] = a.h; dims[3] = a.n;
            strides[0] = pix; strides[1] = pix * a.w; strides[2] = pix * a.w * a.h;
        }
        box[0] = c.ck; box[3] = 1;
        es[0] = 1; es[3] = 1;
        if (c.a_mode == 0) { box[1] = c.pitch; box[2] = c.box_rows; es[1] = es[2] = 1; }
        else { box[1] = c.bw * a.sw; box[2] = c.bh * a.sh; es[1] = a.sw; es[2] = a.sh; }
        CUresult r = enc(&p.tm_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, (void*)base, dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            set_error("ede_conv2d: cuTensorMapEncodeTiled(A) failed with %d (box %u,%u,%u)", (int)r, box[0], box[1],
                      box[2]);
            return EDE_E_DRIVER;
        }
    }
    // ---- weights
    {
        // dim 0 stops at the real K: the zero tail of the last 64-channel chunk is TMA out-of-bounds fill
        // (no L2 traffic) instead of zeros streamed from the padded weight rows
        cuuint64_t dims[3] = {(cuuint64_t)std::min(a.cin_p, a.w_umma_cin_p), (cuuint64_t)a.w_umma_cout_p,
                              (cuuint64_t)(a.kh * a.kw * a.w_passes)};
        cuuint64_t strides[2] = {(cuuint64_t)a.w_umma_cin_p * 2, (cuuint64_t)a.w_umma_cin_p * 2 * a.w_umma_cout_p};
        const int w_rows_box = c.swap ? (c.pack_g > 1 ? 128 : std::min(128, round_up(a.cout_p, 8))) : c.n_pad;
        const int taps_per_mma = (c.pack_g & EDE_PACK_N) ? (c.pack_g & 15) : std::max(1, c.pack_g);
        // taps per weight stage: the stage size the tile picker chose, in units of one tap's box
        int b_group = (c.a_mode == 0 && (w_rows_box * c.ck * 2) % 1024 == 0)
                          ? std::max(1, c.b_stage_bytes / (w_rows_box * c.ck * 2)) : 1;
        // per-tap 1x1 conv with hi + lo weights sharing one pixel box per chunk (umma_config, share_passes)
        if (c.a_mode == 1 && !c.swap && a.kh * a.kw == 1 && a.w_passes == 2 && c.b_stage_bytes == 2 * w_rows_box * c.ck * 2)
            b_group = 2;
        p.b_group = b_group;
        cuuint32_t box[3] = {(cuuint32_t)c.ck, (cuuint32_t)w_rows_box, (cuuint32_t)b_group};
        dims[2] = (cuuint64_t)(a.kh * ceil_div(a.kw, taps_per_mma) * a.w_passes);
        if (a.w_per_image > 0) {
            if (a.w_per_image != (int)dims[2]) { set_error("ede_conv2d: w_per_image must equal the taps of one weight set (%d)", (int)dims[2]); return EDE_E_ARG; }
            p.w_img_taps = a.w_per_image;
            dims[2] *= (cuuint64_t)a.n;
        }
        cuuint32_t es[3] = {1, 1, 1};
        CUresult r = enc(&p.tm_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, (void*)a.w_umma, dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            set_error("ede_conv2d: cuTensorMapEncodeTiled(B) failed with %d", (int)r);
            return EDE_E_DRIVER;
        }
    }
    p.oh = c.flat ? 1 : a.oh;
    p.ow = c.flat ? a.n * a.h * a.w : a.ow;
    p.tiles_w = c.tiles_w; p.tiles_h = c.tiles_h;
    p.n_img = c.flat ? 1 : a.n;
    p.n_tiles = c.n_tiles;
    p.total_tiles = c.tiles_w * c.tiles_h * p.n_img * c.n_tiles;
    p.acc_cols = c.swap ? 256 : c.m_tiles * c.n_pad;
    p.acc_stages = umma_acc_stages(c);
    p.a_mode = c.a_mode; p.ck_bytes = c.ck * 2; p.k_bytes = a.cin_p * 2; p.n_chunks = ceil_div(a.cin_p, c.ck);
    p.w_passes = a.w_passes;
    p.kh = a.kh; p.kw = a.kw; p.sh = a.sh; p.sw = a.sw; p.pt = a.pt; p.pl = a.pl;
    p.bw = c.bw; p.bh = c.bh; p.pitch = c.pitch; p.box_bytes = c.pitch * c.box_rows * c.ck * 2;
    p.m_tiles = c.m_tiles; p.n_pad = c.n_pad;
    p.a_stages = c.a_stages; p.b_stages = c.b_stages;
    p.a_stage_bytes = c.a_stage_bytes; p.b_stage_bytes = c.b_stage_bytes; p.b_tx_bytes = (c.swap ? (c.pack_g > 1 ? 128 : std::min(128, round_up(a.cout_p, 8))) : c.n_pad) * c.ck * 2;
    p.layout_type = c.swizzle == 128 ? 2 : c.swizzle == 64 ? 4 : 6;
    p.sbo_bytes = 8 * c.ck * 2;
    p.tmem_cols = c.tmem_cols;
    p.flat = c.flat;
    const bool on_n = (c.pack_g & EDE_PACK_N) != 0;
    p.pack_g = on_n ? (c.pack_g & 15) : c.pack_g;
    p.npack = on_n ? p.pack_g : 0;
    p.slot = on_n ? c.n_pad / p.pack_g : 0;
    p.b_resident = on_n ? 1 : 0;
    p.kw_steps = ceil_div(a.kw, p.pack_g); p.n_cols = c.swap ? (c.pack_g > 1 ? 240 : 256) : c.n_pad;
    p.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.n_cols >> 3) << 17) | ((128u >> 4) << 24);
    p.swap = c.swap;
    p.y = a.y; p.res = a.res; p.scale = a.scale; p.shift = a.shift;
    p.y_cs = a.y_cs; p.y_co = a.y_co; p.res_cs = a.res_cs; p.res_co = a.res_co;
    p.cout_p = a.cout_p; p.act = a.act; p.act_c_end = a.act_c_end; p.y_dtype = a.y_dtype;

    static int sm_count[64] = {};
    int dev = 0;
    (void)cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (sm_count[dev] == 0) {
        int rc = check_cuda(cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev),
                            "cudaDeviceGetAttribute(MultiProcessorCount)");
        if (rc) return rc;
    }
    // persistent grid: one CTA per SM (the dynamic smem request is kept above half an SM so that two CTAs
    // can never share one SM and fight over the 512 TMEM columns)
    L.n_cta = std::min(p.total_tiles, std::max(1, sm_count[dev]));
    p.tiles_per_cta = p.npack ? ceil_div(p.total_tiles, L.n_cta) : 0;
    L.smem = std::max(c.smem_bytes, 120 * 1024);
    return EDE_OK;
}

static int conv_umma_launch(const EdeConv& a, ConvLaunch& L, cudaStream_t st) {
    UmmaParams& p = L.p;
    const int n_cta = L.n_cta, smem = L.smem;
    int dev = 0;
    (void)cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    static unsigned long long* prof_buf = nullptr;
    static const int prof_on = env_int("EDE_UMMA_PROF", 0);

    if (prof_on) {
        if (!prof_buf && cudaMalloc(&prof_buf, (size_t)4096 * 16 * sizeof(unsigned long long)) != cudaSuccess)
            prof_buf = nullptr;
        p.prof = (prof_buf && n_cta <= 4096) ? prof_buf : nullptr;
        g_prof_last = p.prof;
        g_prof_ctas = p.prof ? (size_t)n_cta : 0;
    }
    // one kernel instantiation per (output dtype, activation, residual): the epilogue has no per-element branches
    int rc = EDE_OK;
#define EDE_UMMA_LAUNCH(TY, ACT, RES) \
    rc = p.npack ? launch_variant_npack<TY, ACT, RES>(p, n_cta, smem, dev, st) : launch_variant<TY, ACT, RES>(p, n_cta, smem, dev, st)
#define EDE_UMMA_ACT(TY, ACT) do { if (a.res) EDE_UMMA_LAUNCH(TY, ACT, true); else EDE_UMMA_LAUNCH(TY, ACT, false); } while (0)
#define EDE_UMMA_TY(TY)                                                      \
    do {                                                                     \
        if (a.act == EDE_ACT_NONE) EDE_UMMA_ACT(TY, EDE_ACT_NONE);           \
        else if (a.act == EDE_ACT_RELU) EDE_UMMA_ACT(TY, EDE_ACT_RELU);      \
        else if (a.act == EDE_ACT_SWISH) EDE_UMMA_ACT(TY, EDE_ACT_SWISH);    \
        else EDE_UMMA_ACT(TY, ACT_RUNTIME);                                  \
    } while (0)
    if (a.y_dtype == EDE_BF16) EDE_UMMA_TY(__nv_bfloat16);
    else if (a.y_dtype == EDE_BF16X2) EDE_UMMA_TY(bf16x2_t);
    else EDE_UMMA_TY(float);
#undef EDE_UMMA_TY
#undef EDE_UMMA_ACT
#undef EDE_UMMA_LAUNCH
    if (rc) return rc;
    EDE_LAUNCH_CHECK(p.npack ? "conv_npack_kernel" : "conv_umma_kernel");
    return EDE_OK;
}

// ---- launch cache: EdeConv bytes (pointers, shapes, flags) + device -> tile configuration + tensor maps
constexpr size_t CONV_KEY_BYTES = offsetof(EdeConv, w_per_image) + sizeof(int32_t);
struct ConvCacheEntry {
    unsigned char key[CONV_KEY_BYTES];
    int dev;
    bool eligible;
    ConvLaunch launch;
};
static uint64_t conv_key_hash(const EdeConv& a, int dev) {
    const unsigned char* b = reinterpret_cast<const unsigned char*>(&a);
    uint64_t h = 1469598103934665603ull ^ (uint64_t)dev;
    for (size_t i = 0; i < CONV_KEY_BYTES; ++i) h = (h ^ b[i]) * 1099511628211ull;
    return h;
}
static ConvCacheEntry* conv_cache_find(const EdeConv& a, int dev, bool insert) {
    // thread-local: DataParallel runs one host thread per device, no locking on the hot path
    thread_local std::unordered_multimap<uint64_t, ConvCacheEntry*> cache;
    const uint64_t h = conv_key_hash(a, dev);
    auto range = cache.equal_range(h);
    for (auto it = range.first; it != range.second; ++it)
        if (it->second->dev == dev && memcmp(it->second->key, &a, CONV_KEY_BYTES) == 0) return it->second;
    if (!insert) return nullptr;
    if (cache.size() >= 8192) {                      // bounded: a process that keeps compiling new plans starts over
        for (auto& kv : cache) delete kv.second;
        cache.clear();
    }
    ConvCacheEntry* e = new ConvCacheEntry();
    memcpy(e->key, &a, CONV_KEY_BYTES);
    e->dev = dev;
    cache.emplace(h, e);
    return e;
}

}  // namespace ede

extern "C" {

/* debug hook (EDE_UMMA_PROF=1): copies the per-CTA globaltimer stamps of the LAST tcgen05 conv launch
 * (8 u64 per CTA: [0] entry, [1] issuer start, [2] issuer end, [3] epilogue end (globaltimer ns); [4] cycles the
 * issuer waited for operand stages, [5] cycles the producer waited for free weight stages, [6] cycles the epilogue
 * waited for accumulators, [7] tiles done) to `host`; returns the number of CTAs (0 when profiling is off). */
int64_t ede_debug_umma_prof(unsigned long long* host, int64_t max_ctas) {
    if (!ede::g_prof_last || !host) return 0;
    cudaDeviceSynchronize();
    size_t n = ede::g_prof_ctas < (size_t)max_ctas ? ede::g_prof_ctas : (size_t)max_ctas;
    cudaMemcpy(host, ede::g_prof_last, n * 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    return (int64_t)n;
}

int ede_umma_available(void) { return ede::get_encode_fn() != nullptr ? 1 : 0; }

int ede_conv2d_umma_config(const EdeConv* a, EdeUmmaCfg* out) {
    EDE_REQUIRE(a != nullptr && out != nullptr, "ede_conv2d_umma_config: null argument");
    EdeConv b = *a;
    if (b.w_passes == 0) b.w_passes = 1;
    ede::umma_config(b, *out);
    return EDE_OK;
}

int ede_conv2d(const EdeConv* a, void* stream) {
    EDE_REQUIRE(a != nullptr, "ede_conv2d: null argument");
    EDE_REQUIRE(a->x && a->y && a->scale && a->shift, "ede_conv2d: null tensor pointer");
    EDE_REQUIRE(a->n > 0 && a->h > 0 && a->w > 0 && a->oh > 0 && a->ow > 0 && a->cin_p > 0 && a->cout_p > 0,
                "ede_conv2d: non-positive shape");
    EDE_REQUIRE(a->kh > 0 && a->kw > 0 && a->sh > 0 && a->sw > 0, "ede_conv2d: bad kernel/stride");
    EDE_REQUIRE((a->oh - 1) * a->sh - a->pt + a->kh - 1 < a->h + a->kh && (a->ow - 1) * a->sw - a->pl + a->kw - 1 < a->w + a->kw,
                "ede_conv2d: output size inconsistent with input size");
    EDE_REQUIRE(a->x_layout == EDE_NHWC || (a->x_dtype == EDE_F32 && a->x_layout == EDE_NCHW),
                "ede_conv2d: NCHW input must be fp32");
    EdeConv b;
    memset(&b, 0, sizeof(b));
    memcpy(&b, a, ede::CONV_KEY_BYTES);
    if (b.w_passes == 0) b.w_passes = 1;
    static const bool cache_on = ede::env_int("EDE_CONV_CACHE", 1) != 0;
    int dev = 0;
    (void)cudaGetDevice(&dev);
    ede::ConvCacheEntry* e = cache_on ? ede::conv_cache_find(b, dev, false) : nullptr;
    if (e == nullptr) {
        EdeUmmaCfg cfg;
        ede::umma_config(b, cfg);
        ede::ConvLaunch tmp;
        if (cfg.eligible) {
            int rc = ede::conv_umma_build(b, cfg, tmp);
            if (rc) return rc;
        }
        if (!cache_on) {
            if (cfg.eligible) return ede::conv_umma_launch(b, tmp, (cudaStream_t)stream);
            return ede::conv_direct_launch(b, (cudaStream_t)stream);
        }
        e = ede::conv_cache_find(b, dev, true);
        e->eligible = cfg.eligible != 0;
        if (e->eligible) e->launch = tmp;
    }
    if (e->eligible) return ede::conv_umma_launch(b, e->launch, (cudaStream_t)stream);
    return ede::conv_direct_launch(b, (cudaStream_t)stream);
}

}  // extern "C"
