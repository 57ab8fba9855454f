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

#include <algorithm>

#include "common.cuh"

namespace ede {

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(bar), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1, int c2,
                                            int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
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
constexpr int MAX_STAGES = 8;

struct alignas(64) UmmaParams {
    CUtensorMap tm_a;     // activations: (C, W, H, N) bf16
    CUtensorMap tm_b;     // weights: (Cin_p, Cout_p, taps) bf16
    int oh, ow, tiles_w, tiles_h;
    int a_mode, ck_bytes, n_chunks, kh, kw, sh, sw, pt, pl, w_passes;
    int bw, bh, pitch, box_bytes;
    int m_tiles, n_pad;
    int a_stages, b_stages, a_stage_bytes, b_stage_bytes, b_tx_bytes;
    int layout_type, sbo_bytes, tmem_cols, base_off_mode, flat;
    uint32_t idesc;
    void* y;
    const void* res;
    const float* scale;
    const float* shift;
    int y_cs, y_co, res_cs, res_co, cout_p, act, act_c_end, y_dtype;
};

__device__ __forceinline__ uint64_t make_smem_desc(uint32_t addr, const UmmaParams& p) {
    uint64_t d = (uint64_t)((addr & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;                                  // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)((uint32_t)p.sbo_bytes >> 4) << 32;       // stride between 8-row groups
    d |= (uint64_t)1 << 46;                                  // descriptor version (Blackwell)
    if (p.base_off_mode) d |= (uint64_t)((addr >> 7) & 7u) << 49;
    d |= (uint64_t)p.layout_type << 61;
    return d;
}

template <typename TY>
__device__ __forceinline__ void epilogue_store8(const UmmaParams& p, long long pix, int c, const uint32_t* acc) {
    float v[8], sc[8], sf[8];
    *reinterpret_cast<float4*>(sc) = __ldg(reinterpret_cast<const float4*>(p.scale + c));
    *reinterpret_cast<float4*>(sc + 4) = __ldg(reinterpret_cast<const float4*>(p.scale + c + 4));
    *reinterpret_cast<float4*>(sf) = __ldg(reinterpret_cast<const float4*>(p.shift + c));
    *reinterpret_cast<float4*>(sf + 4) = __ldg(reinterpret_cast<const float4*>(p.shift + c + 4));
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = fmaf(__uint_as_float(acc[j]), sc[j], sf[j]);
    if (p.res) {
        float r[8];
        load8<TY>(reinterpret_cast<const TY*>(p.res) + pix * p.res_cs + p.res_co + c, r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] += r[j];
    }
    if (p.act != EDE_ACT_NONE) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (c + j < p.act_c_end) v[j] = apply_act(v[j], p.act);
    }
    store8<TY>(reinterpret_cast<TY*>(p.y) + pix * p.y_cs + p.y_co + c, v);
}

__global__ void __launch_bounds__(192) conv_umma_kernel(const __grid_constant__ UmmaParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_base = smem_base;
    const uint32_t b_base = a_base + (uint32_t)p.a_stages * p.a_stage_bytes;
    const uint32_t bar_base = b_base + (uint32_t)p.b_stages * p.b_stage_bytes;
    // barriers: a_full[8] a_empty[8] b_full[8] b_empty[8] tmem_full, then the TMEM base slot
    auto a_full = [&](int s) { return bar_base + 8u * s; };
    auto a_empty = [&](int s) { return bar_base + 8u * (MAX_STAGES + s); };
    auto b_full = [&](int s) { return bar_base + 8u * (2 * MAX_STAGES + s); };
    auto b_empty = [&](int s) { return bar_base + 8u * (3 * MAX_STAGES + s); };
    const uint32_t tmem_full = bar_base + 8u * (4 * MAX_STAGES);
    const uint32_t tmem_slot = tmem_full + 8u;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // tile coordinates
    const int tiles_per_img = p.tiles_w * p.tiles_h;
    const int img = blockIdx.x / tiles_per_img;
    const int t_in = blockIdx.x - img * tiles_per_img;
    const int th = t_in / p.tiles_w, tw = t_in - th * p.tiles_w;
    const int ntile = blockIdx.y;
    const int sp_taps = p.kh * p.kw;              // spatial taps
    const int taps = sp_taps * p.w_passes;        // weight tiles per channel chunk ([w_hi taps ; w_lo taps])
    const int tile_w_pix = p.a_mode == 0 ? p.bw : p.bw * p.m_tiles;   // output columns covered by this CTA
    const int ow0 = tw * tile_w_pix, oh0 = th * p.bh;

    if (threadIdx.x == 0) {
        for (int s = 0; s < MAX_STAGES; ++s) {
            mbar_init(a_full(s), 1); mbar_init(a_empty(s), 1);
            mbar_init(b_full(s), 1); mbar_init(b_empty(s), 1);
        }
        mbar_init(tmem_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                     "r"((uint32_t)p.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_addr;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_addr) : "r"(tmem_slot) : "memory");

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int a_st = 0, a_ph = 0, b_st = 0, b_ph = 0;
            const int ck = p.ck_bytes >> 1;
            for (int chunk = 0; chunk < p.n_chunks; ++chunk) {
                if (p.a_mode == 0) {
                    mbar_wait(a_empty(a_st), a_ph ^ 1);
                    mbar_expect_tx(a_full(a_st), (uint32_t)p.box_bytes);
                    tma_load_4d(a_base + a_st * p.a_stage_bytes, &p.tm_a, a_full(a_st), chunk * ck, ow0 - p.pl,
                                oh0 - p.pt, img);
                    if (++a_st == p.a_stages) { a_st = 0; a_ph ^= 1; }
                }
                for (int tap = 0; tap < taps; ++tap) {
                    if (p.a_mode != 0) {
                        const int stap = tap >= sp_taps ? tap - sp_taps : tap;
                        const int ky = stap / p.kw, kx = stap - ky * p.kw;
                        mbar_wait(a_empty(a_st), a_ph ^ 1);
                        mbar_expect_tx(a_full(a_st), (uint32_t)(p.m_tiles * 128 * p.ck_bytes));
                        for (int m = 0; m < p.m_tiles; ++m)
                            tma_load_4d(a_base + a_st * p.a_stage_bytes + m * 128 * p.ck_bytes, &p.tm_a, a_full(a_st),
                                        chunk * ck, (ow0 + m * p.bw) * p.sw - p.pl + kx, oh0 * p.sh - p.pt + ky,
                                        img);
                        if (++a_st == p.a_stages) { a_st = 0; a_ph ^= 1; }
                    }
                    mbar_wait(b_empty(b_st), b_ph ^ 1);
                    mbar_expect_tx(b_full(b_st), (uint32_t)p.b_tx_bytes);
                    tma_load_3d(b_base + b_st * p.b_stage_bytes, &p.tm_b, b_full(b_st), chunk * ck,
                                ntile * p.n_pad, tap);
                    if (++b_st == p.b_stages) { b_st = 0; b_ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            int a_st = 0, a_ph = 0, b_st = 0, b_ph = 0;
            const int ksteps = p.ck_bytes >> 5;        // 32 bytes of K (16 bf16) per tcgen05.mma
            uint32_t accum = 0;
            for (int chunk = 0; chunk < p.n_chunks; ++chunk) {
                if (p.a_mode == 0) mbar_wait(a_full(a_st), a_ph);
                for (int tap = 0; tap < taps; ++tap) {
                    if (p.a_mode != 0) mbar_wait(a_full(a_st), a_ph);
                    mbar_wait(b_full(b_st), b_ph);
                    tc_fence_after();
                    int shift_rows = 0;
                    if (p.a_mode == 0) {
                        const int stap = tap >= sp_taps ? tap - sp_taps : tap;
                        const int ky = stap / p.kw, kx = stap - ky * p.kw;
                        shift_rows = ky * p.pitch + kx;
                    }
                    const uint32_t a_addr = a_base + a_st * p.a_stage_bytes + shift_rows * p.ck_bytes;
                    const uint32_t b_addr = b_base + b_st * p.b_stage_bytes;
                    for (int m = 0; m < p.m_tiles; ++m) {
                        uint32_t acc_m = accum;
                        for (int k = 0; k < ksteps; ++k) {
                            const uint64_t ad = make_smem_desc(a_addr + m * 128 * p.ck_bytes + k * 32, p);
                            const uint64_t bd = make_smem_desc(b_addr + k * 32, p);
                            tc_mma_bf16(tmem_addr + m * p.n_pad, ad, bd, p.idesc, acc_m);
                            acc_m = 1;
                        }
                    }
                    accum = 1;
                    tc_commit(b_empty(b_st));
                    if (++b_st == p.b_stages) { b_st = 0; b_ph ^= 1; }
                    if (p.a_mode != 0) {
                        tc_commit(a_empty(a_st));
                        if (++a_st == p.a_stages) { a_st = 0; a_ph ^= 1; }
                    }
                }
                if (p.a_mode == 0) {
                    tc_commit(a_empty(a_st));
                    if (++a_st == p.a_stages) { a_st = 0; a_ph ^= 1; }
                }
            }
            tc_commit(tmem_full);
        }
        __syncwarp();
    } else {
        // ===================== epilogue (warps 2..5) =====================
        const int quarter = warp & 3;                 // TMEM lanes [32q, 32q+32) are visible to warp%4 == q
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        const int r = quarter * 32 + lane;
        for (int m = 0; m < p.m_tiles; ++m) {
            // which output pixel does accumulator row (m, r) hold?
            int oh, ow;
            bool valid;
            if (p.a_mode == 0) {
                const int f = m * 128 + r;
                const int fr = f / p.pitch, fc = f - fr * p.pitch;
                oh = oh0 + fr; ow = ow0 + fc;
                valid = fc < p.bw && fr < p.bh;
            } else {
                const int fr = r / p.bw, fc = r - fr * p.bw;
                oh = oh0 + fr; ow = ow0 + m * p.bw + fc;
                valid = true;
            }
            valid = valid && oh < p.oh && ow < p.ow;
            const long long pix = ((long long)img * p.oh + oh) * p.ow + ow;
            for (int c0 = 0; c0 < p.n_pad; c0 += 16) {
                uint32_t acc[16];
                tc_ld16(tmem_addr + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(m * p.n_pad + c0), acc);
                tc_ld_wait();
                if (valid) {
                    const int c = ntile * p.n_pad + c0;
                    if (p.y_dtype == EDE_BF16) {
                        if (c < p.cout_p) epilogue_store8<__nv_bfloat16>(p, pix, c, acc);
                        if (c + 8 < p.cout_p) epilogue_store8<__nv_bfloat16>(p, pix, c + 8, acc + 8);
                    } else if (p.y_dtype == EDE_BF16X2) {
                        if (c < p.cout_p) epilogue_store8<bf16x2_t>(p, pix, c, acc);
                        if (c + 8 < p.cout_p) epilogue_store8<bf16x2_t>(p, pix, c + 8, acc + 8);
                    } else {
                        if (c < p.cout_p) epilogue_store8<float>(p, pix, c, acc);
                        if (c + 8 < p.cout_p) epilogue_store8<float>(p, pix, c + 8, acc + 8);
                    }
                }
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
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        (void)cudaGetLastError();
    }
    return fn;
}

static int env_int(const char* name, int dflt) {
    const char* s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}

constexpr int SMEM_BUDGET = 200 * 1024;      // operand rings; barriers + alignment slack on top
constexpr int SMEM_EXTRA = 1024 + 512;

static bool umma_eligible(const EdeConv& a) {
    if (a.force_direct || a.w_umma == nullptr) return false;
    if (a.x_dtype != EDE_BF16 || a.x_layout != EDE_NHWC || a.relu_in) return false;
    if (a.sh != a.sw || (a.sh != 1 && a.sh != 2)) return false;
    if (a.kh * a.kw > 64 || a.kh > 8 || a.kw > 8) return false;
    if (a.w_passes != 1 && a.w_passes != 2) return false;
    if ((a.x_cs | a.x_co | a.y_cs | a.y_co | a.cin_p | a.cout_p) & 7) return false;
    if (a.res && ((a.res_cs | a.res_co) & 7)) return false;
    if (a.w_umma_cin_p & 7) return false;
    return true;
}

static int pick_ck(int cin_p) {
    // channels per K chunk: trade zero-padded K against TMA/MMA granularity
    const int cand[3] = {64, 32, 16};
    const double pen[3] = {1.00, 1.04, 1.12};
    int best = 64;
    double best_cost = 1e30;
    for (int i = 0; i < 3; ++i) {
        double cost = (double)round_up(cin_p, cand[i]) * pen[i];
        if (cost < best_cost) { best_cost = cost; best = cand[i]; }
    }
    return best;
}

static int pow2_cols(int c) {
    int v = 32;
    while (v < c) v <<= 1;
    return v;
}

constexpr int SMEM_2CTA = 98 * 1024;          // operand budget that still lets two CTAs share an SM

// number of ring stages: prefer a footprint that allows 2 CTAs/SM (one CTA's epilogue then
// overlaps the other's MMAs), else use the whole SM
static int fit_stages(int fixed_bytes, int stage_bytes, int want) {
    int s = (SMEM_2CTA - fixed_bytes) / stage_bytes;
    if (s < want) s = (SMEM_BUDGET - fixed_bytes) / stage_bytes;
    if (s > MAX_STAGES) s = MAX_STAGES;
    return s;
}

static void umma_config(const EdeConv& a, EdeUmmaCfg& c) {
    memset(&c, 0, sizeof(c));
    if (!umma_eligible(a)) return;
    const int force_mode = env_int("EDE_UMMA_MODE", -1);
    const int force_ck = env_int("EDE_UMMA_CK", 0);
    const int ck = (force_ck == 16 || force_ck == 32 || force_ck == 64) ? force_ck : pick_ck(a.cin_p);
    const int ck_bytes = ck * 2;
    const int n16 = round_up(a.cout_p, 16);
    const int n_tiles = ceil_div(n16, 128);
    const int n_pad = round_up(ceil_div(n16, n_tiles), 16);
    const int b_bytes = round_up(n_pad * ck_bytes, 1024);
    const bool is_1x1 = a.kh == 1 && a.kw == 1 && a.sh == 1 && a.pt == 0 && a.pl == 0 && a.oh == a.h && a.ow == a.w;
    const int OH = is_1x1 ? 1 : a.oh;
    const long long OWl = is_1x1 ? (long long)a.n * a.h * a.w : a.ow;
    if (OWl > 0x7fffffffLL) return;
    const int OW = (int)OWl;
    const int max_t = 512 / n_pad;

    EdeUmmaCfg best{};
    double best_score = -1;
    // ---- per-tap candidates
    for (int bw = 128; bw >= 1; bw >>= 1) {
        const int bh = 128 / bw;
        if (bw * a.sw > 256 || bh * a.sh > 256) continue;
        if (is_1x1 && bh != 1) continue;
        for (int T = 4; T >= 1; T >>= 1) {
            if (T > max_t) continue;
            const int a_bytes = T * 128 * ck_bytes;
            const int stages = fit_stages(0, a_bytes + b_bytes, 4);
            if (stages < 3) continue;
            const int tiles_w = ceil_div(OW, bw * T), tiles_h = ceil_div(OH, bh);
            const double frac = (double)OH * OW / ((double)tiles_w * tiles_h * T * 128);
            const double score = frac + 0.004 * T + 0.01 * (bw >= 8) + 0.01 * (bw >= 16) -
                                 0.02 * (pow2_cols(T * n_pad) > 256);
            if (score > best_score) {
                best_score = score;
                best.a_mode = 1; best.bh = bh; best.bw = bw; best.pitch = bw; best.box_rows = bh;
                best.m_tiles = T; best.a_stages = best.b_stages = stages;
                best.a_stage_bytes = a_bytes; best.tiles_w = tiles_w; best.tiles_h = tiles_h;
                best.useful_frac = frac;
            }
        }
    }
    // ---- resident candidates (stride 1, k x k with k > 1)
    if (a.sh == 1 && a.kh * a.kw >= 9 && force_mode != 1 && !is_1x1) {
        EdeUmmaCfg rb{};
        double rb_score = -1;
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
                const int b_stages = fit_stages(2 * a_bytes, b_bytes, 4);
                if (b_stages < 3) continue;
                const int tiles_w = ceil_div(OW, bw), tiles_h = ceil_div(OH, bh);
                const double frac = (double)OH * OW / ((double)tiles_w * tiles_h * T * 128);
                const double amp = (double)box_rows * P / ((double)bh * bw);   // halo read amplification
                const double score = frac - 0.01 * amp - 0.02 * (pow2_cols(T * n_pad) > 256) -
                                     0.02 * (2 * a_bytes + 4 * b_bytes > SMEM_2CTA);
                if (score > rb_score) {
                    rb_score = score;
                    rb.a_mode = 0; rb.bh = bh; rb.bw = bw; rb.pitch = P; rb.box_rows = box_rows;
                    rb.m_tiles = T; rb.a_stages = 2; rb.b_stages = b_stages; rb.a_stage_bytes = a_bytes;
                    rb.tiles_w = tiles_w; rb.tiles_h = tiles_h; rb.useful_frac = frac;
                }
            }
        }
        if (rb_score > 0 && (force_mode == 0 || rb.useful_frac >= 0.8 * best.useful_frac)) best = rb;
    }
    if (best.m_tiles == 0) return;
    c = best;
    c.eligible = 1;
    c.ck = ck; c.swizzle = ck_bytes;
    c.n_pad = n_pad; c.n_tiles = n_tiles;
    c.b_stage_bytes = b_bytes;
    c.tmem_cols = pow2_cols(c.m_tiles * n_pad);
    c.smem_bytes = c.a_stages * c.a_stage_bytes + c.b_stages * c.b_stage_bytes + SMEM_EXTRA;
    c.flat = is_1x1;
    // real MACs / issued MACs also counts channel padding
    const double kfrac = (double)a.cin_p / round_up(a.cin_p, ck);
    const double nfrac = (double)a.cout_p / ((double)n_tiles * n_pad);
    c.useful_frac *= kfrac * nfrac;
}

int conv_direct_launch(const EdeConv& a, cudaStream_t st);

static int conv_umma_launch(const EdeConv& a, const EdeUmmaCfg& c, cudaStream_t st) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        set_error("ede_conv2d: cuTensorMapEncodeTiled driver entry point unavailable");
        return EDE_E_DRIVER;
    }
    UmmaParams p;
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
        cuuint64_t dims[3] = {(cuuint64_t)a.w_umma_cin_p, (cuuint64_t)a.w_umma_cout_p,
                              (cuuint64_t)(a.kh * a.kw * a.w_passes)};
        cuuint64_t strides[2] = {(cuuint64_t)a.w_umma_cin_p * 2, (cuuint64_t)a.w_umma_cin_p * 2 * a.w_umma_cout_p};
        cuuint32_t box[3] = {(cuuint32_t)c.ck, (cuuint32_t)c.n_pad, 1};
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
    p.a_mode = c.a_mode; p.ck_bytes = c.ck * 2; p.n_chunks = ceil_div(a.cin_p, c.ck);
    p.w_passes = a.w_passes;
    p.kh = a.kh; p.kw = a.kw; p.sh = a.sh; p.sw = a.sw; p.pt = a.pt; p.pl = a.pl;
    p.bw = c.bw; p.bh = c.bh; p.pitch = c.pitch; p.box_bytes = c.pitch * c.box_rows * c.ck * 2;
    p.m_tiles = c.m_tiles; p.n_pad = c.n_pad;
    p.a_stages = c.a_stages; p.b_stages = c.b_stages;
    p.a_stage_bytes = c.a_stage_bytes; p.b_stage_bytes = c.b_stage_bytes; p.b_tx_bytes = c.n_pad * c.ck * 2;
    p.layout_type = c.swizzle == 128 ? 2 : c.swizzle == 64 ? 4 : 6;
    p.sbo_bytes = 8 * c.ck * 2;
    p.tmem_cols = c.tmem_cols;
    p.base_off_mode = env_int("EDE_UMMA_BASE_OFFSET", 0);
    p.flat = c.flat;
    p.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(c.n_pad >> 3) << 17) | ((128u >> 4) << 24);
    p.y = a.y; p.res = a.res; p.scale = a.scale; p.shift = a.shift;
    p.y_cs = a.y_cs; p.y_co = a.y_co; p.res_cs = a.res_cs; p.res_co = a.res_co;
    p.cout_p = a.cout_p; p.act = a.act; p.act_c_end = a.act_c_end; p.y_dtype = a.y_dtype;

    static bool smem_attr_set[64] = {};
    int dev = 0;
    (void)cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !smem_attr_set[dev]) {
        int rc = check_cuda(cudaFuncSetAttribute(conv_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 227 * 1024),
                            "cudaFuncSetAttribute(conv_umma_kernel)");
        if (rc) return rc;
        smem_attr_set[dev] = true;
    }
    const int n_img = c.flat ? 1 : a.n;
    dim3 grid((unsigned)(c.tiles_w * c.tiles_h * n_img), (unsigned)c.n_tiles);
    conv_umma_kernel<<<grid, 192, c.smem_bytes, st>>>(p);
    EDE_LAUNCH_CHECK("conv_umma_kernel");
    return EDE_OK;
}

}  // namespace ede

extern "C" {

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
    EdeConv b = *a;
    if (b.w_passes == 0) b.w_passes = 1;
    EdeUmmaCfg cfg;
    ede::umma_config(b, cfg);
    if (cfg.eligible) return ede::conv_umma_launch(b, cfg, (cudaStream_t)stream);
    return ede::conv_direct_launch(b, (cudaStream_t)stream);
}

}  // extern "C"
