// stem.cu -- the network's first convolution, straight from the fp32 NCHW input image.
// replaces: ResNet conv1 7x7 s2 p3 (3 -> 64) + bn1 + ReLU (ReSIDE/models/resnet.py:103-106) and the EfficientNet
//   stem Conv2dStaticSamePadding 3x3 s2 (3 -> 32/48) + bn0, and with them the NCHW -> NHWC
//   layout pass that used to precede them.
// With 3 input channels the GEMM K is 27 / 147: far too thin for the tensor cores (the implicit-GEMM form padded
// every pixel to a 128-byte row and was TMA-bound), so this is a CUDA-core kernel shaped for bandwidth:
//   * a block stages its input patch (3 x rows x cols, zero padded, columns de-interleaved by stride phase so that
//     stride-2 reads are bank-conflict free) and all weights in shared memory;
//   * a thread owns FOUR output pixels and HALF the output channels (4 x COUT/2 fp32 accumulators; the two channel
//     halves are warp-uniform): per tap it reads four inputs and COUT/8 broadcast weight vectors for 2*COUT FMAs --
//     shared-memory reads, not FMAs, bounded the 2-pixel x all-channel form (a broadcast LDS.128 still moves 512 B
//     into the register file);
//   * folded BN + activation, then each pixel's channels leave as contiguous 32-byte vectors in NHWC.
// fp32 inputs and weights, fp32 accumulation: exact in every precision mode.
#include <algorithm>

#include "common.cuh"

namespace ede {

struct StemTile {
    int lw, th;            // column quads per tile row (tile width = 4*lw), tile rows; lw*th <= 128
    int idx_w;             // staged columns per stride phase
    int ih;                // staged input rows
};

// 4-byte asynchronous copy global -> shared; src_bytes = 0 zero-fills (padding) without touching global memory
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc, int src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
                 "l"(gsrc), "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Persistent blocks walk the tiles.  The input patch of tile i+1 streams into the second staging buffer with
// cp.async (every load of the patch in flight at once, padding zero-filled by the copy itself) while tile i is being
// computed; the weights are staged once per block.  Before this the kernel staged with plain loads, one dependent
// batch after the other in front of every tile, and at 1-2 resident blocks per SM (109-190 registers for the
// 2 x COUT accumulators) nothing hid that latency: 1.25 ms for the ResNet stem at batch 64.
template <typename TY, int K, int S, int COUT>
__global__ void __launch_bounds__(256, (COUT <= 32 ? 2 : 1)) stem_conv_kernel(const EdeStem a, const StemTile tl, int tiles_x, int tiles_y) {
    extern __shared__ float sm[];
    const int cin = a.cin;
    float* const ws = sm;                                        // [K*K*cin][COUT]
    const int xs_elems = cin * tl.ih * S * tl.idx_w;
    float* const xs0 = sm + K * K * cin * COUT;                  // 2 x [cin][ih][S][idx_w]
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    pdl_trigger();

    for (int i = t; i < K * K * cin * COUT; i += 256) {
        const int r = i / COUT, c = i - r * COUT;
        ws[i] = c < a.cout_p ? a.w[(long long)r * a.cout_p + c] : 0.f;
    }
    pdl_wait();                                                  // (the weights above are static; x and y are not)
    const int iw_span = tl.idx_w * S;
    const int per_img = tiles_x * tiles_y, ntiles = per_img * a.n;

    // warps own staged rows (channel, input row), lanes sweep the columns, de-interleaved by stride phase
    auto stage = [&](int tile, float* xs) {
        const int n = tile / per_img, rem = tile - n * per_img;
        const int by = rem / tiles_x, bx = rem - by * tiles_x;
        const int ih_base = by * tl.th * S - a.pt, iw_base = bx * (4 * tl.lw) * S - a.pl;
        const int nrows = cin * tl.ih;
        for (int r = warp; r < nrows; r += 8) {
            const int ci = r / tl.ih, hl = r - ci * tl.ih;
            const int ihh = ih_base + hl;
            const bool row_ok = ihh >= 0 && ihh < a.h;
            const float* __restrict__ src = a.x + (((long long)n * cin + ci) * a.h + (row_ok ? ihh : 0)) * a.w_;
            float* __restrict__ dst = xs + (ci * tl.ih + hl) * S * tl.idx_w;
            for (int il = lane; il < iw_span; il += 32) {
                const int iww = iw_base + il;
                const bool ok = row_ok && iww >= 0 && iww < a.w_;
                cp_async4(dst + (il % S) * tl.idx_w + il / S, src + (ok ? iww : 0), ok ? 4 : 0);
            }
        }
    };

    constexpr int PX = 4, CT = COUT / 2;                         // pixels and channels per thread
    const int half = t >> 7, tt = t & 127;                       // channel half (warp-uniform), pixel-quad index
    const int row = tt / tl.lw, cp = tt - row * tl.lw;
    int buf = 0;
    if ((int)blockIdx.x < ntiles) stage(blockIdx.x, xs0);
    cp_async_commit();
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, buf ^= 1) {
        const int next = tile + gridDim.x;
        if (next < ntiles) stage(next, xs0 + (buf ^ 1) * xs_elems);
        cp_async_commit();
        cp_async_wait<1>();                                      // this tile's patch has landed (the next may be in flight)
        __syncthreads();
        const float* const xs = xs0 + buf * xs_elems;
        const int n = tile / per_img, rem = tile - n * per_img;
        const int by = rem / tiles_x, bx = rem - by * tiles_x;
        const int oh = by * tl.th + row, ow0 = bx * (PX * tl.lw);
        if (row < tl.th && oh < a.oh) {
            unsigned long long acc[PX][CT / 2];                  // channel pairs: packed fp32 FMAs (FFMA2)
#pragma unroll
            for (int j = 0; j < PX; ++j)
#pragma unroll
                for (int c = 0; c < CT / 2; ++c) acc[j][c] = 0ull;
            for (int ci = 0; ci < cin; ++ci) {
                // 7x7: the kernel rows stay a rolled loop -- fully unrolled the 49 taps are ~120 KB of straight-line
                // code and the 8 resident warps stalled on instruction fetch (ncu: no_instruction 1.8 per issue)
#pragma unroll(K > 3 ? 1 : K)
                for (int ky = 0; ky < K; ++ky) {
                    const float* xr = xs + ((ci * tl.ih + row * S + ky) * S) * tl.idx_w;
#pragma unroll
                    for (int kx = 0; kx < K; ++kx) {
                        const float* xp = xr + (kx % S) * tl.idx_w + kx / S + cp;
                        unsigned long long x[PX];
#pragma unroll
                        for (int j = 0; j < PX; ++j) {
                            const float xv = xp[j * tl.lw];
                            x[j] = pack_f32x2(xv, xv);
                        }
                        const float4* wv = reinterpret_cast<const float4*>(ws + ((ky * K + kx) * cin + ci) * COUT + half * CT);
#pragma unroll
                        for (int c4 = 0; c4 < CT / 4; ++c4) {
                            const float4 w = wv[c4];
                            const unsigned long long w01 = pack_f32x2(w.x, w.y), w23 = pack_f32x2(w.z, w.w);
#pragma unroll
                            for (int j = 0; j < PX; ++j) {
                                ffma2(acc[j][2 * c4], x[j], w01);
                                ffma2(acc[j][2 * c4 + 1], x[j], w23);
                            }
                        }
                    }
                }
            }
            const int c_base = half * CT;
#pragma unroll
            for (int j = 0; j < PX; ++j) {
                const int ow = ow0 + cp + j * tl.lw;
                if (ow >= a.ow) continue;
                TY* yrow = reinterpret_cast<TY*>(a.y) + (((long long)n * a.oh + oh) * a.ow + ow) * a.y_cs + a.y_co;
#pragma unroll
                for (int c = 0; c < CT; c += 8) {
                    if (c_base + c >= a.cout_p) break;
                    float v[8];
#pragma unroll
                    for (int q = 0; q < 8; q += 2) unpack_f32x2(acc[j][(c + q) / 2], v[q], v[q + 1]);
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        v[q] = apply_act(fmaf(v[q], __ldg(a.scale + c_base + c + q), __ldg(a.shift + c_base + c + q)), a.act);
                    store8<TY>(yrow + c_base + c, v);
                }
            }
        }
        __syncthreads();                                         // everyone is done with `buf` before it is refilled
    }
    cp_async_wait<0>();
}

template <typename TY, int K, int S, int COUT>
static int stem_launch_t(const EdeStem& a, cudaStream_t st) {
    StemTile tl;
    const int wt = ceil_div(a.ow, 64);                           // tiles across: widest tile is 64 columns
    tl.lw = std::min(16, ceil_div(ceil_div(a.ow, wt), 4));
    tl.th = std::max(1, std::min(128 / tl.lw, a.oh));
    tl.idx_w = 4 * tl.lw + (K - 1) / S + 1;
    tl.ih = tl.th * S + K - S;
    const size_t smem = ((size_t)K * K * a.cin * COUT + 2 * (size_t)a.cin * tl.ih * S * tl.idx_w) * sizeof(float);
    if (smem > 200 * 1024) return 1;
    static bool attr_set[64] = {};
    int dev = 0;
    (void)cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!attr_set[dev]) {
        if (cudaFuncSetAttribute(stem_conv_kernel<TY, K, S, COUT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 200 * 1024) != cudaSuccess) {
            (void)cudaGetLastError();
            return 1;
        }
        attr_set[dev] = true;
    }
    const int tiles_x = ceil_div(a.ow, 4 * tl.lw), tiles_y = ceil_div(a.oh, tl.th);
    const long long ntiles = (long long)tiles_x * tiles_y * a.n;
    if (ntiles > 0x7fffffffLL) return 1;
    static int blocks_per_sm[64] = {};
    if (!blocks_per_sm[dev]) {
        int b = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, stem_conv_kernel<TY, K, S, COUT>, 256, smem) != cudaSuccess) {
            (void)cudaGetLastError();
            b = 1;
        }
        blocks_per_sm[dev] = b > 0 ? b : 1;
    }
    int sms = 148;
    (void)cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long resident = (long long)sms * blocks_per_sm[dev];
    dim3 grid((unsigned)(ntiles < resident ? ntiles : resident));
    if (launch_pdl(stem_conv_kernel<TY, K, S, COUT>, grid, dim3(256), smem, st, a, tl, tiles_x, tiles_y) != cudaSuccess) {
        (void)cudaGetLastError();
        return 1;
    }
    return 0;
}

template <typename TY>
static int stem_launch(const EdeStem& a, cudaStream_t st) {
    const int key = a.k * 1000 + a.s * 100 + round_up(a.cout_p, 16);
    switch (key) {
        case 3232: return stem_launch_t<TY, 3, 2, 32>(a, st);    // EfficientNet-B0 stem
        case 3248: return stem_launch_t<TY, 3, 2, 48>(a, st);    // EfficientNet-B4 stem
        case 3264: return stem_launch_t<TY, 3, 2, 64>(a, st);
        case 7264: return stem_launch_t<TY, 7, 2, 64>(a, st);    // ResNet conv1
        default: return 1;
    }
}

}  // namespace ede

using namespace ede;

extern "C" {

int ede_stem_supported(int32_t k, int32_t s, int32_t cin, int32_t cout_p) {
    const int key = k * 1000 + s * 100 + (cout_p + 15) / 16 * 16;
    return cin >= 1 && cin <= 4 && (key == 3232 || key == 3248 || key == 3264 || key == 7264) ? 1 : 0;
}

int ede_stem_conv(const EdeStem* a, void* stream) {
    EDE_REQUIRE(a && a->x && a->y && a->w && a->scale && a->shift, "ede_stem_conv: null argument");
    EDE_REQUIRE(a->n > 0 && a->h > 0 && a->w_ > 0 && a->oh > 0 && a->ow > 0, "ede_stem_conv: bad shape");
    EDE_REQUIRE(((a->cout_p | a->y_cs | a->y_co) & 7) == 0, "ede_stem_conv: channel counts must be multiples of 8");
    EDE_REQUIRE(ede_stem_supported(a->k, a->s, a->cin, a->cout_p), "ede_stem_conv: unsupported stem %dx%d s%d %d->%d",
                a->k, a->k, a->s, a->cin, a->cout_p);
    int rc = 1;
    EDE_DISPATCH_DTYPE(a->y_dtype, TY, (rc = stem_launch<TY>(*a, (cudaStream_t)stream)));
    EDE_REQUIRE(rc == 0, "ede_stem_conv: launch configuration failed");
    EDE_LAUNCH_CHECK("stem_conv_kernel");
    return EDE_OK;
}

}  // extern "C"
