// prepack.cu -- weight prepacking behind the C ABI: reference layout (OIHW fp32 conv weight + BatchNorm vectors) ->
// the operand layouts of ede_conv2d (include/ede.h: w_umma, w_direct, scale, shift).  One kernel per layer, so a
// consumer of libede_sm100.so that is not Python can produce every buffer ede_conv2d needs, and the Python engine
// packs its plans with these kernels instead of eager tensor ops.
#include <algorithm>

#include "common.cuh"

namespace ede {

struct PrepackDims {
    int cout, cin, kh, kw, cin_p, cout_p, passes, g;   // g >= 1 taps of a kernel row per tile
    int co_rows, ci_cols, kws;                         // rows / columns of one w_umma tile, kernel-row steps
    int slot;                                          // rows per tap slot: 128 / g (taps on the M rows) or cout_p16 (on N)
};

__device__ __forceinline__ float src_weight(const EdePrepackConv& a, int lane_o, int lane_i, int ky, int kx) {
    if (lane_o >= a.cout_p || lane_i >= a.cin_p) return 0.f;
    const int so = a.out_map ? __ldg(a.out_map + lane_o) : (lane_o < a.cout ? lane_o : -1);
    const int si = a.in_map ? __ldg(a.in_map + lane_i) : (lane_i < a.cin ? lane_i : -1);
    if (so < 0 || si < 0) return 0.f;
    return __ldg(a.w_oihw + (((size_t)so * a.cin + si) * a.kh + ky) * a.kw + kx);
}

// w_umma: bf16 [passes * kh * kws][co_rows][ci_cols]; every element (padding included) is written
__global__ void prepack_umma_kernel(const EdePrepackConv a, const PrepackDims d) {
    const size_t per_pass = (size_t)d.kh * d.kws * d.co_rows * d.ci_cols;
    __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(a.w_umma);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < per_pass; i += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % d.ci_cols);
        size_t r = i / d.ci_cols;
        const int row = (int)(r % d.co_rows);
        const int tile = (int)(r / d.co_rows);
        const int ky = tile / d.kws, gx = tile - ky * d.kws;
        float v = 0.f;
        if (d.g > 1) {
            const int j = row / d.slot, co = row - j * d.slot, kx = gx * d.g + j;
            if (j < d.g && kx < d.kw && co < d.cout_p) v = src_weight(a, co, c, ky, kx);
        } else {
            v = src_weight(a, row, c, ky, gx);
        }
        const __nv_bfloat16 hi = __float2bfloat16_rn(v);
        out[i] = hi;
        if (d.passes == 2) out[per_pass + i] = __float2bfloat16_rn(v - __bfloat162float(hi));
    }
}

// w_direct: fp32 [kh*kw][cin_p][cout_p]
__global__ void prepack_direct_kernel(const EdePrepackConv a, const PrepackDims d) {
    const size_t total = (size_t)d.kh * d.kw * d.cin_p * d.cout_p;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int co = (int)(i % d.cout_p);
        size_t r = i / d.cout_p;
        const int ci = (int)(r % d.cin_p);
        const int tap = (int)(r / d.cin_p);
        a.w_direct[i] = src_weight(a, co, ci, tap / d.kw, tap % d.kw);
    }
}

// folded BatchNorm (fp64 arithmetic): scale = gamma / sqrt(var + eps), shift = beta - mean * scale
// (+ scale * conv bias, in fp32 like the host fold); pad lanes: 0 / 0
__global__ void prepack_affine_kernel(const EdePrepackConv a) {
    const int lane = blockIdx.x * blockDim.x + threadIdx.x;
    if (lane >= a.cout_p) return;
    const int s = a.out_map ? a.out_map[lane] : (lane < a.cout ? lane : -1);
    double scale = 0.0, shift = 0.0;
    if (s >= 0) {
        scale = 1.0;
        if (a.bn_weight) {
            scale = (double)a.bn_weight[s] / sqrt((double)a.bn_var[s] + (double)a.bn_eps);
            shift = (double)a.bn_bias[s] - (double)a.bn_mean[s] * scale;
        }
        float sf = (float)shift;
        if (a.bias) sf = __fadd_rn(sf, __fmul_rn((float)scale, a.bias[s]));   // two fp32 roundings, like the host fold
        shift = sf;
    }
    if (a.scale) a.scale[lane] = (float)scale;
    if (a.shift) a.shift[lane] = (float)shift;
}

static int prepack_dims(const EdePrepackConv& a, PrepackDims& d) {
    EDE_REQUIRE(a.cout > 0 && a.cin > 0 && a.kh > 0 && a.kw > 0, "ede_prepack_conv: non-positive weight shape");
    EDE_REQUIRE(a.cin_p >= 8 && a.cout_p >= 8 && (a.cin_p & 7) == 0 && (a.cout_p & 7) == 0,
                "ede_prepack_conv: lane counts must be positive multiples of 8");
    EDE_REQUIRE(a.in_map || a.cin <= a.cin_p, "ede_prepack_conv: cin exceeds the lane count (no in_map given)");
    EDE_REQUIRE(a.out_map || a.cout <= a.cout_p, "ede_prepack_conv: cout exceeds the lane count (no out_map given)");
    EDE_REQUIRE(a.passes == 1 || a.passes == 2, "ede_prepack_conv: passes must be 1 or 2");
    const bool on_n = a.pack_g >= (EDE_PACK_N | 2) && a.pack_g <= (EDE_PACK_N | 5);
    EDE_REQUIRE(a.pack_g == 0 || a.pack_g == 1 || a.pack_g == 2 || a.pack_g == 4 || on_n,
                "ede_prepack_conv: pack_g must be 0/1, 2, 4 or EDE_PACK_N | 2..5");
    EDE_REQUIRE((a.bn_weight != nullptr) == (a.bn_bias != nullptr) && (a.bn_weight != nullptr) == (a.bn_mean != nullptr) &&
                    (a.bn_weight != nullptr) == (a.bn_var != nullptr),
                "ede_prepack_conv: give all four BatchNorm vectors or none");
    d.cout = a.cout; d.cin = a.cin; d.kh = a.kh; d.kw = a.kw; d.cin_p = a.cin_p; d.cout_p = a.cout_p;
    d.passes = a.passes;
    d.g = on_n ? (a.pack_g & 15) : a.pack_g > 1 ? a.pack_g : 1;
    d.slot = on_n ? round_up(a.cout_p, 16) : 128 / d.g;
    EDE_REQUIRE(d.g == 1 || on_n || a.cout_p <= 128 / d.g, "ede_prepack_conv: %d output lanes do not fit %d taps in 128 rows",
                a.cout_p, d.g);
    EDE_REQUIRE(!on_n || d.g * d.slot <= 256, "ede_prepack_conv: %d taps of %d output lanes exceed 256 UMMA columns", d.g, d.slot);
    d.kws = ceil_div(a.kw, d.g);
    d.co_rows = on_n ? d.g * d.slot : d.g > 1 ? 128 : round_up(a.cout_p, 16);
    d.ci_cols = round_up(a.cin_p, 64);
    return EDE_OK;
}

}  // namespace ede

extern "C" {

int ede_prepack_conv_sizes(const EdePrepackConv* a, int64_t* bytes4, int32_t* dims2) {
    EDE_REQUIRE(a != nullptr && bytes4 != nullptr, "ede_prepack_conv_sizes: null argument");
    ede::PrepackDims d;
    int rc = ede::prepack_dims(*a, d);
    if (rc) return rc;
    bytes4[0] = (int64_t)d.passes * d.kh * d.kws * d.co_rows * d.ci_cols * 2;
    bytes4[1] = (int64_t)d.kh * d.kw * d.cin_p * d.cout_p * 4;
    bytes4[2] = bytes4[3] = (int64_t)d.cout_p * 4;
    if (dims2) { dims2[0] = d.co_rows; dims2[1] = d.ci_cols; }
    return EDE_OK;
}

int ede_prepack_conv_to(const EdePrepackConv* a, void* stream) {
    EDE_REQUIRE(a != nullptr && a->w_oihw != nullptr, "ede_prepack_conv_to: null weight pointer");
    ede::PrepackDims d;
    int rc = ede::prepack_dims(*a, d);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (a->w_umma) {
        const size_t n = (size_t)d.kh * d.kws * d.co_rows * d.ci_cols;
        const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 16);
        ede::prepack_umma_kernel<<<blocks, 256, 0, st>>>(*a, d);
        EDE_LAUNCH_CHECK("prepack_umma_kernel");
    }
    if (a->w_direct) {
        const size_t n = (size_t)d.kh * d.kw * d.cin_p * d.cout_p;
        const int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 16);
        ede::prepack_direct_kernel<<<blocks, 256, 0, st>>>(*a, d);
        EDE_LAUNCH_CHECK("prepack_direct_kernel");
    }
    if (a->scale || a->shift) {
        ede::prepack_affine_kernel<<<ede::ceil_div(a->cout_p, 128), 128, 0, st>>>(*a);
        EDE_LAUNCH_CHECK("prepack_affine_kernel");
    }
    return EDE_OK;
}

int ede_prepack_conv(const EdePrepackConv* a, EdePackedConv** handle, void* stream) {
    EDE_REQUIRE(a != nullptr && handle != nullptr, "ede_prepack_conv: null argument");
    *handle = nullptr;
    int64_t bytes[4];
    int32_t dims[2];
    int rc = ede_prepack_conv_sizes(a, bytes, dims);
    if (rc) return rc;
    EdePackedConv* h = (EdePackedConv*)calloc(1, sizeof(EdePackedConv));
    EDE_REQUIRE(h != nullptr, "ede_prepack_conv: out of host memory");
    // one device allocation, carved into the four buffers (256-byte aligned)
    size_t off[4], total = 0;
    for (int i = 0; i < 4; ++i) { off[i] = total; total += ((size_t)bytes[i] + 255) & ~(size_t)255; }
    char* base = nullptr;
    rc = ede::check_cuda(cudaMalloc((void**)&base, total), "ede_prepack_conv: cudaMalloc");
    if (rc) { free(h); return rc; }
    h->w_umma = base + off[0];
    h->w_direct = (float*)(base + off[1]);
    h->scale = (float*)(base + off[2]);
    h->shift = (float*)(base + off[3]);
    h->w_umma_cout_p = dims[0]; h->w_umma_cin_p = dims[1];
    h->w_pack_g = a->pack_g > 1 ? a->pack_g : 1; h->w_passes = a->passes;
    h->device_bytes = (int64_t)total;
    EdePrepackConv b = *a;
    b.w_umma = h->w_umma; b.w_direct = h->w_direct; b.scale = h->scale; b.shift = h->shift;
    rc = ede_prepack_conv_to(&b, stream);
    if (rc) { cudaFree(base); free(h); return rc; }
    *handle = h;
    return EDE_OK;
}

int ede_free(EdePackedConv* h) {
    if (h == nullptr) return EDE_OK;
    int rc = ede::check_cuda(cudaFree(h->w_umma), "ede_free");      // w_umma is the base of the single allocation
    free(h);
    return rc;
}

}  // extern "C"
