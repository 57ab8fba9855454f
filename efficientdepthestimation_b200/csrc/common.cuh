// common.cuh -- shared device/host helpers for libede_sm100.so (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdlib.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ede.h"

namespace ede {

// ---------------------------------------------------------------- error plumbing (api.cu)
void set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
void count_launch(int n = 1);

#define EDE_REQUIRE(cond, ...)                                   \
    do {                                                         \
        if (!(cond)) {                                           \
            ::ede::set_error(__VA_ARGS__);                       \
            return EDE_E_ARG;                                    \
        }                                                        \
    } while (0)

#define EDE_LAUNCH_CHECK(what)                                                  \
    do {                                                                        \
        ::ede::count_launch();                                                  \
        int _rc = ::ede::check_cuda(cudaGetLastError(), what);                  \
        if (_rc) return _rc;                                                    \
    } while (0)

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline int round_up(int a, int b) { return ceil_div(a, b) * b; }
static inline size_t dtype_size(int dt) { return dt == EDE_BF16 ? 2 : 4; }

// EDE_BF16X2 ("split bf16"): one logical lane = 4 bytes; every group of 8 logical lanes is stored as
// [8 x bf16 hi][8 x bf16 lo] (32 bytes) with value = hi + lo (16 mantissa bits).  Seen as a plain bf16
// tensor with twice the lanes it feeds the tensor cores directly: weights laid out [w,w] make one
// K=16 MMA compute (hi+lo).w, so the same TMA/UMMA path runs at near-fp32 operand accuracy.
struct bf16x2_t { __nv_bfloat16 hi, lo; };
// EDE_F32_TO_BF16X2 (depthwise convs only): the INPUT lanes are plain fp32 (the expand conv of an MBConv block writes
// fp32 -- the same 4 bytes per lane -- so that neither its epilogue splits nor the depthwise kernel recombines), the
// OUTPUT lanes are split bf16 (the project conv reads them as a tensor-core operand).  A tag type with the layout of
// the output lane: loads go through the float overloads, stores through the split ones.
struct f32in_t : bf16x2_t {};

// ---------------------------------------------------------------- typed element access
template <typename T> struct Elt;
template <> struct Elt<float> {
    static __device__ __forceinline__ float ld(const float* p) { return *p; }
    static __device__ __forceinline__ void st(float* p, float v) { *p = v; }
};
template <> struct Elt<__nv_bfloat16> {
    static __device__ __forceinline__ float ld(const __nv_bfloat16* p) { return __bfloat162float(*p); }
    static __device__ __forceinline__ void st(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};

template <> struct Elt<bf16x2_t> {
    // scalar access to logical lane c of a row: p points at the lane's 4-byte slot index (row base + c)
    static __device__ __forceinline__ float ld(const bf16x2_t* p) {
        // slot index -> group/offset relative to the 32-byte aligned group start
        const uintptr_t a = reinterpret_cast<uintptr_t>(p);
        const __nv_bfloat16* g = reinterpret_cast<const __nv_bfloat16*>(a & ~uintptr_t(31));
        const int j = (int)((a & 31) >> 2);
        return __bfloat162float(g[j]) + __bfloat162float(g[8 + j]);
    }
    static __device__ __forceinline__ void st(bf16x2_t* p, float v) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(p);
        __nv_bfloat16* g = reinterpret_cast<__nv_bfloat16*>(a & ~uintptr_t(31));
        const int j = (int)((a & 31) >> 2);
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        g[j] = h;
        g[8 + j] = __float2bfloat16_rn(v - __bfloat162float(h));
    }
};

// 256-bit global accesses (sm_100: LDG.E.256 / STG.E.256); p must be 32-byte aligned
__device__ __forceinline__ void ldg256(const void* p, uint32_t (&r)[8]) {
    asm volatile("ld.global.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}
__device__ __forceinline__ void stg256(void* p, const uint32_t (&r)[8]) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(r[0]), "r"(r[1]), "r"(r[2]),
                 "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}

// 8 consecutive channels (one 16-byte bf16 vector / one 32-byte fp32 or split-bf16 vector) <-> 8 floats
template <typename T> __device__ __forceinline__ void load8(const T* p, float (&v)[8]);
template <> __device__ __forceinline__ void load8<float>(const float* p, float (&v)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p);
    float4 b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ float bf16lo_f(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi_f(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
template <> __device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
    uint4 u = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 f = __bfloat1622float2(h[i]);
        v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
}
template <> __device__ __forceinline__ void load8<bf16x2_t>(const bf16x2_t* p, float (&v)[8]) {
    uint32_t u[8];                                   // [4 x hi pairs][4 x lo pairs], one 256-bit load
    ldg256(p, u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        v[2 * i] = bf16lo_f(u[i]) + bf16lo_f(u[4 + i]);
        v[2 * i + 1] = bf16hi_f(u[i]) + bf16hi_f(u[4 + i]);
    }
}
template <> __device__ __forceinline__ void load8<f32in_t>(const f32in_t* p, float (&v)[8]) {
    load8<float>(reinterpret_cast<const float*>(p), v);
}
template <typename T> __device__ __forceinline__ void store8(T* p, const float (&v)[8]);
template <> __device__ __forceinline__ void store8<float>(float* p, const float (&v)[8]) {
    uint32_t u[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) u[i] = __float_as_uint(v[i]);
    stg256(p, u);
}
template <> __device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16* p, const float (&v)[8]) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = u;
}

template <> __device__ __forceinline__ void store8<bf16x2_t>(bf16x2_t* p, const float (&v)[8]) {
    uint32_t u[8];                                   // [4 x hi pairs][4 x lo pairs], one 256-bit store
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        u[i] = *reinterpret_cast<const uint32_t*>(&h);
        const __nv_bfloat162 l = __floats2bfloat162_rn(v[2 * i] - bf16lo_f(u[i]), v[2 * i + 1] - bf16hi_f(u[i]));
        u[4 + i] = *reinterpret_cast<const uint32_t*>(&l);
    }
    stg256(p, u);
}

template <> __device__ __forceinline__ void store8<f32in_t>(f32in_t* p, const float (&v)[8]) {
    store8<bf16x2_t>(static_cast<bf16x2_t*>(p), v);
}

// 4 consecutive channels (c % 4 == 0) of one pixel row `p` (lane 0 of the row, 32-byte aligned)
template <typename T> __device__ __forceinline__ void load4(const T* p, int c, float (&v)[4]);
template <typename T> __device__ __forceinline__ void store4(T* p, int c, const float (&v)[4]);
template <> __device__ __forceinline__ void load4<float>(const float* p, int c, float (&v)[4]) {
    const float4 a = *reinterpret_cast<const float4*>(p + c);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
template <> __device__ __forceinline__ void store4<float>(float* p, int c, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p + c) = make_float4(v[0], v[1], v[2], v[3]);
}
template <> __device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, int c, float (&v)[4]) {
    const uint2 u = *reinterpret_cast<const uint2*>(p + c);
    v[0] = bf16lo_f(u.x); v[1] = bf16hi_f(u.x); v[2] = bf16lo_f(u.y); v[3] = bf16hi_f(u.y);
}
template <> __device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, int c, const float (&v)[4]) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    *reinterpret_cast<uint2*>(p + c) = make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}
// split bf16: group of 8 lanes = [8 x hi][8 x lo]; lanes c..c+3 are 8 bytes in each half
template <> __device__ __forceinline__ void load4<bf16x2_t>(const bf16x2_t* p, int c, float (&v)[4]) {
    const uint2* g = reinterpret_cast<const uint2*>(p + (c & ~7)) + ((c & 7) >> 2);
    const uint2 h = g[0], l = g[2];
    v[0] = bf16lo_f(h.x) + bf16lo_f(l.x); v[1] = bf16hi_f(h.x) + bf16hi_f(l.x);
    v[2] = bf16lo_f(h.y) + bf16lo_f(l.y); v[3] = bf16hi_f(h.y) + bf16hi_f(l.y);
}
template <> __device__ __forceinline__ void store4<bf16x2_t>(bf16x2_t* p, int c, const float (&v)[4]) {
    uint2* g = reinterpret_cast<uint2*>(p + (c & ~7)) + ((c & 7) >> 2);
    const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    const uint32_t ua = *reinterpret_cast<const uint32_t*>(&a), ub = *reinterpret_cast<const uint32_t*>(&b);
    const __nv_bfloat162 la = __floats2bfloat162_rn(v[0] - bf16lo_f(ua), v[1] - bf16hi_f(ua));
    const __nv_bfloat162 lb = __floats2bfloat162_rn(v[2] - bf16lo_f(ub), v[3] - bf16hi_f(ub));
    g[0] = make_uint2(ua, ub);
    g[2] = make_uint2(*reinterpret_cast<const uint32_t*>(&la), *reinterpret_cast<const uint32_t*>(&lb));
}

template <typename T> __device__ __forceinline__ float round_as(float v);
template <> __device__ __forceinline__ float round_as<float>(float v) { return v; }
template <> __device__ __forceinline__ float round_as<__nv_bfloat16>(float v) {
    return __bfloat162float(__float2bfloat16_rn(v));
}

template <> __device__ __forceinline__ float round_as<bf16x2_t>(float v) {
    const float h = __bfloat162float(__float2bfloat16_rn(v));
    return h + __bfloat162float(__float2bfloat16_rn(v - h));
}

template <> __device__ __forceinline__ float round_as<f32in_t>(float v) { return round_as<bf16x2_t>(v); }

// ---- programmatic dependent launch.  Kernels of a forward plan are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization: the next kernel's CTAs may become resident (and run their
// prologue: barrier init, TMEM allocation, weight staging) while the current kernel drains.  Every such kernel calls
// pdl_trigger() first and pdl_wait() before it reads activations or writes anything -- the wait returns when all
// prerequisite grids have completed and their memory is visible, so RAW and WAR hazards on the arena are excluded.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
    static const bool on = [] { const char* e = getenv("EDE_PDL"); return !(e && e[0] == '0'); }();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = on ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// the same with a thread-block cluster of `cluster_x` CTAs along x (grid.x must be a multiple of it)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                      unsigned cluster_x, Args&&... args) {
    static const bool on = [] { const char* e = getenv("EDE_PDL"); return !(e && e[0] == '0'); }();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cluster_x;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = on ? 2 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// x * sigmoid(x) with the two SFU approximations (ftz: a flushed exp only matters where the result is ~0 anyway)
__device__ __forceinline__ float swish_fast(float v) {
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(v * -1.4426950408889634f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f + e));
    return v * r;
}

__device__ __forceinline__ float apply_act(float v, int act) {
    switch (act) {
        case EDE_ACT_RELU: return fmaxf(v, 0.f);
        case EDE_ACT_SWISH: return __fdividef(v, 1.f + __expf(-v));
        case EDE_ACT_SIGMOID: return __fdividef(1.f, 1.f + __expf(-v));
        default: return v;
    }
}

// Packed fp32 FMA (Blackwell FFMA2): two independent round-to-nearest fp32 FMAs per issued instruction -- the same
// values as two fmaf(), half the issue slots.  Operands are register pairs (lo, hi).
__device__ __forceinline__ unsigned long long pack_f32x2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack_f32x2(unsigned long long v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void ffma2(unsigned long long& acc, unsigned long long a, unsigned long long b) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}
// acc = a * b + acc on both lanes of a float2 (bit-identical to two fmaf)
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 acc) {
    unsigned long long r = pack_f32x2(acc.x, acc.y);
    ffma2(r, pack_f32x2(a.x, a.y), pack_f32x2(b.x, b.y));
    float2 o;
    unpack_f32x2(r, o.x, o.y);
    return o;
}

// floor and fraction of 0 <= s < 2^22 without the quarter-rate F2I / I2F conversions: adding 2^23 with round-toward-zero
// leaves floor(s) in the low mantissa bits, and (t - 2^23) is that integer as a float, exactly.  Same values as
// `i0 = (int)s; frac = s - i0`, which is how ATen's upsample_bilinear2d computes them.
__device__ __forceinline__ void floor_split(float s, int& i0, float& frac) {
    const float t = __fadd_rz(s, 8388608.f);
    i0 = __float_as_int(t) & 0x7fffff;
    frac = __fsub_rn(s, __fsub_rn(t, 8388608.f));
}

// One axis of ATen's upsample_bilinear2d (align_corners=True): source index pair and weights of output index `o`.
struct LerpTap { int i0, i1; float w0, w1; };
__device__ __forceinline__ LerpTap lerp_tap_ac(float scale, float o, int in_size) {
    LerpTap t;
    floor_split(__fmul_rn(scale, o), t.i0, t.w1);
    t.i1 = t.i0 + (t.i0 < in_size - 1 ? 1 : 0);
    t.w0 = __fsub_rn(1.f, t.w1);
    return t;
}
// explicit roundings (no compiler-chosen FMA contraction): the value does not depend on the surrounding kernel
__device__ __forceinline__ float bilerp_taps(const float* __restrict__ p, int w, const LerpTap& y, const LerpTap& x) {
    const float top = __fmaf_rn(x.w0, p[y.i0 * w + x.i0], __fmul_rn(x.w1, p[y.i0 * w + x.i1]));
    const float bot = __fmaf_rn(x.w0, p[y.i1 * w + x.i0], __fmul_rn(x.w1, p[y.i1 * w + x.i1]));
    return __fmaf_rn(y.w0, top, __fmul_rn(y.w1, bot));
}
// One sample of ATen's upsample_bilinear2d (align_corners=True) on an fp32 plane [h][w] (h*w < 2^31): src = scale *
// dst, the lambdas in fp32, the same expression tree in every kernel that resizes a depth map (bit-identical results).
__device__ __forceinline__ float bilerp_ac_plane(const float* __restrict__ p, int h, int w, float scale_h, float scale_w,
                                                 int oy, int ox) {
    return bilerp_taps(p, w, lerp_tap_ac(scale_h, (float)oy, h), lerp_tap_ac(scale_w, (float)ox, w));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// dispatch helper: call f.template operator()<T>() for the dtype
// depthwise kernels: the three storage types plus fp32-in / split-out
#define EDE_DISPATCH_DTYPE_DW(dt, T, ...)                    \
    do {                                                     \
        if ((dt) == EDE_BF16) { using T = __nv_bfloat16; __VA_ARGS__; } \
        else if ((dt) == EDE_BF16X2) { using T = ::ede::bf16x2_t; __VA_ARGS__; } \
        else if ((dt) == EDE_F32_TO_BF16X2) { using T = ::ede::f32in_t; __VA_ARGS__; } \
        else { using T = float; __VA_ARGS__; }              \
    } while (0)

#define EDE_DISPATCH_DTYPE(dt, T, ...)                       \
    do {                                                     \
        if ((dt) == EDE_BF16) { using T = __nv_bfloat16; __VA_ARGS__; } \
        else if ((dt) == EDE_BF16X2) { using T = ::ede::bf16x2_t; __VA_ARGS__; } \
        else { using T = float; __VA_ARGS__; }              \
    } while (0)

}  // namespace ede
