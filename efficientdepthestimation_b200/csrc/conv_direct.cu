// conv_direct.cu -- shape-generic SIMT implicit-GEMM convolution (fp32 accumulate).
//
// The exact-arithmetic path: used for the fp32 parity mode, for the 3-channel NCHW stem
// in that mode, and for shapes the tcgen05 path (conv_umma.cu) declines.  Any kernel size,
// stride, asymmetric zero padding; NHWC (or NCHW fp32 for a network input) in, NHWC out.
// GEMM view: M = n*oh*ow pixels, N = cout_p, K = kh*kw*cin_p (cin innermost).
#include "common.cuh"

namespace ede {

constexpr int DBM = 64, DBN = 64, DBK = 16;

template <typename TX, typename TY>
__global__ void __launch_bounds__(256) conv_direct_kernel(const EdeConv a) {
    __shared__ float As[DBK][DBM + 4];
    __shared__ float Bs[DBK][DBN + 4];

    const int t = threadIdx.x;
    const int tx = t & 15, ty = t >> 4;
    const long long m_total = (long long)a.n * a.oh * a.ow;
    const long long m0 = (long long)blockIdx.x * DBM;
    const int n0 = blockIdx.y * DBN;
    const int K = a.kh * a.kw * a.cin_p;
    const TX* __restrict__ x = reinterpret_cast<const TX*>(a.x);

    // A-load assignment: k_local = t % 16, pixels m = t/16 + 16*j
    const int ka = t & 15;
    int pn[4], pih[4], piw[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        long long m = m0 + (t >> 4) + 16 * j;
        if (m < m_total) {
            int ow = (int)(m % a.ow);
            long long r = m / a.ow;
            int oh = (int)(r % a.oh);
            pn[j] = (int)(r / a.oh);
            pih[j] = oh * a.sh - a.pt;
            piw[j] = ow * a.sw - a.pl;
        } else {
            pn[j] = -1; pih[j] = 0; piw[j] = 0;
        }
    }
    // B-load assignment: 4 consecutive n for one k
    const int kb = t >> 4, nb = (t & 15) * 4;

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < K; k0 += DBK) {
        {
            const int k = k0 + ka;
            int tap = 0, c = 0, ky = 0, kx = 0;
            const bool kvalid = k < K;
            if (kvalid) { tap = k / a.cin_p; c = k - tap * a.cin_p; ky = tap / a.kw; kx = tap - ky * a.kw; }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float v = 0.f;
                const int ih = pih[j] + ky, iw = piw[j] + kx;
                if (kvalid && pn[j] >= 0 && ih >= 0 && ih < a.h && iw >= 0 && iw < a.w) {
                    if (a.x_layout == EDE_NCHW)
                        v = Elt<TX>::ld(x + (((long long)pn[j] * a.cin_p + c) * a.h + ih) * a.w + iw);
                    else
                        v = Elt<TX>::ld(x + (((long long)pn[j] * a.h + ih) * a.w + iw) * a.x_cs + a.x_co + c);
                    if (a.relu_in) v = fmaxf(v, 0.f);
                }
                As[ka][(t >> 4) + 16 * j] = v;
            }
            const int kk = k0 + kb;
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (kk < K && n0 + nb < a.cout_p)
                bv = *reinterpret_cast<const float4*>(a.w_direct + (long long)kk * a.cout_p + n0 + nb);
            Bs[kb][nb] = bv.x; Bs[kb][nb + 1] = bv.y; Bs[kb][nb + 2] = bv.z; Bs[kb][nb + 3] = bv.w;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < DBK; ++k) {
            float av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = As[k][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) bv[j] = Bs[k][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }

    TY* __restrict__ y = reinterpret_cast<TY*>(a.y);
    const TY* __restrict__ res = reinterpret_cast<const TY*>(a.res);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const long long m = m0 + ty * 4 + i;
        if (m >= m_total) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = n0 + tx * 4 + j;
            if (c >= a.cout_p) continue;
            float v = fmaf(acc[i][j], a.scale[c], a.shift[c]);
            if (res) v += Elt<TY>::ld(res + m * a.res_cs + a.res_co + c);
            if (c < a.act_c_end) v = apply_act(v, a.act);
            Elt<TY>::st(y + m * a.y_cs + a.y_co + c, v);
        }
    }
}

int conv_direct_launch(const EdeConv& a, cudaStream_t st) {
    EDE_REQUIRE(a.w_direct != nullptr, "ede_conv2d: direct path needs w_direct");
    const long long m_total = (long long)a.n * a.oh * a.ow;
    dim3 grid((unsigned)((m_total + DBM - 1) / DBM), (unsigned)ceil_div(a.cout_p, DBN));
    EDE_DISPATCH_DTYPE(a.x_dtype, TX,
                       EDE_DISPATCH_DTYPE(a.y_dtype, TY, (conv_direct_kernel<TX, TY><<<grid, 256, 0, st>>>(a))));
    EDE_LAUNCH_CHECK("conv_direct_kernel");
    return EDE_OK;
}

}  // namespace ede
