// Microbenchmark: what slows tcgen05.mma (M=128, N=256, K=16, 128B swizzle, SS operands) down inside a real conv CTA?
// One issuing thread runs `iters` steps of `ks` MMAs; optional concurrent activity in the same CTA:
//   commit  : tcgen05.commit to an mbarrier after every step (as the conv kernel does to free its stages)
//   shift_b : N-side operand start shifted by whole 128-byte rows (the resident-halo-tile windows of the swapped mode)
//   tma     : a second thread streams cp.async.bulk global->smem copies of `tma_box` bytes, `tma_depth` in flight
//   epi     : four warps loop tcgen05.ld (32x32b.x16) on the other half of TMEM and st.shared what they read
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_contend mma_contend.cu && ./mma_contend
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(bar),
        "r"(parity)
        : "memory");
}

struct Args {
    int n, ks, iters, commit, shift_a, shift_b, tma_box, tma_depth, epi, issue, wait;   // issue: 0 thread 0 alone, 1 whole warp + elected lane
    const uint8_t* src;
    long long* out;      // per CTA: [0] MMA cycles, [1] bytes copied, [2] tcgen05.ld count
};

__global__ void __launch_bounds__(256) k(const Args a) {
    extern __shared__ uint8_t raw[];
    __shared__ uint64_t bars[16];
    __shared__ uint32_t slot;
    __shared__ volatile int done;
    __shared__ float stage[4][32][17];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 16; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[i])));
        asm volatile("fence.mbarrier_init.release.cluster;");
        done = 0;
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = slot;
    if (warp == 0 && a.issue == 1) {
        // the whole warp walks the loop (uniform control flow, uniform registers); one elected lane issues
        uint64_t dc = ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
        const uint32_t hi = (uint32_t)(dc >> 32);
        const uint32_t a_lo = (uint32_t)dc + ((base + a.shift_a * 128) >> 4);
        const uint32_t b_lo = (uint32_t)dc + ((base + 32 * 1024 + a.shift_b * 128) >> 4);
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(a.n >> 3) << 17) | ((128u >> 4) << 24);
        uint32_t leader;
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(leader));
        long long t0 = clock64();
        for (int i = 0; i < a.iters; ++i) {
            if (a.wait && i >= 4) {          // the stage this step reuses was committed four steps ago
                mbar_wait(smem_u32(&bars[1 + (i & 3)]), ((i >> 2) - 1) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            for (int kk = 0; kk < a.ks; ++kk)
                if (leader)
                    asm volatile(
                        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                        "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
                        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}" ::"r"(tmem),
                        "r"(a_lo + 2 * kk), "r"(b_lo + 2 * kk), "r"(hi), "r"(idesc), "r"(1));
            if (a.commit != 0 && leader)
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                    smem_u32(&bars[1 + (i & 3)])));
            __syncwarp();
        }
        if (leader)
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bars[0])));
        mbar_wait(smem_u32(&bars[0]), 0);
        long long t1 = clock64();
        if (lane == 0) { a.out[blockIdx.x * 4] = t1 - t0; done = 1; }
    } else if (threadIdx.x == 0 && a.issue == 0) {
        uint64_t dc = ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
        const uint32_t hi = (uint32_t)(dc >> 32);
        const uint32_t a_lo = (uint32_t)dc + ((base + a.shift_a * 128) >> 4);               // 128 rows: 16 KB
        const uint32_t b_lo = (uint32_t)dc + ((base + 32 * 1024 + a.shift_b * 128) >> 4);   // 256 rows: 32 KB (+ shift)
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(a.n >> 3) << 17) | ((128u >> 4) << 24);
        long long t0 = clock64();
        for (int i = 0; i < a.iters; ++i) {
            if (a.wait && i >= 4) {
                mbar_wait(smem_u32(&bars[1 + (i & 3)]), ((i >> 2) - 1) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            for (int kk = 0; kk < a.ks; ++kk)
                asm volatile(
                    "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                    "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
                    "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}" ::"r"(tmem),
                    "r"(a_lo + 2 * kk), "r"(b_lo + 2 * kk), "r"(hi), "r"(idesc), "r"(1));
            if (a.commit != 0 && (a.commit < 0 || i % a.commit == 0))
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                    smem_u32(&bars[1 + (i & 3)])));
            if (a.commit < 0)
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                    smem_u32(&bars[5])));
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bars[0])));
        mbar_wait(smem_u32(&bars[0]), 0);
        long long t1 = clock64();
        a.out[blockIdx.x * 4] = t1 - t0;
        done = 1;
    } else if (warp == 1 && lane == 0) {
        long long bytes = 0;
        if (a.tma_box > 0) {
            const uint32_t ring = base + 80 * 1024;                 // up to 8 slots of 16 KB
            int issued = 0, waited = 0;
            uint32_t ph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            while (!done) {
                if (issued - waited == a.tma_depth) {
                    const int s = waited % a.tma_depth;
                    mbar_wait(smem_u32(&bars[8 + s]), ph[s]);
                    ph[s] ^= 1;
                    ++waited;
                    bytes += a.tma_box;
                }
                const int s = issued % a.tma_depth;
                const uint32_t bar = smem_u32(&bars[8 + s]);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(a.tma_box) : "memory");
                asm volatile(
                    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                        ring + s * 16384),
                    "l"(a.src + ((size_t)(issued * 37 + blockIdx.x * 11) % 4096) * 16384), "r"(a.tma_box), "r"(bar)
                    : "memory");
                ++issued;
            }
            while (waited < issued) {
                const int s = waited % a.tma_depth;
                mbar_wait(smem_u32(&bars[8 + s]), ph[s]);
                ph[s] ^= 1;
                ++waited;
            }
        }
        a.out[blockIdx.x * 4 + 1] = bytes;
    } else if (warp >= 4 && a.epi) {
        long long n_ld = 0;
        const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + 256;
        while (!done) {
            for (int c = 0; c < 256; c += 16) {
                uint32_t r[16];
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                      "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                    : "r"(taddr + c));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (a.epi > 1)
                    for (int j = 0; j < 16; ++j) stage[warp & 3][lane][j] = __uint_as_float(r[j]);
                ++n_ld;
            }
        }
        if (lane == 0 && warp == 4) a.out[blockIdx.x * 4 + 2] = n_ld;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

static void run(const char* name, Args a) {
    cudaMemset(a.out, 0, 148 * 4 * sizeof(long long));
    k<<<148, 256, 216 * 1024>>>(a);
    long long h[148 * 4];
    if (cudaMemcpy(h, a.out, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) {
        printf("%s: CUDA error %s\n", name, cudaGetErrorString(cudaGetLastError()));
        return;
    }
    double c = 0, b = 0, l = 0;
    for (int i = 0; i < 148; ++i) { c += (double)h[4 * i]; b += (double)h[4 * i + 1]; l += (double)h[4 * i + 2]; }
    c /= 148; b /= 148; l /= 148;
    printf("%-44s N=%3d ks=%d  %7.1f clk/MMA   copy %5.1f B/clk   tcgen05.ld %6.1f B/clk\n", name, a.n, a.ks,
           c / ((double)a.iters * a.ks), b / c, l * 4 * 32 * 16 * 4 / c);
}

int main() {
    long long* d;
    uint8_t* src;
    cudaMalloc(&d, 148 * 4 * sizeof(long long));
    cudaMalloc(&src, (size_t)4096 * 16384 + 16384);
    cudaMemset(src, 0, (size_t)4096 * 16384 + 16384);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024);
    Args z = {256, 4, 2000, 0, 0, 0, 0, 4, 0, 0, 0, src, d};
    for (int n : {256}) {
        Args a = z; a.n = n;
        run("plain", a);
        a.commit = 1; run("commit per step", a);
        for (int ks : {3, 2, 1}) { a.ks = ks; run("commit per step", a); }
        a.commit = 2; run("commit every 2nd step", a);
        a.commit = 4; run("commit every 4th step", a);
        a.commit = 0; run("no commit", a);
        a.commit = -1; run("two commits per step", a);
        a.issue = 1;
        for (int ks : {4, 2, 1}) { a.ks = ks; a.commit = 1; run("warp-uniform loop, commit per step", a); }
        a.commit = 0; run("warp-uniform loop, no commit", a);
        a.commit = 1; a.wait = 1;
        for (int ks : {4, 2, 1}) { a.ks = ks; run("warp-uniform, stage wait + fence + commit", a); }
        a.issue = 0;
        for (int ks : {4, 2, 1}) { a.ks = ks; run("thread 0 alone, stage wait + fence + commit", a); }
        a = z; a.n = n; a.shift_b = 3; run("N operand shifted 3 rows", a);
        a.shift_b = 42; run("N operand shifted 42 rows", a);
        a = z; a.n = n; a.commit = 1;
        for (int box : {4096, 10240, 16384})
            for (int depth : {2, 4, 8}) {
                a.tma_box = box; a.tma_depth = depth;
                char nm[64];
                snprintf(nm, sizeof nm, "bulk copies %5d B x %d in flight", box, depth);
                run(nm, a);
            }
        a = z; a.n = n; a.commit = 1; a.epi = 1; run("tcgen05.ld loop", a);
        a.epi = 2; run("tcgen05.ld + st.shared loop", a);
        a.tma_box = 16384; a.tma_depth = 4; run("all: copies 16K x4 + ld + st.shared", a);
        a.shift_b = 42; run("all + shifted N operand", a);
    }
    return 0;
}
