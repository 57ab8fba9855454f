// Microbenchmark: cycles per tcgen05.mma (kind::f16, cta_group::1, M=128, K=16, SS operands) as a function of
// N, of the A start-address alignment inside the 128B-swizzle atom (the "resident halo tile" trick of
// conv_umma.cu shifts A by whole pixel rows) and of the swizzle width.  No loads: operands are whatever is in smem.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu && ./mma_rate
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128) k(int n, int row_shift, int swz_bytes, int iters, int n_acc, long long* out) {
    extern __shared__ uint8_t raw[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = slot;
    if (threadIdx.x == 0) {
        const int layout = swz_bytes == 128 ? 2 : swz_bytes == 64 ? 4 : 6;
        uint64_t dc = ((uint64_t)1 << 16) | ((uint64_t)((8 * swz_bytes) >> 4) << 32) | ((uint64_t)1 << 46) |
                      ((uint64_t)layout << 61);
        const uint32_t hi = (uint32_t)(dc >> 32);
        const uint32_t a_lo = (uint32_t)dc + ((base + row_shift * swz_bytes) >> 4);
        const uint32_t b_lo = (uint32_t)dc + ((base + 96 * 1024) >> 4);
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
        const int ksteps = swz_bytes / 32;
        long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            for (int m = 0; m < n_acc; ++m)
                for (int kk = 0; kk < ksteps; ++kk) {
                    asm volatile(
                        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\t"
                        "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
                        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}" ::"r"(tmem + m * n),
                        "r"(a_lo + 2 * kk + m * ((128 * swz_bytes) >> 4)), "r"(b_lo + 2 * kk), "r"(hi), "r"(idesc), "r"(1));
                }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)));
        asm volatile(
            "{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(
                smem_u32(&bar)));
        long long t1 = clock64();
        out[blockIdx.x] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

int main() {
    long long* d;
    cudaMalloc(&d, 148 * sizeof(long long));
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 2000;
    printf("swz  N  n_acc shift  cycles/MMA  (grid 148 CTAs, %d iterations)\n", iters);
    int ns[] = {16, 32, 64, 80, 128, 256};
    for (int swz : {128, 64, 32})
        for (int n : ns)
            for (int n_acc : {1, 2})
                for (int shift : {0, 1, 3, 8}) {
                    if (n * n_acc > 512) continue;
                    k<<<148, 128, 200 * 1024>>>(n, shift, swz, iters, n_acc, d);
                    long long h[148];
                    if (cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) {
                        printf("CUDA error %s\n", cudaGetErrorString(cudaGetLastError()));
                        return 1;
                    }
                    double s = 0;
                    for (int i = 0; i < 148; ++i) s += (double)h[i];
                    printf("%4d %4d %3d %4d   %8.1f\n", swz, n, n_acc, shift, s / 148 / (double)(iters * n_acc * (swz / 32)));
                }
    return 0;
}
