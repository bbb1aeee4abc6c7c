// Microbenchmark: cp.async.bulk global->shared throughput per SM as a function of copy size and copies in flight.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_microbench bulk_microbench.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// each CTA streams `iters` batches; a batch = ncopy copies of `bytes`, two batches in flight (double buffer)
__global__ void __launch_bounds__(64) k(const unsigned char* src, long long src_bytes, int bytes, int ncopy, int iters, long long stride,
                                        long long* out) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t bar[2];
    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[i])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const long long batch = (long long)bytes * ncopy;
        long long t0 = clock64();
        for (int it = 0; it < iters + 1; ++it) {
            if (it < iters) {
                const int st = it & 1;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[st])), "r"((uint32_t)batch) : "memory");
                long long off = (((long long)blockIdx.x * iters + it) * stride) % (src_bytes - batch - stride * 8);
                for (int c = 0; c < ncopy; ++c)
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(smem_u32(smem) + (uint32_t)(st * batch + (long long)c * bytes)), "l"(src + off + (long long)c * stride * 3), "r"((uint32_t)bytes),
                                   "r"(smem_u32(&bar[st])) : "memory");
            }
            if (it >= 1) { const int st = (it - 1) & 1; while (!mbar_try(smem_u32(&bar[st]), ((it - 1) >> 1) & 1)) {} }
        }
        long long t1 = clock64();
        if (blockIdx.x == 0) out[0] = t1 - t0;
    }
}
int main() {
    const long long src_bytes = 2LL << 30;
    unsigned char* src; cudaMalloc(&src, src_bytes); cudaMemset(src, 1, src_bytes);
    long long* d; cudaMalloc(&d, 64);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    printf("%-8s %-6s %-10s %-12s %-10s\n", "bytes", "ncopy", "clk/batch", "B/clk/SM", "chip GB/s @1.9GHz");
    const int cfg[][2] = {{57344, 1}, {14336, 4}, {7168, 8}, {3584, 16}, {2048, 28}, {1024, 56}, {28672, 3}, {14336, 6}, {4096, 24}};
    for (auto& c : cfg) {
        const int iters = 200;
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        k<<<148, 64, 2 * c[0] * c[1] + 1024>>>(src, src_bytes, c[0], c[1], iters, 57344 * 7 + 4096, d);
        cudaEventRecord(e0);
        k<<<148, 64, 2 * c[0] * c[1] + 1024>>>(src, src_bytes, c[0], c[1], iters, 57344 * 7 + 4096, d);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        double clk = (double)h / iters, bpc = (double)c[0] * c[1] / clk;
        printf("%-8d %-6d %-10.0f %-12.1f %-10.0f (event: %.0f GB/s)\n", c[0], c[1], clk, bpc, bpc * 148 * 1.9, 148.0 * iters * c[0] * c[1] / ms / 1e6);
    }
    return 0;
}
