// Microbenchmark: tcgen05.mma (kind::f16, cta_group::1, SS) issue/throughput as a function of N and of the
// alignment of the no-swizzle K-major A window, plus tcgen05.ld drain rate.  Design input for cd_conv_mma.cu.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_microbench mma_microbench.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t ad, uint64_t bd, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                 ::"r"(tmem_d), "l"(ad), "l"(bd), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}

// mode 0: MMA throughput. A window starts at row `shift`, rows 16 B apart (8-channel chunk), K chunks lbo apart.
__global__ void __launch_bounds__(192) bench_kernel(int N, int shift, int nmma, int ndistinct, int a_stride_rows,
                                                    long long* out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;  // fp16 1.0
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    // A region: bytes [0, 96 KB): 2 K-chunks of 3000 rows each (lbo = 48000 B).  B region at 96 KB: [kc][n][8]
    const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem) + 96 * 1024;
    if (warp == 1 && lane == 0) {
        long long t0 = clock64();
        uint64_t ads[8];
#pragma unroll
        for (int v = 0; v < 8; ++v) ads[v] = make_desc(a0 + (uint32_t)(shift + (v % ndistinct) * a_stride_rows) * 16u, 48000u, 128u);
        const uint64_t bd = make_desc(b0, (uint32_t)N * 16u, 128u);
        for (int i = 0; i < nmma; i += 8) {
#pragma unroll
            for (int v = 0; v < 8; ++v) umma(tmem_base + (uint32_t)((v & 1) * 256), ads[v], bd, idesc, 1);
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        while (!mbar_try(smem_u32(&bar), 0)) {}
        long long t1 = clock64();
        if (blockIdx.x == 0) out[0] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // mode: TMEM drain rate with 4 warps (warps 2..5 -> lane quarters)
    if (warp >= 2) {
        const int q = warp & 3;
        uint32_t acc = 0;
        __syncwarp();
        long long t0 = clock64();
        for (int it = 0; it < 64; ++it) {
            for (int c = 0; c < 512; c += 32) {
                uint32_t r[32];
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                      "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                      "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                      "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                    : "r"(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 32; ++i) acc ^= r[i];
            }
        }
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) { out[1 + q] = t1 - t0; out[8 + q] = acc; }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

int main() {
    long long* d;
    cudaMalloc(&d, 16 * sizeof(long long));
    const int smem = 160 * 1024;
    cudaFuncSetAttribute(bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int nmma = 4096;
    printf("%-6s %-6s %-10s %-10s %s\n", "N", "shift", "distinct", "clk/mma", "(grid 148; A K-major no-swizzle, 128x16 per MMA)");
    for (int N : {32, 64, 96, 128, 192, 256}) {
        for (int shift : {0, 1, 4}) {
            for (int nd : {1, 8}) {
                cudaMemset(d, 0, 16 * sizeof(long long));
                bench_kernel<<<148, 192, smem>>>(N, shift, nmma, nd, 11, d);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
                long long h[16];
                cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
                printf("%-6d %-6d %-10d %-10.1f tmem-drain clk per 128x512x4B (4 warps): %lld\n", N, shift, nd, (double)h[0] / nmma, h[1] / 64);
            }
        }
    }
    return 0;
}
