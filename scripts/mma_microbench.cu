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
                                                    long long* out, int nacc_rot, int accw, int lbo) {
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
        for (int v = 0; v < 8; ++v) ads[v] = make_desc(a0 + (uint32_t)(shift + (v % ndistinct) * a_stride_rows) * 16u, (uint32_t)lbo, 128u);
        const uint64_t bd = make_desc(b0, (uint32_t)N * 16u, 128u);
        for (int i = 0; i < nmma; i += 8) {
#pragma unroll
            for (int v = 0; v < 8; ++v) umma(tmem_base + (uint32_t)((v % nacc_rot) * accw), ads[v], bd, idesc, 1);
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

// mode B: the conv kernel's descriptor walk: 9 shifts x nq K-steps per 'sub-tile', A offsets spread over a 57 KB window,
// B walking through resident weights (9*nq*N*32 bytes)
__global__ void __launch_bounds__(192) walk_kernel(int N, int nsub, int vary_a, int vary_b, int nq, long long* out, int warpwide) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
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
    const int wrows = 886, plane = 180, Rp = 10;
    const uint32_t chunk_bytes = wrows * 16;
    const uint32_t wbuf = smem_u32(smem), abuf = wbuf + 9 * nq * N * 32;
    if (warp == 1 && (warpwide || lane == 0)) {
        uint32_t leader = 1;
        if (warpwide) asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xffffffff;\nselp.u32 %0, 1, 0, P;\n}" : "=r"(leader));
        uint32_t soff[9];
        for (int s = 0; s < 9; ++s) soff[s] = vary_a ? (uint32_t)(((plane + Rp) + (s / 3 - 1) * plane + (s % 3 - 1) * Rp) * 16) >> 4 : 0u;
        const uint64_t ahi = make_desc(0, chunk_bytes, 128u), bhi = make_desc(0, (uint32_t)N * 16u, 128u);
        const uint32_t qstep = (2u * chunk_bytes) >> 4, bstep = (vary_b & 1) ? (2u * N * 16u) >> 4 : 0u;
        long long t0 = clock64();
        for (int n = 0; n < nsub; ++n) {
            const uint32_t a_lo = (abuf >> 4) + (uint32_t)(vary_a ? (n & 3) * 126 : 0);
            uint32_t b_lo = wbuf >> 4;
            const uint32_t dcol = tmem_base + (uint32_t)(((vary_b & 4) ? 0 : (n % 5)) * N);
#pragma unroll
            for (int s = 0; s < 9; ++s) {
                uint32_t al = a_lo + soff[s];
                for (int q = 0; q < nq; ++q, al += qstep, b_lo += bstep) { if (leader) umma(dcol, ahi | al, bhi | b_lo, idesc, (vary_b & 2) ? 1u : (uint32_t)((s | q) != 0)); }
            }
        }
        if (leader) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        while (!mbar_try(smem_u32(&bar), 0)) {}
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) out[0] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

// mode B: the conv kernel's descriptor walk: 9 shifts x nq K-steps per 'sub-tile', A offsets spread over a 57 KB window,
// B walking through resident weights (9*nq*N*32 bytes)
// mode B: the conv kernel's descriptor walk: 9 shifts x nq K-steps per 'sub-tile', A offsets spread over a 57 KB window,
// B walking through resident weights (9*nq*N*32 bytes)
__global__ void __launch_bounds__(192) walk3_kernel(int N, int nsub, int vary_a, int vary_b, int nq, long long* out, int warpwide) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar[4];
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[i])));
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
    const int wrows = 886, plane = 180, Rp = 10;
    const uint32_t chunk_bytes = wrows * 16;
    const uint32_t wbuf = smem_u32(smem), abuf = wbuf + 9 * nq * N * 32;
    const int nw = warpwide;  // number of issuing warps (1..4): warps 1..nw
    if (warp >= 1 && warp <= nw && lane == 0) {
        const int me = warp - 1;
        uint32_t leader = 1;

        uint32_t soff[9];
        for (int s = 0; s < 9; ++s) soff[s] = vary_a ? (uint32_t)(((plane + Rp) + (s / 3 - 1) * plane + (s % 3 - 1) * Rp) * 16) >> 4 : 0u;
        const uint64_t ahi = make_desc(0, chunk_bytes, 128u), bhi = make_desc(0, (uint32_t)N * 16u, 128u);
        const uint32_t qstep = (2u * chunk_bytes) >> 4, bstep = (vary_b & 1) ? (2u * N * 16u) >> 4 : 0u;
        long long t0 = clock64();
        for (int n = me; n < nsub; n += nw) {
            const uint32_t a_lo = (abuf >> 4) + (uint32_t)(vary_a ? (n & 3) * 126 : 0);
            uint32_t b_lo = wbuf >> 4;
            const uint32_t dcol = tmem_base + (uint32_t)(((vary_b & 4) ? 0 : (n % 5)) * N);
#pragma unroll
            for (int s = 0; s < 9; ++s) {
                uint32_t al = a_lo + soff[s];
                for (int q = 0; q < nq; ++q, al += qstep, b_lo += bstep) { if (leader) umma(dcol, ahi | al, bhi | b_lo, idesc, (vary_b & 2) ? 1u : (uint32_t)((s | q) != 0)); }
            }
        }
        if (leader) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar[me])) : "memory");
        while (!mbar_try(smem_u32(&bar[me]), 0)) {}
        long long t1 = clock64();
        if (blockIdx.x == 0) out[me] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

// mode B: the conv kernel's descriptor walk: 9 shifts x nq K-steps per 'sub-tile', A offsets spread over a 57 KB window,
// B walking through resident weights (9*nq*N*32 bytes)
template <int NQ>
__global__ void __launch_bounds__(192) walk2_kernel(int N, int nsub, int vary_a, int vary_b, int nq, long long* out, int warpwide) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
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
    const int wrows = 886, plane = 180, Rp = 10;
    const uint32_t chunk_bytes = wrows * 16;
    const uint32_t wbuf = smem_u32(smem), abuf = wbuf + 9 * nq * N * 32;
    if (warp == 1 && (warpwide || lane == 0)) {
        uint32_t leader = 1;
        if (warpwide) asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xffffffff;\nselp.u32 %0, 1, 0, P;\n}" : "=r"(leader));
        uint32_t soff[9];
        for (int s = 0; s < 9; ++s) soff[s] = vary_a ? (uint32_t)(((plane + Rp) + (s / 3 - 1) * plane + (s % 3 - 1) * Rp) * 16) >> 4 : 0u;
        const uint64_t ahi = make_desc(0, chunk_bytes, 128u), bhi = make_desc(0, (uint32_t)N * 16u, 128u);
        const uint32_t qstep = (2u * chunk_bytes) >> 4, bstep = (vary_b & 1) ? (2u * N * 16u) >> 4 : 0u;
        long long t0 = clock64();
        for (int n = 0; n < nsub; ++n) {
            const uint32_t a_lo = (abuf >> 4) + (uint32_t)(vary_a ? (n & 3) * 126 : 0);
            uint32_t b_lo = wbuf >> 4;
            const uint32_t dcol = tmem_base + (uint32_t)(((vary_b & 4) ? 0 : (n % 5)) * N);
#pragma unroll
            for (int s = 0; s < 9; ++s) {
                uint32_t al = a_lo + soff[s];
                #pragma unroll
                for (int q = 0; q < NQ; ++q, al += qstep, b_lo += bstep) { if (leader) umma(dcol, ahi | al, bhi | b_lo, idesc, (vary_b & 2) ? 1u : (uint32_t)((s | q) != 0)); }
            }
        }
        if (leader) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        while (!mbar_try(smem_u32(&bar), 0)) {}
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) out[0] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

int main() {
    {
        long long* d;
        cudaMalloc(&d, 16 * sizeof(long long));
        cudaFuncSetAttribute(walk3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        for (int nw = 1; nw <= 4; ++nw) {
            walk3_kernel<<<148, 192, 200 * 1024>>>(96, 400, 1, 1, 2, d, nw);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
            long long h[4]; cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
            long long mx = 0; for (int i = 0; i < nw; ++i) mx = h[i] > mx ? h[i] : mx;
            printf("walk3 N=96 issuing warps=%d: aggregate clk/mma %.1f\n", nw, (double)mx / (400 * 18));
        }
        cudaFuncSetAttribute(walk2_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        for (int ww = 0; ww < 2; ++ww) {
            walk2_kernel<2><<<148, 192, 200 * 1024>>>(96, 400, 1, 1, 2, d, ww);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
            long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
            printf("walk2 (fully unrolled) N=96 warpwide=%d: clk/mma %.1f\n", ww, (double)h / (400 * 18));
        }
        cudaFuncSetAttribute(walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        for (int va = 1; va < 2; ++va)
            for (int vb = 0; vb < 8; ++vb) {
                for (int ww = 1; ww < 2; ++ww) {
                walk_kernel<<<148, 192, 200 * 1024>>>(96, 400, va, vb, 2, d, ww);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
                long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
                printf("walk N=96 vary_a=%d vary_b=%d warpwide=%d: clk/mma %.1f\n", va, vb, ww, (double)h / (400 * 18));
                }
            }
    }
    long long* d;
    cudaMalloc(&d, 16 * sizeof(long long));
    const int smem = 160 * 1024;
    cudaFuncSetAttribute(bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int nmma = 4096;
    printf("%-6s %-6s %-10s %-10s %s\n", "N", "shift", "distinct", "clk/mma", "(grid 148; A K-major no-swizzle, 128x16 per MMA)");
    for (int N : {96, 192}) {
        for (int lbo : {48000, 14176, 14208, 14336, 2048, 4448, 8800}) {
            for (int astr : {11, 61}) {
            cudaMemset(d, 0, 16 * sizeof(long long));
            bench_kernel<<<148, 192, smem>>>(N, 1, nmma, 8, astr, d, 1, N, lbo);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
            long long h[16];
            cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
            printf("N=%-4d lbo=%-6d (mod128=%3d) a_stride_rows=%d clk/mma %-8.1f\n", N, lbo, lbo % 128, astr, (double)h[0] / nmma);
            }
        }
    }
    return 0;
}
