// Microbenchmark: semantics and cost of tcgen05.shift.down (TMEM row shift), a candidate for doing the R-recombination of
// cd_conv_mma.cu inside tensor memory instead of with warp shuffles.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o shift_microbench shift_microbench.cu
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

// out[lane][col] for 16 columns; nshift shifts applied to columns [col0, col0+8)
__global__ void __launch_bounds__(128) shift_kernel(uint32_t* out, int nshift, int col0, long long* clk, int ntime) {
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16);
    uint32_t v[16];
    for (int c = 0; c < 16; ++c) v[c] = threadIdx.x * 100 + c;
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
                   "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (threadIdx.x == 0) {
        long long t0 = clock64();
        for (int i = 0; i < nshift; ++i)
            asm volatile("tcgen05.shift.cta_group::1.down [%0];" ::"r"(tmem_base + (uint32_t)col0) : "memory");
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        while (!mbar_try(smem_u32(&bar), 0)) {}
        long long t1 = clock64();
        if (clk) clk[0] = t1 - t0;
    }
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int c = 0; c < 16; ++c) out[threadIdx.x * 16 + c] = r[c];
    // timing of a long run of shifts (second barrier phase)
    if (ntime > 0 && threadIdx.x == 0) {
        long long t0 = clock64();
        for (int i = 0; i < ntime; ++i)
            asm volatile("tcgen05.shift.cta_group::1.down [%0];" ::"r"(tmem_base + (uint32_t)((i & 1) * 8)) : "memory");
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        while (!mbar_try(smem_u32(&bar), 1)) {}
        long long t1 = clock64();
        clk[1] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tmem_base) : "memory");
}

int main() {
    uint32_t* d; long long* c;
    cudaMalloc(&d, 128 * 16 * 4); cudaMalloc(&c, 16);
    uint32_t h[128 * 16]; long long hc[2];
    for (int cfg = 0; cfg < 3; ++cfg) {
        const int nshift = cfg == 2 ? 2 : 1, col0 = cfg == 1 ? 8 : 0;
        shift_kernel<<<1, 128>>>(d, nshift, col0, c, cfg == 0 ? 2000 : 0);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost); cudaMemcpy(hc, c, 16, cudaMemcpyDeviceToHost);
        printf("== %d shift(s) at column %d (value = 100*lane + col before the shift)\n", nshift, col0);
        for (int l : {0, 1, 2, 3, 30, 31, 32, 33, 63, 64, 65, 95, 96, 97, 126, 127})
            printf("lane %3d: col0 %5u col7 %5u | col8 %5u col15 %5u\n", l, h[l * 16], h[l * 16 + 7], h[l * 16 + 8], h[l * 16 + 15]);
        if (cfg == 0) printf("1 shift + commit round trip %lld clk; 2000 back-to-back shifts %.1f clk each\n", hc[0], (double)hc[1] / 2000);
    }
    return 0;
}
