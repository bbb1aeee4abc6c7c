// Microbenchmark: cost of moving 32 fp32 values per thread to the neighbouring lanes (the R-recombination of conv_mma_kernel's
// epilogue) by (A) rotate shuffles, (B) a shared-memory round trip with 128-bit accesses on 80-byte padded rows.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/shfl_microbench scripts/shfl_microbench.cu && scripts/shfl_microbench
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(float* out, int iters, long long* clk) {
    extern __shared__ __align__(16) float sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = (float)(threadIdx.x * 32 + i);
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
            const int up = (lane + 31) & 31, dn = (lane + 1) & 31;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float u = __shfl_sync(0xffffffffu, v[i], up);
                const float d = __shfl_sync(0xffffffffu, v[16 + i], dn);
                acc[i] += u + d;
            }
        } else {
            // rows of 20 floats (80 B): conflict-free for 128-bit accesses; a set of 4 warps shares a 128-row buffer
            float* buf = sm + (warp >> 2) * (130 * 40);
            const int row = (warp & 3) * 32 + lane + 1;
            float4* w0 = reinterpret_cast<float4*>(buf + row * 40);
#pragma unroll
            for (int i = 0; i < 4; ++i) w0[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
#pragma unroll
            for (int i = 0; i < 4; ++i) w0[5 + i] = make_float4(v[16 + 4 * i], v[17 + 4 * i], v[18 + 4 * i], v[19 + 4 * i]);
            asm volatile("bar.sync %0, 128;" ::"r"(1 + (warp >> 2)) : "memory");
            const float4* ru = reinterpret_cast<const float4*>(buf + (row - 1) * 40);
            const float4* rd = reinterpret_cast<const float4*>(buf + (row + 1) * 40 + 20);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float4 a = ru[i], b = rd[i];
                acc[4 * i] += a.x + b.x; acc[4 * i + 1] += a.y + b.y; acc[4 * i + 2] += a.z + b.z; acc[4 * i + 3] += a.w + b.w;
            }
            if (MODE == 1) asm volatile("bar.sync %0, 128;" ::"r"(1 + (warp >> 2)) : "memory");
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] += acc[i] * 1e-9f;
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

int main() {
    float* out; long long* clk;
    cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&clk, 148 * 8);
    const int iters = 2000;
    const size_t smem = 4 * 130 * 40 * 4;
    cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    for (int warps = 4; warps <= 16; warps *= 2) {
        for (int mode = 0; mode < 3; ++mode) {
            if (mode == 0) k<0><<<148, warps * 32, 0>>>(out, iters, clk);
            if (mode == 1) k<1><<<148, warps * 32, smem>>>(out, iters, clk);
            if (mode == 2) k<2><<<148, warps * 32, smem>>>(out, iters, clk);
            cudaError_t e = cudaDeviceSynchronize();
            long long c[148]; cudaMemcpy(c, clk, sizeof(c), cudaMemcpyDeviceToHost);
            printf("%s warps %2d: %s  %.1f clk per (warp-iteration of 32 values moved) -> %.1f clk per SM per 128-row x 16-channel recombination\n",
                   mode == 0 ? "shuffle   " : mode == 1 ? "smem 2 bar" : "smem 1 bar", warps, cudaGetErrorString(e),
                   (double)c[0] / iters, (double)c[0] / iters * 4.0 / warps * 1.0);
        }
    }
    return 0;
}
