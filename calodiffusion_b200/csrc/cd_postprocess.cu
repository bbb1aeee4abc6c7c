// ReverseNormCaloChall on the device (utils/utils.py:446-573; called from models/diffusion.py:171-191): network-space
// voxels -> deposited energies, for the CaloChallenge '*logit-norm' shower maps (dataset 2 / 3):
//     data = reverse_logit(v * logit_std + logit_mean)
//     ['layer' maps]  data<0 -> 0;  per (shower, layer): data *= target_layer_E / sum(data)   (guards: either < 1e-6 -> 1)
//                     target = reverse_logit(layers*std+mean) / sum_z(...) * (totalE*std+mean)
//     data *= max_deposit * E_inc,  E_inc = emin * (emax/emin)^e;   data < ecut -> 0
// One block per (shower, calorimeter layer): the layer's P*R voxels are read once, reduced with warp shuffles, rescaled
// and written once (8 B/voxel of HBM traffic), replacing a device->host copy plus ~10 numpy passes on the host.
#include "cd_kernels.cuh"

namespace cd {
namespace {

__device__ __forceinline__ float reverse_logit_f(float x) {
    const float e = expf(x);
    const float o = e / (1.f + e);
    return (o - 1e-6f) / (1.f - 2e-6f);
}

__global__ void __launch_bounds__(256) reverse_norm_kernel(RevNormParams p) {
    __shared__ float red[8];
    __shared__ float s_target, s_fac;
    const int z = blockIdx.x, b = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* v = p.vox + ((long long)b * p.Z + z) * p.PR;
    float* o = p.out + ((long long)b * p.Z + z) * p.PR;
    const float energy = p.emin * powf(p.emax / p.emin, p.e[b]);
    if (z == 0 && tid == 0 && p.energy_out) p.energy_out[b] = energy;
    if (p.layerE && warp == 0) {
        // target energy of this layer: normalised reverse-logit of the conditioning layer fractions times the total
        const float* le = p.layerE + (long long)b * (p.Z + 1);
        float s = 0.f, mine = 0.f;
        for (int i = lane; i < p.Z; i += 32) {
            const float t = reverse_logit_f(le[1 + i] * p.layers_std + p.layers_mean);
            s += t;
            if (i == z) mine = t;
        }
#pragma unroll
        for (int k = 16; k > 0; k >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, k); mine += __shfl_xor_sync(0xffffffffu, mine, k); }
        if (lane == 0) s_target = mine / s * (le[0] * p.totalE_std + p.totalE_mean);
    }
    // pass 1: transform (kept in registers for layers up to 256*8 voxels) and layer sum
    constexpr int kMaxPer = 8;
    float d[kMaxPer];
    float sum = 0.f;
    const bool fits = p.PR <= 256 * kMaxPer;
#pragma unroll
    for (int k = 0; k < kMaxPer; ++k) {
        const int i = tid + k * 256;
        float x = 0.f;
        if (i < p.PR) {
            x = reverse_logit_f(v[i] * p.logit_std + p.logit_mean);
            if (p.layerE && x < 0.f) x = 0.f;
        }
        d[k] = x;
        sum += x;
    }
    if (!fits)
        for (int i = tid + kMaxPer * 256; i < p.PR; i += 256) {
            float x = reverse_logit_f(v[i] * p.logit_std + p.logit_mean);
            if (p.layerE && x < 0.f) x = 0.f;
            sum += x;
        }
    float fac = 1.f;
    if (p.layerE) {
#pragma unroll
        for (int k = 16; k > 0; k >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, k);
        if (lane == 0) red[warp] = sum;
        __syncthreads();
        if (tid == 0) {
            float prev = 0.f;
            for (int w = 0; w < 8; ++w) prev += red[w];
            float f = s_target / (prev + 1e-10f);
            if (s_target < 1e-6f || prev < 1e-6f) f = 1.f;      // utils.py:538-540
            s_fac = f;
        }
        __syncthreads();
        fac = s_fac;
    }
    const float scale = fac * p.max_deposit * energy;
#pragma unroll
    for (int k = 0; k < kMaxPer; ++k) {
        const int i = tid + k * 256;
        if (i < p.PR) {
            float y = d[k] * scale;
            if (p.ecut > 0.f && y < p.ecut) y = 0.f;
            o[i] = y;
        }
    }
    if (!fits)
        for (int i = tid + kMaxPer * 256; i < p.PR; i += 256) {
            float x = reverse_logit_f(v[i] * p.logit_std + p.logit_mean);
            if (p.layerE && x < 0.f) x = 0.f;
            float y = x * scale;
            if (p.ecut > 0.f && y < p.ecut) y = 0.f;
            o[i] = y;
        }
}

}  // namespace

void launch_reverse_norm(const RevNormParams& p, int B, cudaStream_t st) {
    dim3 grid(p.Z, B);
    reverse_norm_kernel<<<grid, 256, 0, st>>>(p);
}

}  // namespace cd
