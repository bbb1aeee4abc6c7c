// ReverseNorm on the device: network-space voxels -> deposited energies.
//   ReverseNormCaloChall  utils/utils.py:446-573      (datasets 2 / 3 in network shape; dataset 0 / 1 in their original flat shape)
//   ReverseNormHGCal      utils/HGCal_utils.py:167-292 (HGCal, optionally with the GLaM decoding in the middle)
// Both are the same three steps on every shower, differing only in constants and in how a shower is cut into layers:
//     data = inverse(v * a_scale + a_shift)        inverse = reverse_logit(alpha) | exp(min(., log_cap)) | identity
//     data = data * post_scale + post_shift         (HGCal converter normalisation, HGCal_utils.py:660-661)
//     ['layer' maps]  data<0 -> 0;  per (shower, segment): data *= target / sum(data)   (guards: either < guard_eps -> 1)
//                     target = reverse_logit(layers*std+mean) / sum_z(...) * (totalE*std+mean)
//     data *= max_deposit * E_inc,  E_inc = emin*(emax/emin)^e (logE) or emin + (emax-emin)*e;   data < ecut -> 0
// A "segment" is one calorimeter layer of the shower: Z*P*R grids have uniform segments, the flat Dataset-1 voxel vector has
// the ragged layer boundaries of its binning file (utils.py:549-558), HGCal cell vectors have max_ncell cells per layer.
// One block per (shower, segment): the segment is read once, reduced with warp shuffles, rescaled and written once
// (8 B/voxel of HBM traffic), replacing a device->host copy plus ~10 numpy passes on the host.
#include "cd_kernels.cuh"

namespace cd {
namespace {

__device__ __forceinline__ float reverse_logit_f(float x, float alpha) {
    const float e = expf(x);
    const float o = e / (1.f + e);
    return (o - alpha) / (1.f - 2.f * alpha);
}

__device__ __forceinline__ float inverse_map(const RevNormParams& p, float v) {
    float x = v * p.a_scale + p.a_shift;
    if (p.map_kind == 0) x = reverse_logit_f(x, p.alpha);
    else if (p.map_kind == 1) x = expf(fminf(x, p.log_cap));
    x = x * p.post_scale + p.post_shift;
    if (p.layerE && x < 0.f) x = 0.f;
    return x;
}

__global__ void __launch_bounds__(256) reverse_norm_kernel(RevNormParams p) {
    __shared__ float red[8];
    __shared__ float s_target, s_fac;
    const int z = blockIdx.x, b = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = p.total / p.nseg;
    const int lo = p.seg ? p.seg[z] : z * per, hi = p.seg ? p.seg[z + 1] : (z + 1) * per;
    const int n = hi - lo;
    const float* v = p.vox + (long long)b * p.total + lo;
    float* o = p.out + (long long)b * p.total + lo;
    const float ev = p.e[(long long)b * p.e_stride];
    const float energy = p.energy_mode == 0 ? p.emin * powf(p.emax / p.emin, ev) : p.emin + (p.emax - p.emin) * ev;
    if (z == 0 && tid == 0 && p.energy_out) p.energy_out[b] = energy;
    if (p.layerE && warp == 0) {
        // target energy of this layer: normalised reverse-logit of the conditioning layer fractions times the total
        const float* le = p.layerE + (long long)b * (p.nseg + 1);
        float s = 0.f, mine = 0.f;
        for (int i = lane; i < p.nseg; i += 32) {
            const float t = reverse_logit_f(le[1 + i] * p.layers_std + p.layers_mean, p.alpha);
            s += t;
            if (i == z) mine = t;
        }
#pragma unroll
        for (int k = 16; k > 0; k >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, k); mine += __shfl_xor_sync(0xffffffffu, mine, k); }
        if (lane == 0) s_target = mine / s * (le[0] * p.totalE_std + p.totalE_mean);
    }
    // pass 1: transform (kept in registers for segments up to 256*8 values) and segment sum
    constexpr int kMaxPer = 8;
    float d[kMaxPer];
    float sum = 0.f;
    const bool fits = n <= 256 * kMaxPer;
#pragma unroll
    for (int k = 0; k < kMaxPer; ++k) {
        const int i = tid + k * 256;
        const float x = i < n ? inverse_map(p, v[i]) : 0.f;
        d[k] = x;
        sum += x;
    }
    if (!fits)
        for (int i = tid + kMaxPer * 256; i < n; i += 256) sum += inverse_map(p, v[i]);
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
            if (s_target < p.guard_eps || prev < p.guard_eps) f = 1.f;      // utils.py:538-540, HGCal_utils.py:262-264
            s_fac = f;
        }
        __syncthreads();
        fac = s_fac;
    }
    const float scale = p.final_scale ? fac * p.max_deposit * energy : fac;
#pragma unroll
    for (int k = 0; k < kMaxPer; ++k) {
        const int i = tid + k * 256;
        if (i < n) {
            float y = d[k] * scale;
            if (p.ecut > 0.f && y < p.ecut) y = 0.f;
            o[i] = y;
        }
    }
    if (!fits)
        for (int i = tid + kMaxPer * 256; i < n; i += 256) {
            float y = inverse_map(p, v[i]) * scale;
            if (p.ecut > 0.f && y < p.ecut) y = 0.f;
            o[i] = y;
        }
}

}  // namespace

void launch_reverse_norm(const RevNormParams& p, int B, cudaStream_t st) {
    dim3 grid(p.nseg, B);
    reverse_norm_kernel<<<grid, 256, 0, st>>>(p);
}

}  // namespace cd
