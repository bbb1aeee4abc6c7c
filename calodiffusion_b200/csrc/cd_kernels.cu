// Hand-written CUDA kernels (sm_100a) for the CaloDiffusion denoiser hot path — everything except the
// tcgen05 implicit-GEMM convolution, which lives in cd_conv_mma.cu.
#include "cd_kernels.cuh"

namespace cd {

// =================================================================================================
// prep_input: x (B,Z,P,R) fp32 -> PLCP tensor with 8 channels: [x*c_in, R, Z, (phi), 0...]
// Replaces calodiffusion.py:121-142 add_RZPhi (repeat+cat) and the c_in multiply of :159.
// =================================================================================================
__global__ void __launch_bounds__(256) prep_input_kernel(Geom g, const float* __restrict__ x, const float* __restrict__ sigma,
                                                         int sigma_stride, float sigma_data, const float* __restrict__ r_vals,
                                                         int rz_input, int phi_input, __half* __restrict__ dst, int out_chunks) {
    const int b = blockIdx.y;
    const uint32_t rem = blockIdx.x * 256 + threadIdx.x;
    if (rem >= (uint32_t)g.rows_per_sample) return;
    const SubRow sr = decode_subrow(g, rem);
    const long long row = (long long)b * g.rows_per_sample + rem;
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (sr.kind != 2) {
        int p = sr.pp - 1;
        if (p < 0) p += g.P;
        if (p >= g.P) p -= g.P;
        const float s = sigma[(long long)b * sigma_stride];
        const float c_in = sigma_data > 0.f ? rsqrtf(s * s + sigma_data * sigma_data) : 1.f;
        const long long vox = (((long long)b * g.Z + sr.z) * g.P + p) * g.R + sr.rr;
        f[0] = x[vox] * c_in;
        int c = 1;
        if (rz_input) { f[c++] = r_vals[sr.rr]; f[c++] = (float)sr.z / (float)g.Z; }
        if (phi_input) { f[c++] = (g.P > 1) ? (float)p / (float)(g.P - 1) : 0.f; }
    }
    store8(dst + plcp_offset(g, 0, row), pack8(f));
    // channel padding up to the tensor-core K granularity (the buffer may have been used by another layer)
    for (int k = 1; k < out_chunks; ++k) *reinterpret_cast<uint4*>(dst + plcp_offset(g, k, row)) = make_uint4(0u, 0u, 0u, 0u);
}

void launch_prep_input(const Geom& g, const float* x, const float* sigma, int sigma_stride, float sigma_data,
                       const float* r_vals, int rz_input, int phi_input, __half* dst, int out_chunks, int B, cudaStream_t st) {
    dim3 grid((g.rows_per_sample + 255) / 256, B);
    prep_input_kernel<<<grid, 256, 0, st>>>(g, x, sigma, sigma_stride, sigma_data, r_vals, rz_input, phi_input, dst, out_chunks);
}

// prep_input with the in-loop geometry embedding: the grid value is  sum_k val[k] * x_flat[col[k]]  over the CSR row of
// the grid cell (NNConverter.enc utils/utils.py:610-634; HGCal Embeder.forward HGCal_utils.py:315-320).
__global__ void __launch_bounds__(256) prep_input_geom_kernel(Geom g, GeomMap enc, const float* __restrict__ x,
                                                              const float* __restrict__ sigma, int sigma_stride, float sigma_data,
                                                              const float* __restrict__ r_vals, int rz_input, int phi_input,
                                                              __half* __restrict__ dst, int out_chunks) {
    const int b = blockIdx.y;
    const uint32_t rem = blockIdx.x * 256 + threadIdx.x;
    if (rem >= (uint32_t)g.rows_per_sample) return;
    const SubRow sr = decode_subrow(g, rem);
    const long long row = (long long)b * g.rows_per_sample + rem;
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (sr.kind != 2) {
        int p = sr.pp - 1;
        if (p < 0) p += g.P;
        if (p >= g.P) p -= g.P;
        const float s = sigma[(long long)b * sigma_stride];
        const float c_in = sigma_data > 0.f ? rsqrtf(s * s + sigma_data * sigma_data) : 1.f;
        const int cell = (sr.z * g.P + p) * g.R + sr.rr;
        const float* xb = x + (long long)b * enc.n_cols;
        float acc = 0.f;
        for (int k = enc.rowptr[cell], ke = enc.rowptr[cell + 1]; k < ke; ++k) acc = fmaf(enc.val[k], xb[enc.col[k]], acc);
        f[0] = acc * c_in;
        int c = 1;
        if (rz_input) { f[c++] = r_vals[sr.rr]; f[c++] = (float)sr.z / (float)g.Z; }
        if (phi_input) { f[c++] = (g.P > 1) ? (float)p / (float)(g.P - 1) : 0.f; }
    }
    store8(dst + plcp_offset(g, 0, row), pack8(f));
    for (int k = 1; k < out_chunks; ++k) *reinterpret_cast<uint4*>(dst + plcp_offset(g, k, row)) = make_uint4(0u, 0u, 0u, 0u);
}

void launch_prep_input_geom(const Geom& g, const GeomMap& enc, const float* x, const float* sigma, int sigma_stride,
                            float sigma_data, const float* r_vals, int rz_input, int phi_input, __half* dst, int out_chunks,
                            int B, cudaStream_t st) {
    dim3 grid((g.rows_per_sample + 255) / 256, B);
    prep_input_geom_kernel<<<grid, 256, 0, st>>>(g, enc, x, sigma, sigma_stride, sigma_data, r_vals, rz_input, phi_input, dst,
                                                 out_chunks);
}

// =================================================================================================
// Generic direct convolution (SIMT, fp32 accumulate).  Handles every CylindricalConv /
// CylindricalConvTrans variant of models/models.py:25-96 by explicit coordinate arithmetic (circular
// phi, zero Z/R), independent of the PLCP halo convention, so it doubles as the on-device cross-check
// for the tcgen05 kernel.  Thread = (output row, 8 output channels).
// =================================================================================================
template <int NSRC>
__global__ void conv_generic_kernel(ConvParams p, long long nrows) {
    long long row = (long long)blockIdx.x * 32 + threadIdx.x;
    int cog = threadIdx.y;  // output chunk
    bool active = row < nrows;
    RowInfo ri;
    ri.kind = 2; ri.b = 0; ri.z = 0; ri.pp = 0; ri.rr = 0; ri.src = 0;
    if (active) ri = decode_row(p.gout, row);
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    bool compute = active && (ri.kind == 0 || (ri.kind == 1 && p.write_halo));
    if (compute) {
        int zo = ri.z, ro = ri.rr;
        int po = ri.pp - 1;
        if (po < 0) po += p.gout.P;
        if (po >= p.gout.P) po -= p.gout.P;
        const Geom& gi = p.gin;
        for (int a = 0; a < p.kz; ++a) {
            int zi;
            if (!p.transposed) {
                zi = zo * p.sz + a - p.pz;
            } else {
                int t = zo + p.pz - a;
                if (t < 0 || (t % p.sz) != 0) continue;
                zi = t / p.sz;
            }
            if (zi < 0 || zi >= gi.Z) continue;
            for (int bq = 0; bq < p.kp; ++bq) {
                int pi;
                if (!p.transposed) {
                    int q = po * p.sp + bq;        // index into the circularly padded axis
                    pi = q - p.cp;
                } else {
                    int t = po + (p.kp - 1) - bq;  // ConvTranspose padding[1] = kp-1 (models.py:45)
                    if (t < 0 || (t % p.sp) != 0) continue;
                    int q = t / p.sp;              // index into the padded (P+2) input axis
                    if (q >= gi.P + 2 * p.cp) continue;
                    pi = q - p.cp;
                }
                pi %= gi.P;
                if (pi < 0) pi += gi.P;
                for (int c = 0; c < p.kr; ++c) {
                    int rin;
                    if (!p.transposed) {
                        rin = ro * p.sr + c - p.pr;
                    } else {
                        int t = ro + p.pr - c;
                        if (t < 0 || (t % p.sr) != 0) continue;
                        rin = t / p.sr;
                    }
                    if (rin < 0 || rin >= gi.R) continue;
                    long long irow = interior_row(gi, ri.b, zi, pi, rin);
                    int tap = (a * p.kp + bq) * p.kr + c;
                    const float* wt = p.w + ((long long)tap * p.Cin_total) * p.Cout + cog * 8;
                    int ci0 = 0;
#pragma unroll
                    for (int s = 0; s < NSRC; ++s) {
                        const __half* sb = p.src[s].base;
                        int nch = p.src[s].C >> 3;
                        for (int k = 0; k < nch; ++k) {
                            Half8 v = load8(sb + plcp_offset(gi, k, irow));
                            float f[8];
                            unpack8(v, f);
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const float4* w4 = reinterpret_cast<const float4*>(wt + (long long)(ci0 + k * 8 + j) * p.Cout);
                                float4 w0 = __ldg(w4), w1 = __ldg(w4 + 1);
                                acc[0] += f[j] * w0.x; acc[1] += f[j] * w0.y; acc[2] += f[j] * w0.z; acc[3] += f[j] * w0.w;
                                acc[4] += f[j] * w1.x; acc[5] += f[j] * w1.y; acc[6] += f[j] * w1.z; acc[7] += f[j] * w1.w;
                            }
                        }
                        ci0 += p.src[s].C;
                    }
                }
            }
        }
        if (p.bias) {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += p.bias[cog * 8 + j];
        }
    }
    if (active && (compute || p.write_halo)) {
        // zero rows are written as zeros only in write_halo mode; raw mode leaves halos untouched
        store8(p.dst + plcp_offset(p.gout, cog, row), pack8(acc));
    }
    if (p.stats) {
        // per-(sample, channel) sum / sum of squares over interior rows, of the fp16-rounded values
        bool contrib = active && ri.kind == 0;
        float v[8];
        Half8 hv = pack8(acc);
        unpack8(hv, v);
        unsigned mask = 0xffffffffu;
        int b0 = __shfl_sync(mask, ri.b, 0);
        bool uniform = __all_sync(mask, (!contrib) || ri.b == b0);
        int bfirst = b0;
        if (uniform) {
            // all contributing lanes share one sample; find it (lane 0 may be inactive)
            unsigned bal = __ballot_sync(mask, contrib);
            if (bal == 0) return;
            int src_lane = __ffs(bal) - 1;
            bfirst = __shfl_sync(mask, ri.b, src_lane);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float s = contrib ? v[j] : 0.f, q = contrib ? v[j] * v[j] : 0.f;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(mask, s, o); q += __shfl_xor_sync(mask, q, o); }
                if (threadIdx.x == 0) {
                    double* d = p.stats + ((long long)bfirst * p.Cout + cog * 8 + j) * 2;
                    atomicAdd(d, (double)s);
                    atomicAdd(d + 1, (double)q);
                }
            }
        } else if (contrib) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                double* d = p.stats + ((long long)ri.b * p.Cout + cog * 8 + j) * 2;
                atomicAdd(d, (double)v[j]);
                atomicAdd(d + 1, (double)(v[j] * v[j]));
            }
        }
    }
}

void launch_conv_generic(const ConvParams& p, cudaStream_t st) {
    long long nrows = (long long)p.B * p.gout.rows_per_sample;
    dim3 block(32, p.Cout / 8);
    long long blocks = (nrows + 31) / 32;
    if (p.nsrc == 1) conv_generic_kernel<1><<<(unsigned)blocks, block, 0, st>>>(p, nrows);
    else conv_generic_kernel<2><<<(unsigned)blocks, block, 0, st>>>(p, nrows);
}

// =================================================================================================
// GroupNorm finalize + apply (models/models.py:155,162 GroupNorm(8,C); :293,325 GroupNorm(1,C); eps 1e-5).
// Statistics are per-(sample, channel) double sums (sum, sum of squares) produced by the conv epilogues or by
// the producer noted at the call site; every block folds its sample's sums per group into a per-channel
// affine y = x*scale + shift (shared memory), then streams rows:
//     out = act(x*scale+shift) + addvec[b][c] + residual          written for every row incl. halos
// Covers Block's GN+SiLU (models.py:160-169), the additive conditioning (:194-197), the ResnetBlock residual
// (:200), PreNorm (:327-329) and Residual(to_out GN) (:117,:293).  HBM-bound: one thread owns one padded-linear
// row (decoded once with 32-bit fast divisions) and moves all of its 16-byte channel chunks, four loads in
// flight per operand; a warp's access to one chunk plane is 512 contiguous bytes.
// Optionally accumulates the per-sample sum / sum of squares of its own (fp16-rounded) output over interior rows
// into out_stats[b][channel 0] for the GroupNorm(1, C) of the attention block that follows (models.py:325).
// =================================================================================================
constexpr int kGnThreads = 256;
constexpr int kGnRowsPerThread = 4;      // rows per thread: amortises the per-block statistics prologue
constexpr int kGnMaxC = 256;

// SiLU through one MUFU: x*sigmoid(x) = 0.5*x*(1 + tanh(x/2)); tanh.approx error (2^-11) is below the fp16 store rounding
__device__ __forceinline__ float silu_fast(float v) {
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * v));
    const float h = 0.5f * v;
    return fmaf(h, t, h);
}

__global__ void __launch_bounds__(kGnThreads, 4) gn_apply_kernel(
    Geom g, const __half* __restrict__ src, const double* __restrict__ stats, const float* __restrict__ gamma,
    const float* __restrict__ beta, int groups, double count_per_channel, const float* __restrict__ addvec, int addvec_stride,
    const __half* __restrict__ residual, const __half* __restrict__ residual2, int res_split, __half* __restrict__ dst, int C,
    int act, double* __restrict__ out_stats) {
    __shared__ __align__(16) float s_scale[kGnMaxC];
    __shared__ __align__(16) float s_shift[kGnMaxC];
    __shared__ __align__(16) float s_add[kGnMaxC];
    __shared__ float red[2][kGnThreads / 32];
    const int b = blockIdx.y;
    const int nch = C >> 3;
    for (int c = threadIdx.x; c < C; c += kGnThreads) {
        const int cpg = C / groups;
        const int g0 = (c / cpg) * cpg;
        double s = 0.0, q = 0.0;
        for (int j = 0; j < cpg; ++j) {
            const double* d = stats + ((long long)b * C + g0 + j) * 2;
            s += d[0]; q += d[1];
        }
        const double n = count_per_channel * cpg;
        const float mean = (float)(s / n);
        float var = (float)(q / n - (s / n) * (s / n));
        if (var < 0.f) var = 0.f;
        const float sc = gamma[c] * rsqrtf(var + 1e-5f);
        const float av = addvec ? addvec[(long long)b * addvec_stride + c] : 0.f;
        s_scale[c] = sc;
        // without an activation the additive conditioning folds into the shift
        s_shift[c] = beta[c] - mean * sc + (act ? 0.f : av);
        s_add[c] = act ? av : 0.f;
    }
    __syncthreads();
    float ssum = 0.f, ssq = 0.f;
    const uint32_t rem0 = blockIdx.x * (kGnThreads * kGnRowsPerThread) + threadIdx.x;
#pragma unroll 1
    for (int it = 0; it < kGnRowsPerThread; ++it) {
        const uint32_t rem = rem0 + it * kGnThreads;
        if (rem >= (uint32_t)g.rows_per_sample) break;
        const SubRow sr = decode_subrow(g, rem);
        const long long row = (long long)b * g.rows_per_sample + rem;
        const long long srow = row + sr.src_delta;
        if (sr.kind == 2) {
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);
            for (int k = 0; k < nch; ++k) *reinterpret_cast<uint4*>(dst + plcp_offset(g, k, row)) = z;
            continue;
        }
        for (int k0 = 0; k0 < nch; k0 += 4) {
            Half8 v[4], rv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (k0 + i < nch) v[i] = load8(src + plcp_offset(g, k0 + i, srow));
            if (residual) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (k0 + i < nch) {
                        const int k = k0 + i;
                        const __half* rb = (k < res_split) ? residual + plcp_offset(g, k, srow) : residual2 + plcp_offset(g, k - res_split, srow);
                        rv[i] = load8(rb);
                    }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (k0 + i >= nch) break;
                const int k = k0 + i;
                float f[8], sc[8], sh[8];
                unpack8(v[i], f);
                *reinterpret_cast<float4*>(sc) = *reinterpret_cast<const float4*>(s_scale + k * 8);
                *reinterpret_cast<float4*>(sc + 4) = *reinterpret_cast<const float4*>(s_scale + k * 8 + 4);
                *reinterpret_cast<float4*>(sh) = *reinterpret_cast<const float4*>(s_shift + k * 8);
                *reinterpret_cast<float4*>(sh + 4) = *reinterpret_cast<const float4*>(s_shift + k * 8 + 4);
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = fmaf(f[j], sc[j], sh[j]);
                if (act) {
                    float ad[8];
                    *reinterpret_cast<float4*>(ad) = *reinterpret_cast<const float4*>(s_add + k * 8);
                    *reinterpret_cast<float4*>(ad + 4) = *reinterpret_cast<const float4*>(s_add + k * 8 + 4);
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] = silu_fast(f[j]) + ad[j];
                }
                if (residual) {
                    float r[8];
                    unpack8(rv[i], r);
#pragma unroll
                    for (int j = 0; j < 8; ++j) f[j] += r[j];
                }
                const Half8 o = pack8(f);
                store8(dst + plcp_offset(g, k, row), o);
                if (out_stats && sr.kind == 0) {
                    unpack8(o, f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) { ssum += f[j]; ssq = fmaf(f[j], f[j], ssq); }
                }
            }
        }
    }
    if (out_stats) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { ssum += __shfl_xor_sync(0xffffffffu, ssum, o); ssq += __shfl_xor_sync(0xffffffffu, ssq, o); }
        const int w = threadIdx.x >> 5;
        if ((threadIdx.x & 31) == 0) { red[0][w] = ssum; red[1][w] = ssq; }
        __syncthreads();
        if (threadIdx.x < 2) {
            double t = 0.0;
            for (int i = 0; i < kGnThreads / 32; ++i) t += (double)red[threadIdx.x][i];
            atomicAdd(out_stats + (long long)b * C * 2 + threadIdx.x, t);
        }
    }
}

void launch_gn_apply(const Geom& g, const __half* src, const double* stats, const float* gamma, const float* beta, int groups,
                     double count_per_channel, const float* addvec, int addvec_stride, const __half* residual,
                     const __half* residual2, int res_split, __half* dst, int C, int act, double* out_stats, int B,
                     cudaStream_t st) {
    const int rpb = kGnThreads * kGnRowsPerThread;
    dim3 grid((g.rows_per_sample + rpb - 1) / rpb, B);
    gn_apply_kernel<<<grid, kGnThreads, 0, st>>>(g, src, stats, gamma, beta, groups, count_per_channel, addvec, addvec_stride,
                                                residual, residual2, residual2 ? res_split : (C >> 3), dst, C, act, out_stats);
}

// =================================================================================================
// Conditioning: time MLP + cond MLP (models/models.py:575-608, :704-707) and every ResnetBlock's
// SiLU -> Linear(cond_dim, C) projection (:176-178, :195), one block per sample.
// time input = do_time_embed(sigma) (calodiffusion.py:144-155).
// =================================================================================================
// out[o] = act(bias[o] + sum_i W[o][i] in[i]): one warp per output row so the weight reads are coalesced
// The kernel is a chain of small matrix-vector products, one block per shower: what it costs is the number of dependent
// global-load round trips, so a warp works on FOUR output rows at a time (four independent load streams and reductions).
__device__ __forceinline__ void warp_rows4(const float* __restrict__ W, const float* in, int n_in, int n_out, int o0, int stride,
                                           float (&a)[4]) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int u = 0; u < 4; ++u) a[u] = 0.f;
    for (int i = lane; i < n_in; i += 32) {
        const float x = in[i];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int o = o0 + u * stride;
            if (o < n_out) a[u] = fmaf(__ldg(W + (long long)o * n_in + i), x, a[u]);
        }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1)
#pragma unroll
        for (int u = 0; u < 4; ++u) a[u] += __shfl_xor_sync(0xffffffffu, a[u], s);
}
__device__ void dense_layer(const float* __restrict__ W, const float* __restrict__ bias, const float* in, float* out,
                            int n_in, int n_out, int act) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int o0 = warp; o0 < n_out; o0 += 4 * nwarps) {
        float a[4];
        warp_rows4(W, in, n_in, n_out, o0, nwarps, a);
        if (lane == 0) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int o = o0 + u * nwarps;
                if (o >= n_out) break;
                float v = a[u] + (bias ? bias[o] : 0.f);
                if (act == 1) v = gelu_erf(v);
                out[o] = v;
            }
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256) embed_kernel(EmbedParams p, const float* __restrict__ sigma, int sigma_stride,
                                                    const float* __restrict__ cond, float* __restrict__ addvec) {
    __shared__ float buf0[256], buf1[256], cvec[256];
    int b = blockIdx.x;
    int half_dim = p.cond_dim / 2;
    if (threadIdx.x == 0) {
        float s = sigma[(long long)b * sigma_stride];
        buf0[0] = (p.time_mode == 0) ? s / sqrtf(1.f + s * s) : (p.time_mode == 1 ? 0.5f * logf(s) : s);
    }
    __syncthreads();
    dense_layer(p.tw[0], p.tb[0], buf0, buf1, 1, half_dim / 2, 1);
    dense_layer(p.tw[1], p.tb[1], buf1, buf0, half_dim / 2, half_dim, 1);
    dense_layer(p.tw[2], p.tb[2], buf0, cvec, half_dim, half_dim, 0);  // conditions[0:half] = t
    for (int i = threadIdx.x; i < p.cond_size; i += blockDim.x) buf0[i] = cond[(long long)b * p.cond_size + i];
    __syncthreads();
    dense_layer(p.cw[0], p.cb[0], buf0, buf1, p.cond_size, p.cond_hidden, 1);
    dense_layer(p.cw[1], p.cb[1], buf1, buf0, p.cond_hidden, half_dim, 1);
    dense_layer(p.cw[2], p.cb[2], buf0, cvec + half_dim, half_dim, half_dim, 0);  // conditions[half:] = c
    for (int i = threadIdx.x; i < p.cond_dim; i += blockDim.x) cvec[i] = silu_f(cvec[i]);
    __syncthreads();
    // every ResnetBlock's Linear(cond_dim, C) on the shared SiLU(conditions) (models.py:176-178,195): warp per output
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int k = 0; k < p.n_proj; ++k) {
        const float* W = p.pw[k];
        const float* bias = p.pb[k];
        const int C = p.pC[k];
        float* out = addvec + (long long)b * p.addvec_stride + p.poff[k];
        for (int o0 = warp; o0 < C; o0 += 4 * nwarps) {
            float a[4];
            warp_rows4(W, cvec, p.cond_dim, C, o0, nwarps, a);
            if (lane == 0) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int o = o0 + u * nwarps;
                    if (o < C) out[o] = a[u] + bias[o];
                }
            }
        }
    }
}

void launch_embed(const EmbedParams& p, const float* sigma, int sigma_stride, const float* cond, float* addvec, int B,
                  cudaStream_t st) {
    embed_kernel<<<B, 256, 0, st>>>(p, sigma, sigma_stride, cond, addvec);
}

// =================================================================================================
// Philox4x32-10 -> standard normal (Box-Muller), keyed by (seed, global shower index, step, voxel) so a
// shower's noise does not depend on how the batch is sharded over GPUs.
// =================================================================================================
__device__ __forceinline__ void philox_round(uint32_t& c0, uint32_t& c1, uint32_t& c2, uint32_t& c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
}
__device__ __forceinline__ float philox_normal(uint64_t seed, uint32_t shower, uint32_t step, uint32_t voxel) {
    uint32_t c0 = voxel, c1 = shower, c2 = step, c3 = 0x43414C4Fu;
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        philox_round(c0, c1, c2, c3, k0, k1);
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    float u1 = ((float)c0 + 1.0f) * 2.3283064365386963e-10f;  // (0,1]
    float u2 = (float)c1 * 2.3283064365386963e-10f;
    return sqrtf(-2.0f * __logf(u1)) * __cosf(6.283185307179586f * u2);
}

// shared tail of the final kernels: EDM pre-conditioning output + sampler update for caller-facing voxel v = b*vps + rem
__device__ __forceinline__ void final_tail(const FinalParams& p, long long v, int b, int rem, float pred) {
    float s = p.sigma[(long long)b * p.sigma_stride];
    float sd = p.sigma_data;
    float den = s * s + sd * sd;
    float xin = p.x ? p.x[v] : 0.f;
    float x0 = sd > 0.f ? (sd * sd / den) * xin + (s * sd * rsqrtf(den)) * pred : pred;
    if (p.pred_out) p.pred_out[v] = pred;
    if (p.x0_out) p.x0_out[v] = x0;
    if (p.d_out) p.d_out[v] = p.coef[7] * (xin - x0);
    if (p.fault && !isfinite(pred)) atomicMin(p.fault, p.eval_index);   // activations left the fp16 range somewhere upstream
    if (p.x_out) {
        const float* k = p.coef;
        float o = k[0] * x0 + k[1] * xin;
        if (p.a1) o += k[2] * p.a1[v];
        else if (p.rng_enabled && k[2] != 0.f) {
            const uint64_t seed = p.rng_key ? (((uint64_t)p.rng_key[1] << 32) | p.rng_key[0]) : p.seed;
            const int so = p.rng_key ? (int)p.rng_key[2] : p.shower_offset;
            o += k[2] * philox_normal(seed, (uint32_t)(so + b), (uint32_t)p.step, (uint32_t)rem);
        }
        if (p.a2) o += k[3] * p.a2[v];
        if (p.a3) o += k[4] * p.a3[v];
        p.x_out[v] = o;
    }
}

// =================================================================================================
// Final 1x1x1 conv (C->1, models.py:696) fused with the EDM pre-conditioning output
//   x0 = c_skip*x + c_out*pred          (calodiffusion.py:166-167, loss.py:33-37)
// and the sampler's per-step update
//   x_out = k0*x0 + k1*x + k2*a1 + k3*a2 + k4*a3      (sample.py:90-104 DDIM/DDPM; :784-821 Euler/Heun)
// where a1 may be in-kernel Philox noise.  Thread = interior voxel.
// =================================================================================================
__global__ void final_kernel(FinalParams p, long long nvox) {
    long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nvox) return;
    const Geom& g = p.g;
    int vps = g.Z * g.P * g.R;
    int b = (int)(v / vps);
    int rem = (int)(v - (long long)b * vps);
    int z = rem / (g.P * g.R);
    int rem2 = rem - z * g.P * g.R;
    int ph = rem2 / g.R;
    int r = rem2 - ph * g.R;
    long long row = interior_row(g, b, z, ph, r);
    float pred = p.bias_ptr ? *p.bias_ptr : p.bias;
    int nch = p.C >> 3;
    for (int k = 0; k < nch; ++k) {
        Half8 hv = load8(p.h + plcp_offset(g, k, row));
        float f[8];
        unpack8(hv, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) pred += f[j] * p.w[k * 8 + j];
    }
    final_tail(p, v, b, rem, pred);
}

void launch_final(const FinalParams& p, int B, cudaStream_t st) {
    long long nvox = (long long)B * p.g.Z * p.g.P * p.g.R;
    final_kernel<<<(unsigned)((nvox + 255) / 256), 256, 0, st>>>(p, nvox);
}

// Geometry decoding fused with the pre-conditioning output and the sampler update: thread = flat voxel j of shower b,
//   pred[b,j] = sum_k val[k] * grid_pred[b, col[k]]      (NNConverter.dec utils/utils.py:636-653; HGCal Decoder.forward
//   HGCal_utils.py:340-349 without sparse decoding), then the same tail as final_kernel on the (B, n_flat) tensors.
__global__ void __launch_bounds__(256) geom_decode_kernel(FinalParams p, GeomMap dec, const float* __restrict__ grid_pred,
                                                          long long nvox) {
    long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nvox) return;
    const int b = (int)(v / dec.n_rows);
    const int j = (int)(v - (long long)b * dec.n_rows);
    const float* gb = grid_pred + (long long)b * dec.n_cols;
    float pred = 0.f;
    for (int k = dec.rowptr[j], ke = dec.rowptr[j + 1]; k < ke; ++k) pred = fmaf(dec.val[k], gb[dec.col[k]], pred);
    final_tail(p, v, b, j, pred);
}

// stand-alone application of one geometry map: y[b, i] = sum_k val[k] * x[b, col[k]]  (the converters' enc / dec outside the loop)
__global__ void __launch_bounds__(256) geom_apply_kernel(GeomMap m, const float* __restrict__ x, float* __restrict__ y, long long n) {
    long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const int b = (int)(v / m.n_rows);
    const int i = (int)(v - (long long)b * m.n_rows);
    const float* xb = x + (long long)b * m.n_cols;
    float acc = 0.f;
    for (int k = m.rowptr[i], ke = m.rowptr[i + 1]; k < ke; ++k) acc = fmaf(m.val[k], xb[m.col[k]], acc);
    y[v] = acc;
}

void launch_geom_apply(const GeomMap& m, const float* x, float* y, int B, cudaStream_t st) {
    long long n = (long long)B * m.n_rows;
    geom_apply_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(m, x, y, n);
}

// Sampled GLaM decoding (generate_sparse_mat, HGCal_utils.py:355-407, applied as in Decoder.forward :344-346).
// Thread = (shower b, grid bin c).  Over the bin's positive decoder entries (cells n with w > 1e-6): r = u + w with
// u ~ U[0,1); the entries with r > 1 and the one with the largest r are kept, and the bin's energy is split equally
// between them (atomicAdd into the cell vector: several bins feed one cell).
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint32_t a, uint32_t b, uint32_t c) {
    uint32_t c0 = c, c1 = a, c2 = b, c3 = 0x53504152u;
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        philox_round(c0, c1, c2, c3, k0, k1);
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return (float)(c0 >> 8) * 5.9604644775390625e-8f;     // [0, 1): 24 random bits, like torch.rand for float32
}

__global__ void __launch_bounds__(256) sparse_decode_kernel(SparseDecParams p, long long n) {
    long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const int b = (int)(v / p.n_cols);
    const int c = (int)(v - (long long)b * p.n_cols);
    const int k0 = p.colptr[c], k1 = p.colptr[c + 1];
    if (k0 == k1) return;                                  // bin without cells: its energy is dropped (mask, :393)
    const float x = p.x[v];
    const int rb = p.per_batch ? 0 : b;
    const int l = c / p.bins_per_layer, e = c - l * p.bins_per_layer;
    auto draw = [&](int k) -> float {
        const int nloc = p.row[k] - l * p.cells_per_layer;
        float u;
        if (p.u_bank) u = p.u_bank[(((long long)rb * (p.n_cols / p.bins_per_layer) + l) * p.cells_per_layer + nloc) * p.bins_per_layer + e];
        else u = philox_uniform(p.seed, (uint32_t)rb, (uint32_t)c, (uint32_t)nloc);
        return u + p.val[k];
    };
    int best = k0, count = 0;
    float rbest = -1.f;
    for (int k = k0; k < k1; ++k) {                        // entries are stored in increasing cell order: first maximum wins (argmax)
        const float r = draw(k);
        if (r > rbest) { rbest = r; best = k; }
        if (r > 1.f) ++count;
    }
    if (!(rbest > 1.f)) ++count;                           // the forced entry was not selected by itself
    const float share = x / (float)count;
    float* ob = p.out + (long long)b * p.n_rows;
    for (int k = k0; k < k1; ++k) {
        const float r = draw(k);
        if (r > 1.f || k == best) atomicAdd(ob + p.row[k], share);
    }
}

void launch_sparse_decode(const SparseDecParams& p, int B, cudaStream_t st) {
    cudaMemsetAsync(p.out, 0, (size_t)B * p.n_rows * sizeof(float), st);
    long long n = (long long)B * p.n_cols;
    sparse_decode_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p, n);
}

void launch_geom_decode(const FinalParams& p, const GeomMap& dec, const float* grid_pred, int B, cudaStream_t st) {
    long long nvox = (long long)B * dec.n_rows;
    geom_decode_kernel<<<(unsigned)((nvox + 255) / 256), 256, 0, st>>>(p, dec, grid_pred, nvox);
}

// =================================================================================================
// Dense ResNet denoiser of the layer-energy model (LayerDiffusion.layer_model: models/layerdiffusion.py:34-37, 111-114;
// ResNet / ResDense models/models.py:372-457), fp32 throughout, fused with the pre-conditioning output and the sampler
// update (same tail as final_kernel).  One block carries kLmS showers through the whole network: every weight matrix is
// read once per block ([in][out] layout, thread = output column, coalesced) and applied to all kLmS showers held in
// shared memory.  dim_in <= 64 values per shower, hidden <= 256.
// =================================================================================================
constexpr int kLmS = 8;
__device__ __forceinline__ void lm_dense(const float* __restrict__ Wt, const float* __restrict__ bias, const float* in, int in_stride,
                                         float* out, int out_stride, int n_in, int n_out, int act, const float* add, int add_stride,
                                         const float* res, int res_stride) {
    // out[s][o] = act(bias[o] + sum_i in[s][i] * Wt[i][o]) (+ add[s][o]) (+ res[s][o]);  act: 0 none, 1 GELU
    for (int o = threadIdx.x; o < n_out; o += blockDim.x) {
        float acc[kLmS];
        const float bo = bias ? __ldg(bias + o) : 0.f;
#pragma unroll
        for (int s = 0; s < kLmS; ++s) acc[s] = bo;
        for (int i = 0; i < n_in; ++i) {
            const float w = __ldg(Wt + (long long)i * n_out + o);
#pragma unroll
            for (int s = 0; s < kLmS; ++s) acc[s] = fmaf(in[s * in_stride + i], w, acc[s]);
        }
#pragma unroll
        for (int s = 0; s < kLmS; ++s) {
            float v = act ? gelu_erf(acc[s]) : acc[s];
            if (add) v += add[s * add_stride + o];
            if (res) v += res[s * res_stride + o];
            out[s * out_stride + o] = v;
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256) layer_mlp_kernel(LayerMlpParams m, FinalParams p, const float* __restrict__ cond, int B,
                                                        int time_is_raw) {
    __shared__ float xin[kLmS][64];
    __shared__ float cvec[kLmS][128];      // [cond_mlp output | time_mlp output] (models.py:443-445), then GELU of it
    __shared__ float t0[kLmS][64], t1[kLmS][64];
    __shared__ float h[kLmS][256], a[kLmS][256], e[kLmS][256];
    const int b0 = blockIdx.x * kLmS;
    const int D = m.dim_in, H = m.hidden, half = m.cond_dim / 2;
    for (int i = threadIdx.x; i < kLmS * 64; i += blockDim.x) {
        const int s = i >> 6, j = i & 63, b = b0 + s;
        float v = 0.f;
        if (b < B && j < D) {
            const float sg = p.sigma[(long long)b * p.sigma_stride];
            const float c_in = p.sigma_data > 0.f ? rsqrtf(sg * sg + p.sigma_data * p.sigma_data) : 1.f;
            v = p.x[(long long)b * D + j] * c_in;
        }
        xin[s][j] = v;
    }
    if (threadIdx.x < kLmS) {
        const int b = b0 + threadIdx.x;
        float tv = 0.f;
        if (b < B) {
            const float sg = p.sigma[(long long)b * p.sigma_stride];
            tv = time_is_raw ? sg : (m.time_mode == 0 ? sg / sqrtf(1.f + sg * sg) : 0.5f * logf(sg));
        }
        t0[threadIdx.x][0] = tv;
    }
    __syncthreads();
    // time_mlp: Linear(1,h/2) GELU Linear(h/2,h) GELU Linear(h,h) -> cvec[:, half:]
    lm_dense(m.tw[0], m.tb[0], &t0[0][0], 64, &t1[0][0], 64, 1, half / 2, 1, nullptr, 0, nullptr, 0);
    lm_dense(m.tw[1], m.tb[1], &t1[0][0], 64, &t0[0][0], 64, half / 2, half, 1, nullptr, 0, nullptr, 0);
    lm_dense(m.tw[2], m.tb[2], &t0[0][0], 64, &cvec[0][half], 128, half, half, 0, nullptr, 0, nullptr, 0);
    // cond_mlp: Linear(cs,h/2) GELU Linear(h/2,h) GELU Linear(h,h) -> cvec[:, :half]
    for (int i = threadIdx.x; i < kLmS * m.cond_size; i += blockDim.x) {
        const int s = i / m.cond_size, j = i - s * m.cond_size, b = b0 + s;
        t0[s][j] = b < B ? cond[(long long)b * m.cond_size + j] : 0.f;
    }
    __syncthreads();
    lm_dense(m.cw[0], m.cb[0], &t0[0][0], 64, &t1[0][0], 64, m.cond_size, half / 2, 1, nullptr, 0, nullptr, 0);
    lm_dense(m.cw[1], m.cb[1], &t1[0][0], 64, &t0[0][0], 64, half / 2, half, 1, nullptr, 0, nullptr, 0);
    lm_dense(m.cw[2], m.cb[2], &t0[0][0], 64, &cvec[0][0], 128, half, half, 0, nullptr, 0, nullptr, 0);
    // every ResDense embeder starts with GELU(cond) (models.py:377): apply it once
    for (int i = threadIdx.x; i < kLmS * 128; i += blockDim.x) cvec[i >> 7][i & 127] = gelu_erf(cvec[i >> 7][i & 127]);
    __syncthreads();
    lm_dense(m.in_w, m.in_b, &xin[0][0], 64, &h[0][0], 256, D, H, 0, nullptr, 0, nullptr, 0);
    for (int l = 0; l < m.n_hidden; ++l) {
        lm_dense(m.ew[l], m.eb[l], &cvec[0][0], 128, &e[0][0], 256, m.cond_dim, H, 0, nullptr, 0, nullptr, 0);       // embed
        lm_dense(m.w1[l], m.b1[l], &h[0][0], 256, &a[0][0], 256, H, H, 1, &e[0][0], 256, nullptr, 0);                // GELU(dense1) + embed
        lm_dense(m.w2[l], m.b2[l], &a[0][0], 256, &e[0][0], 256, H, H, 1, nullptr, 0, &h[0][0], 256);                // GELU(dense2) + x
        for (int i = threadIdx.x; i < kLmS * H; i += blockDim.x) h[i / H][i % H] = e[i / H][i % H];
        __syncthreads();
    }
    lm_dense(m.out_w, m.out_b, &h[0][0], 256, &a[0][0], 256, H, D, 0, nullptr, 0, nullptr, 0);
    for (int i = threadIdx.x; i < kLmS * D; i += blockDim.x) {
        const int s = i / D, j = i - s * D, b = b0 + s;
        if (b < B) final_tail(p, (long long)b * D + j, b, j, a[s][j]);
    }
}

void launch_layer_mlp(const LayerMlpParams& m, const FinalParams& p, const float* cond, int B, int time_is_raw, cudaStream_t st) {
    layer_mlp_kernel<<<(B + kLmS - 1) / kLmS, 256, 0, st>>>(m, p, cond, B, time_is_raw);
}

// x_out = a*x + b*noise (start scaling  x = start*sigma_0, and the EDM churn pre-noise, sample.py:633,651)
__global__ void axpby_kernel(const float* __restrict__ x, const float* __restrict__ noise, float* __restrict__ out,
                             const float* __restrict__ coef, long long n, int vps, uint64_t seed, int shower_offset,
                             int step, int rng_enabled, const uint32_t* __restrict__ rng_key) {
    long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    if (rng_key) { seed = ((uint64_t)rng_key[1] << 32) | rng_key[0]; shower_offset = (int)rng_key[2]; }
    float o = coef[0] * x[v];
    float kb = coef[1];
    if (kb != 0.f) {
        if (noise) o += kb * noise[v];
        else if (rng_enabled) {
            int b = (int)(v / vps);
            o += kb * philox_normal(seed, (uint32_t)(shower_offset + b), (uint32_t)step, (uint32_t)(v - (long long)b * vps));
        }
    }
    out[v] = o;
}

void launch_axpby(const float* x, const float* noise, float* out, const float* coef, long long n, int vps, unsigned long long seed,
                  int shower_offset, int step, int rng_enabled, cudaStream_t st, const uint32_t* rng_key) {
    axpby_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, noise, out, coef, n, vps, seed, shower_offset, step,
                                                              rng_enabled, rng_key);
}

// plain device copy / clear as kernels: inside the captured sampling graph these replace cudaMemcpyAsync / cudaMemsetAsync
// nodes (a graph holding large copy-engine memcpy nodes crashed cuGraphLaunch when NCCL peer mappings were active)
__global__ void copy16_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, long long n16, const unsigned char* __restrict__ src_tail,
                              unsigned char* __restrict__ dst_tail, int tail) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) dst[i] = src ? src[i] : make_uint4(0u, 0u, 0u, 0u);
    if (blockIdx.x == 0 && (int)threadIdx.x < tail) dst_tail[threadIdx.x] = src_tail ? src_tail[threadIdx.x] : 0;
}
__global__ void copy4_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, long long n4) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) dst[i] = src ? src[i] : 0u;
}
void launch_copy_bytes(const void* src, void* dst, size_t bytes, cudaStream_t st) {
    // src == nullptr clears dst.  16-byte vectors when both pointers allow it; record slots of odd-sized tensors
    // (e.g. B x 46 layer-energy vectors) are only 4-byte aligned and take the word form.
    if ((((uintptr_t)src | (uintptr_t)dst) & 15u) != 0 && (((uintptr_t)src | (uintptr_t)dst | bytes) & 3u) == 0) {
        const long long n4 = (long long)(bytes / 4);
        long long blocks = (n4 + 255) / 256;
        if (blocks > 148 * 16) blocks = 148 * 16;
        if (blocks < 1) blocks = 1;
        copy4_kernel<<<(unsigned)blocks, 256, 0, st>>>((const uint32_t*)src, (uint32_t*)dst, n4);
        return;
    }
    const long long n16 = (long long)(bytes / 16);
    const int tail = (int)(bytes % 16);
    long long blocks = (n16 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    const unsigned char* st_ = src ? (const unsigned char*)src + n16 * 16 : nullptr;
    copy16_kernel<<<(unsigned)blocks, 256, 0, st>>>((const uint4*)src, (uint4*)dst, n16, st_, (unsigned char*)dst + n16 * 16, tail);
}

__global__ void fill_normal_kernel(float* out, long long n, int vps, uint64_t seed, int shower_offset, int step) {
    long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    int b = (int)(v / vps);
    out[v] = philox_normal(seed, (uint32_t)(shower_offset + b), (uint32_t)step, (uint32_t)(v - (long long)b * vps));
}
void launch_fill_normal(float* out, long long n, int vps, unsigned long long seed, int shower_offset, int step, cudaStream_t st) {
    fill_normal_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(out, n, vps, seed, shower_offset, step);
}

}  // namespace cd
