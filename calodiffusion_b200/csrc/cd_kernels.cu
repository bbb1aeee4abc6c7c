// Hand-written CUDA kernels (sm_100a) for the CaloDiffusion denoiser hot path — everything except the
// tcgen05 implicit-GEMM convolution, which lives in cd_conv_mma.cu.
#include "cd_kernels.cuh"

namespace cd {

// =================================================================================================
// prep_input: x (B,Z,P,R) fp32 -> PLCP tensor with 8 channels: [x*c_in, R, Z, (phi), 0...]
// Replaces calodiffusion.py:121-142 add_RZPhi (repeat+cat) and the c_in multiply of :159.
// =================================================================================================
__global__ void prep_input_kernel(Geom g, const float* __restrict__ x, const float* __restrict__ sigma, int sigma_stride,
                                  float sigma_data, const float* __restrict__ r_vals, int rz_input, int phi_input,
                                  __half* __restrict__ dst, long long nrows) {
    long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    RowInfo ri = decode_row(g, row);
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (ri.kind != 2) {
        int p = ri.pp - 1;
        if (p < 0) p += g.P;
        if (p >= g.P) p -= g.P;
        float s = sigma[(long long)ri.b * sigma_stride];
        float c_in = sigma_data > 0.f ? rsqrtf(s * s + sigma_data * sigma_data) : 1.f;
        long long vox = (((long long)ri.b * g.Z + ri.z) * g.P + p) * g.R + ri.rr;
        f[0] = x[vox] * c_in;
        int c = 1;
        if (rz_input) { f[c++] = r_vals[ri.rr]; f[c++] = (float)ri.z / (float)g.Z; }
        if (phi_input) { f[c++] = (g.P > 1) ? (float)p / (float)(g.P - 1) : 0.f; }
    }
    *reinterpret_cast<Half8*>(dst + plcp_offset(g, 0, row)) = pack8(f);
}

void launch_prep_input(const Geom& g, const float* x, const float* sigma, int sigma_stride, float sigma_data,
                       const float* r_vals, int rz_input, int phi_input, __half* dst, int B, cudaStream_t st) {
    long long nrows = (long long)B * g.rows_per_sample;
    int threads = 256;
    long long blocks = (nrows + threads - 1) / threads;
    prep_input_kernel<<<(unsigned)blocks, threads, 0, st>>>(g, x, sigma, sigma_stride, sigma_data, r_vals, rz_input,
                                                            phi_input, dst, nrows);
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
                            Half8 v = *reinterpret_cast<const Half8*>(sb + plcp_offset(gi, k, irow));
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
        *reinterpret_cast<Half8*>(p.dst + plcp_offset(p.gout, cog, row)) = pack8(acc);
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
// GroupNorm pieces (models/models.py:155,162 GroupNorm(8,C); :293,325 GroupNorm(1,C); eps 1e-5).
// Statistics are per-(sample, channel) double sums produced by the conv epilogues (or row_stats);
// gn_finalize folds them per group into a per-(sample, channel) affine  y = x*scale + shift.
// =================================================================================================
__global__ void gn_finalize_kernel(const double* __restrict__ stats, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float2* __restrict__ affine, int B, int C, int groups,
                                   double count_per_channel, float eps) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= B * C) return;
    int b = idx / C, c = idx - b * C;
    int cpg = C / groups;
    int g0 = (c / cpg) * cpg;
    double s = 0.0, q = 0.0;
    for (int j = 0; j < cpg; ++j) {
        const double* d = stats + ((long long)b * C + g0 + j) * 2;
        s += d[0]; q += d[1];
    }
    double n = count_per_channel * cpg;
    double mean = s / n;
    double var = q / n - mean * mean;
    if (var < 0.0) var = 0.0;
    float rstd = (float)(1.0 / sqrt(var + (double)eps));
    float sc = gamma[c] * rstd;
    affine[idx] = make_float2(sc, beta[c] - (float)mean * sc);
}

void launch_gn_finalize(const double* stats, const float* gamma, const float* beta, float2* affine, int B, int C,
                        int groups, double count_per_channel, cudaStream_t st) {
    int n = B * C;
    gn_finalize_kernel<<<(n + 127) / 128, 128, 0, st>>>(stats, gamma, beta, affine, B, C, groups, count_per_channel, 1e-5f);
}

// out = act(x*scale+shift) + addvec[b][c] + residual ; written for every row incl. halos (gather form)
// Covers Block's GN+SiLU (models.py:160-169), the additive conditioning (:194-197), the ResnetBlock
// residual (:200), PreNorm (:327-329) and Residual(to_out GN) (:117,:293).
__global__ void gn_apply_kernel(Geom g, const __half* __restrict__ src, const float2* __restrict__ affine,
                                const float* __restrict__ addvec, int addvec_stride, const __half* __restrict__ residual,
                                const __half* __restrict__ residual2, int res_split,
                                __half* __restrict__ dst, int C, int act, long long nrows) {
    long long row = (long long)blockIdx.x * 32 + threadIdx.x;
    int k = threadIdx.y + blockIdx.y * blockDim.y;
    if (row >= nrows || k >= (C >> 3)) return;
    RowInfo ri = decode_row(g, row);
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (ri.kind != 2) {
        Half8 v = *reinterpret_cast<const Half8*>(src + plcp_offset(g, k, ri.src));
        unpack8(v, f);
        const float2* af = affine + (long long)ri.b * C + k * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float2 a = af[j];
            float y = f[j] * a.x + a.y;
            if (act) y = silu_f(y);
            f[j] = y;
        }
        if (addvec) {
            const float* av = addvec + (long long)ri.b * addvec_stride + k * 8;
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] += av[j];
        }
        if (residual) {
            // the residual may be the channel concatenation of two tensors (torch.cat skip, models.py:741)
            const __half* rb = (k < res_split) ? residual + plcp_offset(g, k, ri.src) : residual2 + plcp_offset(g, k - res_split, ri.src);
            Half8 rv = *reinterpret_cast<const Half8*>(rb);
            float r[8];
            unpack8(rv, r);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] += r[j];
        }
    }
    *reinterpret_cast<Half8*>(dst + plcp_offset(g, k, row)) = pack8(f);
}

void launch_gn_apply(const Geom& g, const __half* src, const float2* affine, const float* addvec, int addvec_stride,
                     const __half* residual, const __half* residual2, int res_split, __half* dst, int C, int act, int B,
                     cudaStream_t st) {
    long long nrows = (long long)B * g.rows_per_sample;
    int chunks = C >> 3;
    int by = chunks < 8 ? chunks : 8;
    dim3 block(32, by);
    dim3 grid((unsigned)((nrows + 31) / 32), (chunks + by - 1) / by);
    gn_apply_kernel<<<grid, block, 0, st>>>(g, src, affine, addvec, addvec_stride, residual, residual2,
                                            residual2 ? res_split : (C >> 3), dst, C, act, nrows);
}

// per-(sample, channel) sum / sumsq over interior rows of an existing tensor (GroupNorm(1,C) PreNorm input)
__global__ void row_stats_kernel(Geom g, const __half* __restrict__ src, double* __restrict__ stats, int C,
                                 int rows_per_block) {
    int b = blockIdx.y;
    int k = threadIdx.y;
    long long r0 = (long long)b * g.rows_per_sample + (long long)blockIdx.x * rows_per_block;
    long long r1 = r0 + rows_per_block;
    long long rend = (long long)(b + 1) * g.rows_per_sample;
    if (r1 > rend) r1 = rend;
    float s[8], q[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] = 0.f; q[j] = 0.f; }
    for (long long row = r0 + threadIdx.x; row < r1; row += 32) {
        RowInfo ri = decode_row(g, row);
        if (ri.kind != 0) continue;
        Half8 v = *reinterpret_cast<const Half8*>(src + plcp_offset(g, k, row));
        float f[8];
        unpack8(v, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] += f[j] * f[j]; }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s[j] += __shfl_xor_sync(0xffffffffu, s[j], o);
            q[j] += __shfl_xor_sync(0xffffffffu, q[j], o);
        }
        if (threadIdx.x == 0) {
            double* d = stats + ((long long)b * C + k * 8 + j) * 2;
            atomicAdd(d, (double)s[j]);
            atomicAdd(d + 1, (double)q[j]);
        }
    }
}

void launch_row_stats(const Geom& g, const __half* src, double* stats, int C, int B, cudaStream_t st) {
    int rows_per_block = 1024;
    dim3 block(32, C >> 3);
    dim3 grid((g.rows_per_sample + rows_per_block - 1) / rows_per_block, B);
    row_stats_kernel<<<grid, block, 0, st>>>(g, src, stats, C, rows_per_block);
}

// =================================================================================================
// Linear attention (models/models.py:301-318; heads=1, dim_head=32).  qkv is a raw 96-channel PLCP
// tensor: chunks 0-3 q, 4-7 k, 8-11 v.
//   k_max[b][d]          = max_n k[d,n]                       (softmax over voxels, stabiliser)
//   S[b][d]              = sum_n exp(k[d,n]-k_max)
//   Ctx[b][d][e]         = sum_n exp(k[d,n]-k_max) v[e,n]
//   out[e,n]             = 32^-0.5 * sum_d Ctx[d][e]/S[d] * softmax_d(q[:,n])[d]
// =================================================================================================
__device__ __forceinline__ unsigned f2ord(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned k) {
    unsigned u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    return __uint_as_float(u);
}

__global__ void attn_kmax_kernel(Geom g, const __half* __restrict__ qkv, unsigned* __restrict__ kmax, int rows_per_block) {
    int b = blockIdx.y;
    int k = threadIdx.y;  // 0..3 -> k chunk
    long long r0 = (long long)b * g.rows_per_sample + (long long)blockIdx.x * rows_per_block;
    long long r1 = r0 + rows_per_block;
    long long rend = (long long)(b + 1) * g.rows_per_sample;
    if (r1 > rend) r1 = rend;
    float m[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) m[j] = -INFINITY;
    for (long long row = r0 + threadIdx.x; row < r1; row += 32) {
        RowInfo ri = decode_row(g, row);
        if (ri.kind != 0) continue;
        Half8 v = *reinterpret_cast<const Half8*>(qkv + plcp_offset(g, 4 + k, row));
        float f[8];
        unpack8(v, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) m[j] = fmaxf(m[j], f[j]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m[j] = fmaxf(m[j], __shfl_xor_sync(0xffffffffu, m[j], o));
        if (threadIdx.x == 0 && m[j] > -INFINITY) atomicMax(kmax + b * 32 + k * 8 + j, f2ord(m[j]));
    }
}

// ctx layout: [B][33][32] floats: rows 0..31 = Ctx[d][e], row 32 = S[d]
__global__ void __launch_bounds__(256) attn_context_kernel(Geom g, const __half* __restrict__ qkv,
                                                           const unsigned* __restrict__ kmax, float* __restrict__ ctx,
                                                           int rows_per_block, int use_atomics) {
    __shared__ float ek[128][33];
    __shared__ float vv[128][33];
    __shared__ float km[32];
    int b = blockIdx.y;
    int tid = threadIdx.x;
    if (tid < 32) km[tid] = ord2f(kmax[b * 32 + tid]);
    long long r0 = (long long)b * g.rows_per_sample + (long long)blockIdx.x * rows_per_block;
    long long r1 = r0 + rows_per_block;
    long long rend = (long long)(b + 1) * g.rows_per_sample;
    if (r1 > rend) r1 = rend;
    int d = tid >> 3;          // 0..31
    int e0 = (tid & 7) * 4;    // 4 consecutive e
    float acc[4] = {0, 0, 0, 0};
    float sacc = 0.f;
    __syncthreads();
    for (long long base = r0; base < r1; base += 128) {
        // stage 128 rows: thread -> (row = tid&127, half = tid>>7 : 0 -> k, 1 -> v)
        {
            int lr = tid & 127;
            int which = tid >> 7;
            long long row = base + lr;
            bool ok = row < r1;
            if (ok) { RowInfo ri = decode_row(g, row); ok = (ri.kind == 0); }
            float (*dstm)[33] = which ? vv : ek;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                if (ok) {
                    Half8 v = *reinterpret_cast<const Half8*>(qkv + plcp_offset(g, (which ? 8 : 4) + k, row));
                    unpack8(v, f);
                    if (!which) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) f[j] = __expf(f[j] - km[k * 8 + j]);
                    }
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) dstm[lr][k * 8 + j] = f[j];
            }
        }
        __syncthreads();
#pragma unroll 4
        for (int n = 0; n < 128; ++n) {
            float e = ek[n][d];
            acc[0] += e * vv[n][e0]; acc[1] += e * vv[n][e0 + 1]; acc[2] += e * vv[n][e0 + 2]; acc[3] += e * vv[n][e0 + 3];
            sacc += e;
        }
        __syncthreads();
    }
    float* c = ctx + (long long)b * 33 * 32;
    if (use_atomics) {
#pragma unroll
        for (int i = 0; i < 4; ++i) atomicAdd(c + d * 32 + e0 + i, acc[i]);
        if ((tid & 7) == 0) atomicAdd(c + 32 * 32 + d, sacc);
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) c[d * 32 + e0 + i] = acc[i];
        if ((tid & 7) == 0) c[32 * 32 + d] = sacc;
    }
}

__global__ void __launch_bounds__(128) attn_out_kernel(Geom g, const __half* __restrict__ qkv, const float* __restrict__ ctx,
                                                       __half* __restrict__ dst, int rows_per_block) {
    __shared__ float M[32][33];  // M[d][e] = Ctx[d][e]/S[d]*scale
    int b = blockIdx.y;
    const float* c = ctx + (long long)b * 33 * 32;
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) {
        int d = i >> 5, e = i & 31;
        M[d][e] = c[i] / c[1024 + d] * 0.17677669529663687f;  // 32^-0.5
    }
    __syncthreads();
    long long r0 = (long long)b * g.rows_per_sample + (long long)blockIdx.x * rows_per_block;
    long long r1 = r0 + rows_per_block;
    long long rend = (long long)(b + 1) * g.rows_per_sample;
    if (r1 > rend) r1 = rend;
    for (long long row = r0 + threadIdx.x; row < r1; row += blockDim.x) {
        RowInfo ri = decode_row(g, row);
        if (ri.kind != 0) continue;
        float q[32];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            Half8 v = *reinterpret_cast<const Half8*>(qkv + plcp_offset(g, k, row));
            unpack8(v, q + k * 8);
        }
        float mx = q[0];
#pragma unroll
        for (int j = 1; j < 32; ++j) mx = fmaxf(mx, q[j]);
        float sum = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) { q[j] = __expf(q[j] - mx); sum += q[j]; }
        float inv = 1.0f / sum;
        float o[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) o[e] = 0.f;
#pragma unroll
        for (int d = 0; d < 32; ++d) {
            float qd = q[d] * inv;
#pragma unroll
            for (int e = 0; e < 32; ++e) o[e] += qd * M[d][e];
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
            *reinterpret_cast<Half8*>(dst + plcp_offset(g, k, row)) = pack8(o + k * 8);
    }
}

void launch_attention(const Geom& g, const __half* qkv, unsigned* kmax, float* ctx, __half* dst, int B, cudaStream_t st) {
    // kmax and ctx must be zero on entry (the engine clears its scratch once per forward)
    int rpb = 2048;
    {
        dim3 block(32, 4), grid((g.rows_per_sample + rpb - 1) / rpb, B);
        attn_kmax_kernel<<<grid, block, 0, st>>>(g, qkv, kmax, rpb);
    }
    {
        int nsplit = 1;
        while ((long long)B * nsplit < 296 && (g.rows_per_sample / (nsplit * 2)) >= 512) nsplit *= 2;
        int rows = (g.rows_per_sample + nsplit - 1) / nsplit;
        rows = (rows + 127) / 128 * 128;
        int nblk = (g.rows_per_sample + rows - 1) / rows;
        dim3 grid(nblk, B);
        attn_context_kernel<<<grid, 256, 0, st>>>(g, qkv, kmax, ctx, rows, nblk > 1 ? 1 : 0);
    }
    {
        int rows = 512;
        dim3 grid((g.rows_per_sample + rows - 1) / rows, B);
        attn_out_kernel<<<grid, 128, 0, st>>>(g, qkv, ctx, dst, rows);
    }
}

// =================================================================================================
// Conditioning: time MLP + cond MLP (models/models.py:575-608, :704-707) and every ResnetBlock's
// SiLU -> Linear(cond_dim, C) projection (:176-178, :195), one block per sample.
// time input = do_time_embed(sigma) (calodiffusion.py:144-155).
// =================================================================================================
__device__ void dense_layer(const float* __restrict__ W, const float* __restrict__ bias, const float* in, float* out,
                            int n_in, int n_out, int act) {
    for (int o = threadIdx.x; o < n_out; o += blockDim.x) {
        float a = bias ? bias[o] : 0.f;
        const float* w = W + (long long)o * n_in;
        for (int i = 0; i < n_in; ++i) a += w[i] * in[i];
        if (act == 1) a = gelu_erf(a);
        out[o] = a;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(128) embed_kernel(EmbedParams p, const float* __restrict__ sigma, int sigma_stride,
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
    for (int k = 0; k < p.n_proj; ++k) {
        const float* W = p.pw[k];
        const float* bias = p.pb[k];
        int C = p.pC[k];
        float* out = addvec + (long long)b * p.addvec_stride + p.poff[k];
        for (int o = threadIdx.x; o < C; o += blockDim.x) {
            float a = bias[o];
            const float* w = W + (long long)o * p.cond_dim;
            for (int i = 0; i < p.cond_dim; ++i) a += w[i] * cvec[i];
            out[o] = a;
        }
    }
}

void launch_embed(const EmbedParams& p, const float* sigma, int sigma_stride, const float* cond, float* addvec, int B,
                  cudaStream_t st) {
    embed_kernel<<<B, 128, 0, st>>>(p, sigma, sigma_stride, cond, addvec);
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
    float pred = p.bias;
    int nch = p.C >> 3;
    for (int k = 0; k < nch; ++k) {
        Half8 hv = *reinterpret_cast<const Half8*>(p.h + plcp_offset(g, k, row));
        float f[8];
        unpack8(hv, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) pred += f[j] * p.w[k * 8 + j];
    }
    float s = p.sigma[(long long)b * p.sigma_stride];
    float sd = p.sigma_data;
    float den = s * s + sd * sd;
    float xin = p.x[v];
    float x0 = sd > 0.f ? (sd * sd / den) * xin + (s * sd * rsqrtf(den)) * pred : pred;
    if (p.pred_out) p.pred_out[v] = pred;
    if (p.x0_out) p.x0_out[v] = x0;
    if (p.x_out) {
        const float* k = p.coef;
        float o = k[0] * x0 + k[1] * xin;
        if (p.a1) o += k[2] * p.a1[v];
        else if (p.rng_enabled && k[2] != 0.f)
            o += k[2] * philox_normal(p.seed, (uint32_t)(p.shower_offset + b), (uint32_t)p.step, (uint32_t)rem);
        if (p.a2) o += k[3] * p.a2[v];
        if (p.a3) o += k[4] * p.a3[v];
        p.x_out[v] = o;
    }
}

void launch_final(const FinalParams& p, int B, cudaStream_t st) {
    long long nvox = (long long)B * p.g.Z * p.g.P * p.g.R;
    final_kernel<<<(unsigned)((nvox + 255) / 256), 256, 0, st>>>(p, nvox);
}

// x_out = a*x + b*noise (start scaling  x = start*sigma_0, and the EDM churn pre-noise, sample.py:633,651)
__global__ void axpby_kernel(const float* __restrict__ x, const float* __restrict__ noise, float* __restrict__ out,
                             const float* __restrict__ coef, long long n, int vps, uint64_t seed, int shower_offset,
                             int step, int rng_enabled) {
    long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
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
                  int shower_offset, int step, int rng_enabled, cudaStream_t st) {
    axpby_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, noise, out, coef, n, vps, seed, shower_offset, step,
                                                              rng_enabled);
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
