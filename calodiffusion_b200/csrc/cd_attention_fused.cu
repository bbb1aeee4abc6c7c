// PreNorm + to_qkv + linear attention + to_out, without ever writing the 96-channel qkv tensor
// (models/models.py:301-329: Residual(PreNorm(dim, LinearAttention(dim))), heads = 1, dim_head = 32).
//
// The unfused form (cd_conv_mma.cu folded-PreNorm 1x1x1 conv + cd_attention.cu) moves 9 activation-sized streams per
// block (x in, qkv out = 3, k twice, v, q, y out); here the q/k/v projections are recomputed from x inside each of the
// three passes, so the block moves 4 (x three times, y out):
//   fa_kmax : K^T = Wk' x^T per 32-row group (mma.sync m16n8k16, fp16 x fp16 -> fp32), running max_n per channel
//   fa_ctx  : K^T, V^T again; e = exp(rstd*(K^T - max));  Ctx[d][e] += e V (the accumulators of the projection ARE the
//             A / B fragments of the second MMA: no shuffles, no shared memory), S[d] += sum_n e
//   fa_out  : Q = x Wq'^T, softmax over d inside the lane quad, y = softmax(q) (Wout (Ctx/S)^T 32^-1/2)^T + bias as one
//             more MMA, fp16 PLCP store + GroupNorm(1,C) statistics
// PreNorm's GroupNorm(1,C) is folded as in the conv:  W(gamma*(x-mean)*rstd+beta) = rstd*(W')x + (W beta - mean*rstd*rowsum(W')),
// W' = W diag(gamma) in fp16.  The x fragments are loaded straight from global memory: in the PLCP layout one 8-row x
// 8-channel fragment is 128 contiguous bytes, so every warp-level load is one fully used 128-byte line.
#include <cstdlib>
#include "cd_kernels.cuh"

namespace cd {
namespace {

__device__ __forceinline__ unsigned f2ord(float f) {
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned k) {
    unsigned u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    return __uint_as_float(u);
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t* r) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint32_t ldg_u32(const __half* p) { return __ldg(reinterpret_cast<const unsigned int*>(p)); }

__device__ __forceinline__ float ex2f(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// interior-row mask of the 32-row group starting at `base` (bit = row base+lane is an interior voxel of this block's range)
__device__ __forceinline__ unsigned group_mask(const Geom& g, uint32_t base, uint32_t r1, int lane) {
    const uint32_t rem = base + lane;
    return __ballot_sync(0xffffffffu, rem < r1 && decode_subrow(g, rem).kind == 0);
}
// half2 lane mask for the row pair (2t, 2t+1) of row tile j
__device__ __forceinline__ uint32_t pair_mask(unsigned vmask, int j, int tq) {
    const unsigned v = vmask >> (8 * j + 2 * tq);
    return ((v & 1u) ? 0x0000FFFFu : 0u) | ((v & 2u) ? 0xFFFF0000u : 0u);
}

constexpr int kWarps = 8;
constexpr float kLog2e = 1.4426950408889634f;

// per-sample scalars of the folded PreNorm
struct PnScal { float a, m; };   // y = a*acc + (wb + m*wsum),  a = rstd, m = -mean*rstd
__device__ __forceinline__ PnScal pn_scalars(const double* pn_stats, int b, int C, float inv_count) {
    const double* st = pn_stats + (long long)b * C * 2;
    const double mean = st[0] * (double)inv_count;
    double var = st[1] * (double)inv_count - mean * mean;
    if (var < 0.0) var = 0.0;
    PnScal s;
    s.a = rsqrtf((float)var + 1e-5f);
    s.m = -(float)mean * s.a;
    return s;
}

// 32 rows x (16*KS) channels of x as mma fragments: f[j][s][0|1] = x[row base+8j+g][channels 16s+2t,+1 | 16s+8+2t,+1].
// As the col-major B operand (N = 8 rows of tile j) of W x^T, and as halves of the row-major A operand (M = rows) of x W^T.
template <int KS>
__device__ __forceinline__ void load_x32(const __half* __restrict__ xb, long long cs, uint32_t base, uint32_t r1, int lane,
                                         uint32_t (&f)[4][KS][2]) {
    const int gq = lane >> 2, tq = lane & 3;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t rem = base + 8 * j + gq;
        const bool ok = rem < r1;
        const __half* p = xb + (long long)rem * 8 + 2 * tq;
#pragma unroll
        for (int s = 0; s < KS; ++s) {
            f[j][s][0] = ok ? ldg_u32(p + (2 * s) * cs) : 0u;
            f[j][s][1] = ok ? ldg_u32(p + (2 * s + 1) * cs) : 0u;
        }
    }
}

// stage `nrows` rows of the folded fp16 weights [.][C] into shared memory with a conflict-free row stride
template <int KS>
__device__ __forceinline__ void stage_w(const __half* __restrict__ w, unsigned char* sm, int nrows) {
    constexpr int C = 16 * KS, STR = C * 2 + 16;
    for (int i = threadIdx.x; i < nrows * (C / 8); i += blockDim.x) {
        const int r = i / (C / 8), c8 = i - r * (C / 8);
        *reinterpret_cast<uint4*>(sm + r * STR + c8 * 16) = __ldg(reinterpret_cast<const uint4*>(w + (long long)r * C) + c8);
    }
}
// A fragments (16 rows x 16 k) of a staged row-major matrix: rows m0.., k columns 16s..
template <int KS>
__device__ __forceinline__ void ldsm_a(const unsigned char* sm, int m0, int s, int lane, uint32_t* a) {
    constexpr int STR = 16 * KS * 2 + 16;
    ldsm_x4(smem_u32(sm) + (m0 + (lane & 7) + 8 * ((lane >> 3) & 1)) * STR + s * 32 + (lane >> 4) * 16, a);
}

// ---------------------------------------------------------------------------------------------------------
// pass 1: raw (pre-affine) maximum over the interior voxels of K^T = Wk' x^T, per (sample, channel d).
// rstd > 0, so  max_n k[d,n] = rstd * rawmax[d] + const[d]  and  k - max = rstd * (acc - rawmax): pass 2 needs nothing else.
template <int KS>
__global__ void __launch_bounds__(256, KS <= 2 ? 4 : 3) fa_kmax_kernel(Geom g, const __half* __restrict__ x, const __half* __restrict__ wf,
                                                         unsigned* __restrict__ kmax, int rows_per_block) {
    constexpr int C = 16 * KS, STR = C * 2 + 16;
    __shared__ __align__(16) unsigned char sw[32 * STR];
    __shared__ unsigned smax[32];
    const int b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gq = lane >> 2, tq = lane & 3;
    stage_w<KS>(wf + 32 * C, sw, 32);                 // rows 32..63 of [96][C]: the k projection
    if (threadIdx.x < 32) smax[threadIdx.x] = 0u;
    __syncthreads();
    uint32_t wk[2][KS][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int s = 0; s < KS; ++s) ldsm_a<KS>(sw, 16 * i, s, lane, wk[i][s]);
    const uint32_t r0 = blockIdx.x * rows_per_block;
    uint32_t r1 = r0 + rows_per_block;
    if (r1 > (uint32_t)g.rows_per_sample) r1 = g.rows_per_sample;
    const __half* xb = x + plcp_offset(g, 0, (long long)b * g.rows_per_sample);
    const long long cs = g.chunk_stride * 8;
    float mx[2][2] = {{-INFINITY, -INFINITY}, {-INFINITY, -INFINITY}};   // [m tile][row g | g+8]
    // light on registers on purpose: 4 blocks (32 warps) per SM keep enough loads in flight without software pipelining
    for (uint32_t base = r0 + warp * 32; base < r1; base += kWarps * 32) {
        const unsigned cm = group_mask(g, base, r1, lane);
        if (!cm) continue;
        uint32_t xf[4][KS][2];
        load_x32<KS>(xb, cs, base, r1, lane, xf);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const bool v0 = (cm >> (8 * j + 2 * tq)) & 1u, v1 = (cm >> (8 * j + 2 * tq + 1)) & 1u;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                float c4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int s = 0; s < KS; ++s) mma16816(c4, wk[i][s], xf[j][s][0], xf[j][s][1]);
                if (v0) { mx[i][0] = fmaxf(mx[i][0], c4[0]); mx[i][1] = fmaxf(mx[i][1], c4[2]); }
                if (v1) { mx[i][0] = fmaxf(mx[i][0], c4[1]); mx[i][1] = fmaxf(mx[i][1], c4[3]); }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float m = mx[i][h];
            m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
            m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
            if (tq == 0 && m > -INFINITY) atomicMax(&smax[16 * i + 8 * h + gq], f2ord(m));
        }
    __syncthreads();
    if (threadIdx.x < 32 && smax[threadIdx.x]) atomicMax(kmax + b * 32 + threadIdx.x, smax[threadIdx.x]);
}

// ---------------------------------------------------------------------------------------------------------
// pass 2: Ctx[d][e] = sum_n exp(k[d,n]-kmax[d]) v[e,n],  S[d] = sum_n exp(k[d,n]-kmax[d])   (ctx layout [B][33][32], row 32 = S)
template <int KS>
__global__ void __launch_bounds__(256, 2) fa_ctx_kernel(Geom g, const __half* __restrict__ x, const __half* __restrict__ wf,
                                                        const float* __restrict__ wsum, const float* __restrict__ wb,
                                                        const double* __restrict__ pn_stats, float pn_inv_count,
                                                        const unsigned* __restrict__ kmax, float* __restrict__ ctx,
                                                        int rows_per_block, int use_atomics) {
    constexpr int C = 16 * KS, STR = C * 2 + 16;
    __shared__ __align__(16) unsigned char sw[64 * STR];     // rows 0..31 Wk', 32..63 Wv'
    __shared__ float red[33 * 32];
    const int b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gq = lane >> 2, tq = lane & 3;
    stage_w<KS>(wf + 32 * C, sw, 64);
    for (int i = threadIdx.x; i < 33 * 32; i += 256) red[i] = 0.f;
    __syncthreads();
    const PnScal ps = pn_scalars(pn_stats, b, C, pn_inv_count);
    // this lane's channels: d (and e) = 16i + 8h + g
    float km[2][2], cv[2][2];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int d = 16 * i + 8 * h + gq;
            km[i][h] = ord2f(kmax[b * 32 + d]);
            cv[i][h] = fmaf(ps.m, __ldg(wsum + 64 + d), __ldg(wb + 64 + d));
        }
    const uint32_t r0 = blockIdx.x * rows_per_block;
    uint32_t r1 = r0 + rows_per_block;
    if (r1 > (uint32_t)g.rows_per_sample) r1 = g.rows_per_sample;
    const __half* xb = x + plcp_offset(g, 0, (long long)b * g.rows_per_sample);
    const long long cs = g.chunk_stride * 8;
    float acc[2][4][4];     // Ctx: [m tile d][n tile e][frag]
    float sacc[2][4];       // S: column 0 of the same MMA against a ones column (identically rounded weights)
#pragma unroll
    for (int i = 0; i < 2; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int f = 0; f < 4; ++f) acc[i][j][f] = 0.f;
#pragma unroll
        for (int f = 0; f < 4; ++f) sacc[i][f] = 0.f;
    }
    const uint32_t ones = gq == 0 ? 0x3C003C00u : 0u;
    // exp(rstd*(acc - rawmax)) = 2^(a2*acc + nk)
    const float a2 = ps.a * kLog2e;
    float nk[2][2];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int h = 0; h < 2; ++h) nk[i][h] = -a2 * km[i][h];
    uint32_t xf[4][KS][2];
    uint32_t base = r0 + warp * 32;
    unsigned vmask = 0;
    if (base < r1) {
        vmask = group_mask(g, base, r1, lane);
        if (vmask) load_x32<KS>(xb, cs, base, r1, lane, xf);
    }
    while (base < r1) {
        const unsigned cm = vmask;
        uint32_t ek[2][4][2];     // [m tile d][row tile j]: packed (c0,c1) row g, (c2,c3) row g+8
        uint32_t vt[2][4][2];     // [m tile e][row tile j]
        if (cm) {
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                uint32_t a[KS][4];
#pragma unroll
                for (int s = 0; s < KS; ++s) ldsm_a<KS>(sw, 16 * i, s, lane, a[s]);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float c4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int s = 0; s < KS; ++s) mma16816(c4, a[s], xf[j][s][0], xf[j][s][1]);
                    const uint32_t pm = pair_mask(cm, j, tq);
                    ek[i][j][0] = pack_h2(ex2f(fmaf(a2, c4[0], nk[i][0])), ex2f(fmaf(a2, c4[1], nk[i][0]))) & pm;
                    ek[i][j][1] = pack_h2(ex2f(fmaf(a2, c4[2], nk[i][1])), ex2f(fmaf(a2, c4[3], nk[i][1]))) & pm;
                }
#pragma unroll
                for (int s = 0; s < KS; ++s) ldsm_a<KS>(sw, 32 + 16 * i, s, lane, a[s]);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float c4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int s = 0; s < KS; ++s) mma16816(c4, a[s], xf[j][s][0], xf[j][s][1]);
                    vt[i][j][0] = pack_h2(fmaf(ps.a, c4[0], cv[i][0]), fmaf(ps.a, c4[1], cv[i][0]));
                    vt[i][j][1] = pack_h2(fmaf(ps.a, c4[2], cv[i][1]), fmaf(ps.a, c4[3], cv[i][1]));
                }
            }
        }
        // x of the next group goes in flight now (its registers are free), under the second contraction
        base += kWarps * 32;
        if (base < r1) {
            vmask = group_mask(g, base, r1, lane);
            if (vmask) load_x32<KS>(xb, cs, base, r1, lane, xf);
        }
        if (cm) {
            // Ctx += eK (d x 32 rows) * V (32 rows x e): two k16 steps over the rows; fragments are the packed accumulators
#pragma unroll
            for (int ks = 0; ks < 2; ++ks)
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const uint32_t a[4] = {ek[i][2 * ks][0], ek[i][2 * ks][1], ek[i][2 * ks + 1][0], ek[i][2 * ks + 1][1]};
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj)      // e tile jj: rows 8jj..8jj+7 of V^T = m tile jj/2, half jj%2
                        mma16816(acc[i][jj], a, vt[jj >> 1][2 * ks][jj & 1], vt[jj >> 1][2 * ks + 1][jj & 1]);
                    mma16816(sacc[i], a, ones, ones);
                }
        }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int d0 = i * 16 + gq, e0 = j * 8 + 2 * tq;
            atomicAdd(&red[d0 * 32 + e0], acc[i][j][0]);
            atomicAdd(&red[d0 * 32 + e0 + 1], acc[i][j][1]);
            atomicAdd(&red[(d0 + 8) * 32 + e0], acc[i][j][2]);
            atomicAdd(&red[(d0 + 8) * 32 + e0 + 1], acc[i][j][3]);
        }
        if (tq == 0) {
            atomicAdd(&red[32 * 32 + i * 16 + gq], sacc[i][0]);
            atomicAdd(&red[32 * 32 + i * 16 + gq + 8], sacc[i][2]);
        }
    }
    __syncthreads();
    float* c = ctx + (long long)b * 33 * 32;
    for (int i = threadIdx.x; i < 33 * 32; i += 256) {
        if (use_atomics) atomicAdd(c + i, red[i]);
        else c[i] = red[i];
    }
}

// ---------------------------------------------------------------------------------------------------------
// pass 1+2 in one ("online" softmax over the voxels): every warp keeps a running raw maximum per channel and rescales its
// Ctx / S accumulators when it grows, so the separate max pass (one more read of x) disappears.  Each block writes its
// partial (Ctx, S) together with the maximum they are relative to: part[b][blk] = {33 x 32 floats, 32 maxima};
// fa_fold_online merges the blocks.  No atomics on global memory: the result is independent of block scheduling.
constexpr int kPartFloats = 33 * 32 + 32;
template <int KS>
__global__ void __launch_bounds__(256, 2) fa_ctx_online_kernel(Geom g, const __half* __restrict__ x, const __half* __restrict__ wf,
                                                               const float* __restrict__ wsum, const float* __restrict__ wb,
                                                               const double* __restrict__ pn_stats, float pn_inv_count,
                                                               float* __restrict__ part, int rows_per_block) {
    constexpr int C = 16 * KS, STR = C * 2 + 16;
    __shared__ __align__(16) unsigned char sw[64 * STR];     // rows 0..31 Wk', 32..63 Wv'
    __shared__ float red[33 * 32];
    __shared__ float wmax[kWarps][32];
    const int b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gq = lane >> 2, tq = lane & 3;
    stage_w<KS>(wf + 32 * C, sw, 64);
    for (int i = threadIdx.x; i < 33 * 32; i += 256) red[i] = 0.f;
    __syncthreads();
    const PnScal ps = pn_scalars(pn_stats, b, C, pn_inv_count);
    const float a2 = ps.a * kLog2e;
    float cv[2][2], mr[2][2], nk[2][2];     // this lane's channels: d (and e) = 16i + 8h + g
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int d = 16 * i + 8 * h + gq;
            cv[i][h] = fmaf(ps.m, __ldg(wsum + 64 + d), __ldg(wb + 64 + d));
            mr[i][h] = -INFINITY;
            nk[i][h] = 0.f;
        }
    const uint32_t r0 = blockIdx.x * rows_per_block;
    uint32_t r1 = r0 + rows_per_block;
    if (r1 > (uint32_t)g.rows_per_sample) r1 = g.rows_per_sample;
    const __half* xb = x + plcp_offset(g, 0, (long long)b * g.rows_per_sample);
    const long long cs = g.chunk_stride * 8;
    float acc[2][4][4], sacc[2][4];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int f = 0; f < 4; ++f) acc[i][j][f] = 0.f;
#pragma unroll
        for (int f = 0; f < 4; ++f) sacc[i][f] = 0.f;
    }
    const uint32_t ones = gq == 0 ? 0x3C003C00u : 0u;
    uint32_t xf[4][KS][2];
    uint32_t base = r0 + warp * 32;
    unsigned vmask = 0;
    if (base < r1) {
        vmask = group_mask(g, base, r1, lane);
        if (vmask) load_x32<KS>(xb, cs, base, r1, lane, xf);
    }
    while (base < r1) {
        const unsigned cm = vmask;
        uint32_t ek[2][4][2], vt[2][4][2];
        if (cm) {
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                uint32_t a[KS][4];
#pragma unroll
                for (int s = 0; s < KS; ++s) ldsm_a<KS>(sw, 16 * i, s, lane, a[s]);
                float kr[4][4];
                float gm0 = -INFINITY, gm1 = -INFINITY;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    kr[j][0] = kr[j][1] = kr[j][2] = kr[j][3] = 0.f;
#pragma unroll
                    for (int s = 0; s < KS; ++s) mma16816(kr[j], a[s], xf[j][s][0], xf[j][s][1]);
                    const bool v0 = (cm >> (8 * j + 2 * tq)) & 1u, v1 = (cm >> (8 * j + 2 * tq + 1)) & 1u;
                    if (v0) { gm0 = fmaxf(gm0, kr[j][0]); gm1 = fmaxf(gm1, kr[j][2]); }
                    if (v1) { gm0 = fmaxf(gm0, kr[j][1]); gm1 = fmaxf(gm1, kr[j][3]); }
                }
                gm0 = fmaxf(gm0, __shfl_xor_sync(0xffffffffu, gm0, 1)); gm0 = fmaxf(gm0, __shfl_xor_sync(0xffffffffu, gm0, 2));
                gm1 = fmaxf(gm1, __shfl_xor_sync(0xffffffffu, gm1, 1)); gm1 = fmaxf(gm1, __shfl_xor_sync(0xffffffffu, gm1, 2));
                if (__any_sync(0xffffffffu, gm0 > mr[i][0] || gm1 > mr[i][1])) {
                    // the maximum grew for some channel: bring this m tile's accumulators to the new reference
                    const float n0 = fmaxf(mr[i][0], gm0), n1 = fmaxf(mr[i][1], gm1);
                    const float f0 = mr[i][0] == -INFINITY ? 0.f : ex2f(a2 * (mr[i][0] - n0));
                    const float f1 = mr[i][1] == -INFINITY ? 0.f : ex2f(a2 * (mr[i][1] - n1));
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) { acc[i][jj][0] *= f0; acc[i][jj][1] *= f0; acc[i][jj][2] *= f1; acc[i][jj][3] *= f1; }
                    sacc[i][0] *= f0; sacc[i][1] *= f0; sacc[i][2] *= f1; sacc[i][3] *= f1;
                    mr[i][0] = n0; mr[i][1] = n1;
                    nk[i][0] = -a2 * n0; nk[i][1] = -a2 * n1;
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t pm = pair_mask(cm, j, tq);
                    ek[i][j][0] = pack_h2(ex2f(fmaf(a2, kr[j][0], nk[i][0])), ex2f(fmaf(a2, kr[j][1], nk[i][0]))) & pm;
                    ek[i][j][1] = pack_h2(ex2f(fmaf(a2, kr[j][2], nk[i][1])), ex2f(fmaf(a2, kr[j][3], nk[i][1]))) & pm;
                }
#pragma unroll
                for (int s = 0; s < KS; ++s) ldsm_a<KS>(sw, 32 + 16 * i, s, lane, a[s]);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float c4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int s = 0; s < KS; ++s) mma16816(c4, a[s], xf[j][s][0], xf[j][s][1]);
                    vt[i][j][0] = pack_h2(fmaf(ps.a, c4[0], cv[i][0]), fmaf(ps.a, c4[1], cv[i][0]));
                    vt[i][j][1] = pack_h2(fmaf(ps.a, c4[2], cv[i][1]), fmaf(ps.a, c4[3], cv[i][1]));
                }
            }
        }
        base += kWarps * 32;
        if (base < r1) {
            vmask = group_mask(g, base, r1, lane);
            if (vmask) load_x32<KS>(xb, cs, base, r1, lane, xf);
        }
        if (cm) {
#pragma unroll
            for (int ks = 0; ks < 2; ++ks)
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const uint32_t a[4] = {ek[i][2 * ks][0], ek[i][2 * ks][1], ek[i][2 * ks + 1][0], ek[i][2 * ks + 1][1]};
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj)
                        mma16816(acc[i][jj], a, vt[jj >> 1][2 * ks][jj & 1], vt[jj >> 1][2 * ks + 1][jj & 1]);
                    mma16816(sacc[i], a, ones, ones);
                }
        }
    }
    // block maximum per channel, then every warp rescales to it and adds into the block partial
    if (tq == 0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) { wmax[warp][16 * i + gq] = mr[i][0]; wmax[warp][16 * i + 8 + gq] = mr[i][1]; }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        float f[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int d = 16 * i + 8 * h + gq;
            float bm = wmax[0][d];
#pragma unroll
            for (int w = 1; w < kWarps; ++w) bm = fmaxf(bm, wmax[w][d]);
            f[h] = mr[i][h] == -INFINITY ? 0.f : ex2f(a2 * (mr[i][h] - bm));
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int d0 = i * 16 + gq, e0 = j * 8 + 2 * tq;
            atomicAdd(&red[d0 * 32 + e0], acc[i][j][0] * f[0]);
            atomicAdd(&red[d0 * 32 + e0 + 1], acc[i][j][1] * f[0]);
            atomicAdd(&red[(d0 + 8) * 32 + e0], acc[i][j][2] * f[1]);
            atomicAdd(&red[(d0 + 8) * 32 + e0 + 1], acc[i][j][3] * f[1]);
        }
        if (tq == 0) {
            atomicAdd(&red[32 * 32 + i * 16 + gq], sacc[i][0] * f[0]);
            atomicAdd(&red[32 * 32 + i * 16 + gq + 8], sacc[i][2] * f[1]);
        }
    }
    __syncthreads();
    float* o = part + ((long long)b * gridDim.x + blockIdx.x) * kPartFloats;
    for (int i = threadIdx.x; i < 33 * 32; i += 256) o[i] = red[i];
    if (threadIdx.x < 32) {
        float bm = wmax[0][threadIdx.x];
#pragma unroll
        for (int w = 1; w < kWarps; ++w) bm = fmaxf(bm, wmax[w][threadIdx.x]);
        o[33 * 32 + threadIdx.x] = bm;
    }
}

// merge the block partials of one sample and fold:  fold[c][d] = sum_e Wout[c][e] * Ctx[d][e] / S[d] * 32^-1/2
__global__ void __launch_bounds__(256) fa_fold_online_kernel(const float* __restrict__ part, int nblk, const double* __restrict__ pn_stats,
                                                             float pn_inv_count, int C, const float* __restrict__ wout /* [Cout][32] */,
                                                             __half* __restrict__ fold, int CO) {
    __shared__ float M[32][33];
    __shared__ float wgt[32][32];      // [blk][d] weight of block blk for channel d (nblk <= 32)
    __shared__ float S[32];
    const int b = blockIdx.x;
    const float a2 = pn_scalars(pn_stats, b, C, pn_inv_count).a * kLog2e;
    const float* pb = part + (long long)b * nblk * kPartFloats;
    if (threadIdx.x < 32) {
        const int d = threadIdx.x;
        float gm = -INFINITY;
        for (int k = 0; k < nblk; ++k) gm = fmaxf(gm, pb[(long long)k * kPartFloats + 33 * 32 + d]);
        float s = 0.f;
        for (int k = 0; k < nblk; ++k) {
            const float mk = pb[(long long)k * kPartFloats + 33 * 32 + d];
            const float w = mk == -INFINITY ? 0.f : ex2f(a2 * (mk - gm));
            wgt[k][d] = w;
            s = fmaf(w, pb[(long long)k * kPartFloats + 32 * 32 + d], s);
        }
        S[d] = s;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 1024; i += 256) {
        const int d = i >> 5, e = i & 31;
        float c = 0.f;
        for (int k = 0; k < nblk; ++k) c = fmaf(wgt[k][d], pb[(long long)k * kPartFloats + i], c);
        M[d][e] = c / S[d] * 0.17677669529663687f;  // 32^-0.5
    }
    __syncthreads();
    for (int i = threadIdx.x; i < CO * 32; i += 256) {
        const int co = i >> 5, d = i & 31;
        float a = 0.f;
#pragma unroll 8
        for (int e = 0; e < 32; ++e) a = fmaf(__ldg(wout + co * 32 + e), M[d][e], a);
        fold[(long long)b * CO * 32 + i] = __float2half_rn(a);
    }
}

// ---------------------------------------------------------------------------------------------------------
// once per sample: fold[c][d] = sum_e Wout[c][e] * Ctx[d][e] / S[d] * 32^-1/2   (second einsum + to_out conv as one 32 x Cout matrix)
__global__ void __launch_bounds__(256) fa_fold_kernel(const float* __restrict__ ctx, const float* __restrict__ wout /* [Cout][32] */,
                                                      __half* __restrict__ fold, int CO) {
    __shared__ float M[32][33];
    const int b = blockIdx.x;
    const float* c = ctx + (long long)b * 33 * 32;
    for (int i = threadIdx.x; i < 1024; i += 256) {
        const int d = i >> 5, e = i & 31;
        M[d][e] = c[i] / c[1024 + d] * 0.17677669529663687f;  // 32^-0.5
    }
    __syncthreads();
    for (int i = threadIdx.x; i < CO * 32; i += 256) {
        const int co = i >> 5, d = i & 31;
        float a = 0.f;
#pragma unroll 8
        for (int e = 0; e < 32; ++e) a = fmaf(__ldg(wout + co * 32 + e), M[d][e], a);
        fold[(long long)b * CO * 32 + i] = __float2half_rn(a);
    }
}

// ---------------------------------------------------------------------------------------------------------
// pass 3: q projection, softmax over d, (Ctx/S, to_out) as one folded 32 x Cout GEMM, bias, store, GroupNorm(1,Cout) sums
template <int KS, int NT>   // input channels 16*KS, output channel 8-blocks NT (Cout = 8*NT = input channels)
__global__ void __launch_bounds__(256, 2) fa_out_kernel(Geom g, const __half* __restrict__ x, const __half* __restrict__ wf,
                                                        const float* __restrict__ wsum, const float* __restrict__ wb,
                                                        const double* __restrict__ pn_stats, float pn_inv_count,
                                                        const __half* __restrict__ fold /* [B][Cout][32] */,
                                                        const float* __restrict__ bout, __half* __restrict__ dst,
                                                        double* __restrict__ stats, int rows_per_block) {
    constexpr int C = 16 * KS, STR = C * 2 + 16, CO = NT * 8, SBS = 80;
    __shared__ __align__(16) unsigned char sw[32 * STR];        // Wq' [d][c]
    __shared__ __align__(16) unsigned char sb[CO * SBS];        // folded B operand [c_out][d]
    __shared__ float red[2][kWarps];
    const int b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gq = lane >> 2, tq = lane & 3;
    stage_w<KS>(wf, sw, 32);
    {
        const __half* fb = fold + (long long)b * CO * 32;          // [c_out][d] fp16, written by fa_fold_kernel
        for (int i = threadIdx.x; i < CO * 4; i += 256) {
            const int co = i >> 2, q4 = i & 3;
            *reinterpret_cast<uint4*>(sb + co * SBS + q4 * 16) = *(reinterpret_cast<const uint4*>(fb + co * 32) + q4);
        }
        __syncthreads();
    }
    uint32_t bfrag[NT][4];
#pragma unroll
    for (int j = 0; j < NT; ++j) ldsm_x4(smem_u32(sb) + (j * 8 + (lane & 7)) * SBS + (lane >> 3) * 16, bfrag[j]);
    uint32_t wq[4][KS][2];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int sp = 0; sp < KS / 2; ++sp) {
            uint32_t r[4];
            ldsm_x4(smem_u32(sw) + (j * 8 + (lane & 7)) * STR + sp * 64 + (lane >> 3) * 16, r);
            wq[j][2 * sp][0] = r[0]; wq[j][2 * sp][1] = r[1]; wq[j][2 * sp + 1][0] = r[2]; wq[j][2 * sp + 1][1] = r[3];
        }
    const PnScal ps = pn_scalars(pn_stats, b, C, pn_inv_count);
    float cq[4][2], bias[NT][2];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int u = 0; u < 2; ++u) cq[j][u] = fmaf(ps.m, __ldg(wsum + 8 * j + 2 * tq + u), __ldg(wb + 8 * j + 2 * tq + u));
#pragma unroll
    for (int j = 0; j < NT; ++j) { bias[j][0] = __ldg(bout + j * 8 + 2 * tq); bias[j][1] = __ldg(bout + j * 8 + 2 * tq + 1); }
    const uint32_t r0 = blockIdx.x * rows_per_block;
    uint32_t r1 = r0 + rows_per_block;
    if (r1 > (uint32_t)g.rows_per_sample) r1 = g.rows_per_sample;
    const long long srow0 = (long long)b * g.rows_per_sample;
    const __half* xb = x + plcp_offset(g, 0, srow0);
    const long long cs = g.chunk_stride * 8;
    float ssum = 0.f, ssq = 0.f;
    // softmax in the log2 domain: t = log2e*(rstd*acc + const)
    const float a2 = ps.a * kLog2e;
#pragma unroll
    for (int j = 0; j < 4; ++j) { cq[j][0] *= kLog2e; cq[j][1] *= kLog2e; }
    uint32_t xf[4][KS][2];
    uint32_t base = r0 + warp * 32;
    if (base < r1) load_x32<KS>(xb, cs, base, r1, lane, xf);
    while (base < r1) {
        const uint32_t cur = base;
        const uint32_t rem = cur + lane;
        const bool inb = rem < r1;
        const unsigned vmask = __ballot_sync(0xffffffffu, inb && decode_subrow(g, rem).kind == 0);
        const unsigned imask = __ballot_sync(0xffffffffu, inb);     // rows to store: all of them (whole sectors); zeros outside the interior
        float q[2][4][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                q[mt][j][0] = q[mt][j][1] = q[mt][j][2] = q[mt][j][3] = 0.f;
#pragma unroll
                for (int s = 0; s < KS; ++s) {
                    const uint32_t a[4] = {xf[2 * mt][s][0], xf[2 * mt + 1][s][0], xf[2 * mt][s][1], xf[2 * mt + 1][s][1]};
                    mma16816(q[mt][j], a, wq[j][s][0], wq[j][s][1]);
                }
            }
        // the next group's x goes in flight under the softmax, the second GEMM and the stores
        base += kWarps * 32;
        if (base < r1) load_x32<KS>(xb, cs, base, r1, lane, xf);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {           // 16 rows per m tile
            float (&t)[4][4] = q[mt];
            // softmax over d: this lane holds 8 of the 32 channels of rows g and g+8; the lane quad holds the rest
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                t[j][0] = fmaf(a2, t[j][0], cq[j][0]); t[j][1] = fmaf(a2, t[j][1], cq[j][1]);
                t[j][2] = fmaf(a2, t[j][2], cq[j][0]); t[j][3] = fmaf(a2, t[j][3], cq[j][1]);
            }
            float m0 = t[0][0], m1 = t[0][2];
#pragma unroll
            for (int j = 0; j < 4; ++j) { m0 = fmaxf(m0, fmaxf(t[j][0], t[j][1])); m1 = fmaxf(m1, fmaxf(t[j][2], t[j][3])); }
            m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
            m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                t[j][0] = ex2f(t[j][0] - m0); t[j][1] = ex2f(t[j][1] - m0); s0 += t[j][0] + t[j][1];
                t[j][2] = ex2f(t[j][2] - m1); t[j][3] = ex2f(t[j][3] - m1); s1 += t[j][2] + t[j][3];
            }
            s0 += __shfl_xor_sync(0xffffffffu, s0, 1); s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
            s1 += __shfl_xor_sync(0xffffffffu, s1, 1); s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
            const float i0 = __fdividef(1.0f, s0), i1 = __fdividef(1.0f, s1);
            uint32_t a[2][4];
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
                a[ks][0] = pack_h2(t[2 * ks][0] * i0, t[2 * ks][1] * i0);
                a[ks][1] = pack_h2(t[2 * ks][2] * i1, t[2 * ks][3] * i1);
                a[ks][2] = pack_h2(t[2 * ks + 1][0] * i0, t[2 * ks + 1][1] * i0);
                a[ks][3] = pack_h2(t[2 * ks + 1][2] * i1, t[2 * ks + 1][3] * i1);
            }
            const bool v0 = (vmask >> (mt * 16 + gq)) & 1u, v1 = (vmask >> (mt * 16 + gq + 8)) & 1u;
            const bool st0 = (imask >> (mt * 16 + gq)) & 1u, st1 = (imask >> (mt * 16 + gq + 8)) & 1u;
            __half* o = dst + plcp_offset(g, 0, srow0 + cur + mt * 16 + gq) + 2 * tq;
#pragma unroll
            for (int j = 0; j < NT; ++j) {
                float c4[4] = {bias[j][0], bias[j][1], bias[j][0], bias[j][1]};
                mma16816(c4, a[0], bfrag[j][0], bfrag[j][1]);
                mma16816(c4, a[1], bfrag[j][2], bfrag[j][3]);
                // exact zeros outside the interior (a select, not a multiply: non-finite values must not leak into the zero rows)
                c4[0] = v0 ? c4[0] : 0.f; c4[1] = v0 ? c4[1] : 0.f; c4[2] = v1 ? c4[2] : 0.f; c4[3] = v1 ? c4[3] : 0.f;
                if (st0) *reinterpret_cast<__half2*>(o + j * cs) = __floats2half2_rn(c4[0], c4[1]);
                if (st1) *reinterpret_cast<__half2*>(o + j * cs + 64) = __floats2half2_rn(c4[2], c4[3]);      // row + 8
                ssum += (c4[0] + c4[1]) + (c4[2] + c4[3]);
                ssq = fmaf(c4[0], c4[0], fmaf(c4[1], c4[1], fmaf(c4[2], c4[2], fmaf(c4[3], c4[3], ssq))));
            }
        }
    }
    if (stats) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { ssum += __shfl_xor_sync(0xffffffffu, ssum, o); ssq += __shfl_xor_sync(0xffffffffu, ssq, o); }
        if (lane == 0) { red[0][warp] = ssum; red[1][warp] = ssq; }
        __syncthreads();
        if (threadIdx.x < 2) {
            double t = 0.0;
            for (int i = 0; i < kWarps; ++i) t += (double)red[threadIdx.x][i];
            atomicAdd(stats + (long long)b * CO * 2 + threadIdx.x, t);   // GroupNorm(1,C): one group, channel-0 slot
        }
    }
}

template <int KS>
void launch_fused_t(const Geom& g, const __half* x, const __half* wf, const float* wsum, const float* wb, const double* pn_stats,
                    float inv_count, unsigned* kmax, float* ctx, float* part, __half* fold, const float* wout, const float* bout,
                    __half* dst, double* stats, int B, cudaStream_t st) {
    // rows per block: as many as leaves >= ~4 waves of blocks (2 x 148 resident): the per-block set-up (weights, per-sample
    // scalars, partial write-out) is amortised over the block's rows.  CD_ATTN_ROWS / CD_ATTN_OROWS override (diagnostics).
    auto pick = [&](int minrows, const char* env) {
        if (const char* e = getenv(env)) { const int v = atoi(e); if (v >= 256) return (v + 255) / 256 * 256; }
        int rows = (g.rows_per_sample + 255) / 256 * 256;
        while (rows > minrows && (long long)B * ((g.rows_per_sample + rows - 1) / rows) < 1184) {
            int half = (rows / 2 + 255) / 256 * 256;
            if (half >= rows) break;
            rows = half < minrows ? minrows : half;
        }
        return rows;
    };
    int rows = pick(512, "CD_ATTN_ROWS");
    // the online form merges at most kAttnMaxBlocks partials per sample
    while ((g.rows_per_sample + rows - 1) / rows > kAttnMaxBlocks) rows += 32 * kWarps;
    const int nblk = (g.rows_per_sample + rows - 1) / rows;
    dim3 grid(nblk, B);
    if (part) {
        fa_ctx_online_kernel<KS><<<grid, 256, 0, st>>>(g, x, wf, wsum, wb, pn_stats, inv_count, part, rows);
        fa_fold_online_kernel<<<B, 256, 0, st>>>(part, nblk, pn_stats, inv_count, 16 * KS, wout, fold, 16 * KS);
    } else {
        fa_kmax_kernel<KS><<<grid, 256, 0, st>>>(g, x, wf, kmax, rows);
        fa_ctx_kernel<KS><<<grid, 256, 0, st>>>(g, x, wf, wsum, wb, pn_stats, inv_count, kmax, ctx, rows, nblk > 1 ? 1 : 0);
        fa_fold_kernel<<<B, 256, 0, st>>>(ctx, wout, fold, 16 * KS);
    }
    const int orows = pick(512, "CD_ATTN_OROWS");
    dim3 ogrid((g.rows_per_sample + orows - 1) / orows, B);
    fa_out_kernel<KS, 2 * KS><<<ogrid, 256, 0, st>>>(g, x, wf, wsum, wb, pn_stats, inv_count, fold, bout, dst, stats, orows);
}

}  // namespace

bool attention_fused_supported(int C) { return C == 32 || C == 64; }
int attn_part_floats() { return kPartFloats; }

// x: raw (un-normalised) C-channel PLCP input of the attention block; wf: fp16 [96][C] = to_qkv weight * diag(gamma);
// wsum[96] = row sums of wf as fp16, wb[96] = W beta; pn_stats: GroupNorm(1,C) sums of x (channel-0 slot per sample).
// kmax, ctx and stats must be zero on entry; fold is [B][C][32] fp16 scratch; part (optional) is
// [B][kAttnMaxBlocks][attn_part_floats()] floats and selects the online-softmax form (no max pass).  Returns 0 if launched.
int launch_attention_fused(const Geom& g, const __half* x, const __half* wf, const float* wsum, const float* wb,
                           const double* pn_stats, double pn_count, unsigned* kmax, float* ctx, float* part, __half* fold,
                           const float* wout, const float* bout, __half* dst, int C, double* stats, int B, cudaStream_t st) {
    const float inv = (float)(1.0 / pn_count);
    if (C == 32) launch_fused_t<2>(g, x, wf, wsum, wb, pn_stats, inv, kmax, ctx, part, fold, wout, bout, dst, stats, B, st);
    else if (C == 64) launch_fused_t<4>(g, x, wf, wsum, wb, pn_stats, inv, kmax, ctx, part, fold, wout, bout, dst, stats, B, st);
    else return 1;
    return 0;
}

}  // namespace cd
