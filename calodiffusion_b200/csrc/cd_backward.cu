// Backward kernels of the training step (Loss.__call__ models/loss.py:118-142, hybrid_weight :168-179,
// training_loop train/train_diffusion.py:42-66).  The forward pass is the inference engine run without buffer reuse, so every
// activation a backward kernel needs is still in its PLCP fp16 tensor.  Activation gradients are PLCP fp16 as well, scaled by
// a constant loss scale (fp16 keeps 11 significant bits but only 5 exponent bits: the 1/numel of the loss would underflow);
// parameter gradients accumulate in fp32 in a flat buffer laid out like the master parameters and carry the same scale,
// which the optimizer divides out.
//   * stride-1 3x3x3 and 1x1x1 convolutions: data gradient = the SAME tcgen05 implicit-GEMM kernel (cd_conv_mma.cu) run on
//     the gradient tensor with tap-flipped, channel-transposed weights (circular phi / zero Z,R padding are self-adjoint);
//     weight gradient = wgrad_mma_kernel below (mma.sync m16n8k16, K = voxels x batch).
//   * strided / transposed convolutions (4 per network, 5 % of the FLOPs): exact adjoints by inverse index enumeration (SIMT).
//   * GroupNorm + SiLU + conditioning + residual, linear attention, the conditioning MLPs, the final 1x1x1 conv + loss.
#include <cstdlib>
#include "cd_kernels.cuh"

namespace cd {
namespace {

__device__ __forceinline__ float sigmoid_f(float v) { return 1.f / (1.f + __expf(-v)); }
__device__ __forceinline__ float block_reduce_sum(float v, float* red) {   // red: >= 32 floats of shared memory
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    v = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
    if (warp == 0) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    }
    return v;   // valid in thread 0
}

// =================================================================================================
// Loss + final 1x1x1 conv backward.  hybrid_weight (loss.py:168-179) with the l2 form (:103-104):
//   L = sum_b w_b sum_v (x0 - data)^2 / (mean(w) * numel),  w_b = 1 + 1/sigma_b^2,  x0 = c_skip*x + c_out*pred
// coef[b] = loss_scale * 2 * w_b / (mean(w) * numel) is prepared by the caller; lossw[b] = w_b / (mean(w) * numel).
// Thread = interior voxel: writes the gradient of the 32-channel tensor feeding the final conv and reduces dW / db / L.
// =================================================================================================
__global__ void __launch_bounds__(256) loss_backward_kernel(LossBwdParams p, long long nvox) {
    __shared__ float red[32];
    __shared__ float s_w[kBwdMaxC];
    for (int c = threadIdx.x; c < p.C; c += blockDim.x) s_w[c] = p.w[c];
    __syncthreads();
    const long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const Geom& g = p.g;
    float dpred = 0.f, lossv = 0.f;
    float hw[kBwdMaxC / 8][8];
    const int nch = p.C >> 3;
    long long row = 0;
    if (v < nvox) {
        const int vps = g.Z * g.P * g.R;
        const int b = (int)(v / vps);
        const int rem = (int)(v - (long long)b * vps);
        const int z = rem / (g.P * g.R);
        const int rem2 = rem - z * g.P * g.R;
        const int ph = rem2 / g.R, r = rem2 - ph * g.R;
        row = interior_row(g, b, z, ph, r);
        float pred = *p.bias;
        for (int k = 0; k < nch; ++k) {
            unpack8(load8(p.h + plcp_offset(g, k, row)), hw[k]);
#pragma unroll
            for (int j = 0; j < 8; ++j) pred += hw[k][j] * s_w[k * 8 + j];
        }
        const float s = p.sigma[b], sd = p.sigma_data, den = s * s + sd * sd;
        const float c_skip = sd * sd / den, c_out = s * sd * rsqrtf(den);
        const float x0 = c_skip * p.x[v] + c_out * pred;
        const float diff = x0 - p.data[v];
        lossv = p.lossw[b] * diff * diff;
        dpred = p.coef[b] * diff * c_out;
        for (int k = 0; k < nch; ++k) {
            float f[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = s_w[k * 8 + j] * dpred;
            store8(p.dh + plcp_offset(g, k, row), pack8(f));
        }
        if (p.x0_out) p.x0_out[v] = x0;
    }
    // reductions: dW[c] = sum h[c]*dpred, db = sum dpred, loss
    for (int k = 0; k < nch; ++k)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float t = block_reduce_sum(v < nvox ? hw[k][j] * dpred : 0.f, red);
            if (threadIdx.x == 0) atomicAdd(p.dw + k * 8 + j, t);
        }
    float t = block_reduce_sum(dpred, red);
    if (threadIdx.x == 0) atomicAdd(p.db, t);
    t = block_reduce_sum(lossv, red);
    if (threadIdx.x == 0) atomicAdd(p.loss, (double)t);
}

// =================================================================================================
// GroupNorm(+SiLU)(+conditioning)(+residual) backward:  y = act(gamma*xh + beta) + addvec[b,c] + res,  xh = (x - mu)*rstd
//   pass 1 (reduce):   per (b,c): A = sum dz*xh, Bc = sum dz, D = sum dy, X = sum xh      (dz = dy*act'(z))
//   finalize:          per (b,group): m1 = sum_c gamma*Bc / n, m2 = sum_c gamma*A / n;  dgamma += A, dbeta += Bc, daddvec += D,
//                      d(conv bias)[c] += rstd*(gamma*Bc - V*m1 - m2*X)
//   pass 2 (apply):    dx = rstd*(gamma*dz - m1 - xh*m2) for every row incl. halos (the tcgen05 data-gradient conv reads
//                      neighbours);  residual gradient: dres (+)= dy
// =================================================================================================
constexpr int kGbThreads = 256;

__device__ __forceinline__ void gn_group_stats(const double* stats, int b, int C, int groups, int c, double count_per_channel,
                                               float& mean, float& rstd) {
    const int cpg = C / groups, g0 = (c / cpg) * cpg;
    double s = 0.0, q = 0.0;
    for (int j = 0; j < cpg; ++j) { const double* d = stats + ((long long)b * C + g0 + j) * 2; s += d[0]; q += d[1]; }
    const double n = count_per_channel * cpg;
    mean = (float)(s / n);
    float var = (float)(q / n - (s / n) * (s / n));
    if (var < 0.f) var = 0.f;
    rstd = rsqrtf(var + 1e-5f);
}

// one block per sample: mean / rstd of every channel's group from the forward pass's fp64 sums (once, instead of in every
// block of the two streaming passes), and the accumulators cleared
__global__ void __launch_bounds__(128) gn_bwd_prep_kernel(GnBwdParams p) {
    const int b = blockIdx.x;
    for (int c = threadIdx.x; c < p.C; c += blockDim.x) {
        float mean, rstd;
        gn_group_stats(p.stats, b, p.C, p.groups, c, p.count_per_channel, mean, rstd);
        *reinterpret_cast<float2*>(p.mr + ((long long)b * p.C + c) * 2) = make_float2(mean, rstd);
        *reinterpret_cast<float4*>(p.acc + ((long long)b * p.C + c) * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// grid (row chunks, B); thread layout: chunk-major so that a warp reads 512 contiguous bytes of one chunk plane
__global__ void __launch_bounds__(kGbThreads) gn_bwd_reduce_kernel(GnBwdParams p) {
    __shared__ float s_mean[kBwdMaxC], s_rstd[kBwdMaxC], s_gamma[kBwdMaxC], s_beta[kBwdMaxC];
    __shared__ float s_acc[kBwdMaxC][4];
    const int b = blockIdx.y, C = p.C, nch = C >> 3;
    const Geom& g = p.g;
    for (int c = threadIdx.x; c < C; c += kGbThreads) {
        const float2 mr = *reinterpret_cast<const float2*>(p.mr + ((long long)b * C + c) * 2);
        s_mean[c] = mr.x; s_rstd[c] = mr.y;
        s_gamma[c] = p.gamma[c]; s_beta[c] = p.beta[c];
        s_acc[c][0] = s_acc[c][1] = s_acc[c][2] = s_acc[c][3] = 0.f;
    }
    __syncthreads();
    const int tpc = kGbThreads / nch;               // threads per chunk plane
    const int k = threadIdx.x / tpc, t = threadIdx.x - k * tpc;
    float A[8], Bc[8], D[8], X[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) A[j] = Bc[j] = D[j] = X[j] = 0.f;
    if (k < nch) {
        const uint32_t r0 = blockIdx.x * p.rows_per_block;
        uint32_t r1 = r0 + p.rows_per_block;
        if (r1 > (uint32_t)g.rows_per_sample) r1 = g.rows_per_sample;
        for (uint32_t rem = r0 + t; rem < r1; rem += tpc) {
            if (decode_subrow(g, rem).kind != 0) continue;
            const long long row = (long long)b * g.rows_per_sample + rem;
            float x[8], dy[8];
            unpack8(load8(p.x + plcp_offset(g, k, row)), x);
            unpack8(load8(p.dy + plcp_offset(g, k, row)), dy);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int c = k * 8 + j;
                const float xh = (x[j] - s_mean[c]) * s_rstd[c];
                float dz = dy[j];
                if (p.act) {
                    const float z = fmaf(s_gamma[c], xh, s_beta[c]);
                    const float sg = sigmoid_f(z);
                    dz *= sg * (1.f + z * (1.f - sg));
                }
                A[j] = fmaf(dz, xh, A[j]); Bc[j] += dz; D[j] += dy[j]; X[j] += xh;
            }
        }
    }
    // lanes of one chunk plane are contiguous (tpc is 8, 16, 32, 64 ...): butterfly over min(32, tpc) lanes, then one lane per
    // group adds to the block's accumulators (all lanes take part in the shuffles, also those with k >= nch)
    // (a width whose chunk count does not divide the block - 96 channels: 21 threads per plane - has no lane-aligned groups:
    //  span 1, every thread adds its own partial sums)
    const int span = (tpc & (tpc - 1)) != 0 ? 1 : (tpc < 32 ? tpc : 32);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float a = A[j], bsum = Bc[j], d = D[j], x = X[j];
        for (int o = span >> 1; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o); bsum += __shfl_xor_sync(0xffffffffu, bsum, o);
            d += __shfl_xor_sync(0xffffffffu, d, o); x += __shfl_xor_sync(0xffffffffu, x, o);
        }
        if (k < nch && (threadIdx.x & (span - 1)) == 0) {
            atomicAdd(&s_acc[k * 8 + j][0], a); atomicAdd(&s_acc[k * 8 + j][1], bsum);
            atomicAdd(&s_acc[k * 8 + j][2], d); atomicAdd(&s_acc[k * 8 + j][3], x);
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += kGbThreads)
#pragma unroll
        for (int i = 0; i < 4; ++i) atomicAdd(p.acc + ((long long)b * C + c) * 4 + i, s_acc[c][i]);
}

// one block per sample
__global__ void __launch_bounds__(256) gn_bwd_finalize_kernel(GnBwdParams p) {
    __shared__ float s_m1[kBwdMaxC], s_m2[kBwdMaxC];
    const int b = blockIdx.x, C = p.C, cpg = C / p.groups;
    const float V = (float)p.count_per_channel, n = V * cpg;
    for (int gi = threadIdx.x; gi < p.groups; gi += blockDim.x) {
        float a = 0.f, bb = 0.f;
        for (int j = 0; j < cpg; ++j) {
            const int c = gi * cpg + j;
            const float* acc = p.acc + ((long long)b * C + c) * 4;
            a += p.gamma[c] * acc[0]; bb += p.gamma[c] * acc[1];
        }
        s_m2[gi] = a / n; s_m1[gi] = bb / n;
        p.m12[((long long)b * p.groups + gi) * 2] = bb / n;
        p.m12[((long long)b * p.groups + gi) * 2 + 1] = a / n;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float* acc = p.acc + ((long long)b * C + c) * 4;
        atomicAdd(p.dgamma + c, acc[0]);
        atomicAdd(p.dbeta + c, acc[1]);
        if (p.daddvec) p.daddvec[(long long)b * p.addvec_stride + c] += acc[2];     // one block per sample: no race
        if (p.dbias) {
            const float rstd = p.mr[((long long)b * C + c) * 2 + 1];
            const int gi = c / cpg;
            atomicAdd(p.dbias + c, rstd * (p.gamma[c] * acc[1] - V * s_m1[gi] - s_m2[gi] * acc[3]));
        }
    }
}

// grid (row chunks, B); thread = padded-linear row, all chunks (like gn_apply_kernel)
__global__ void __launch_bounds__(kGbThreads) gn_bwd_apply_kernel(GnBwdParams p) {
    __shared__ float s_mean[kBwdMaxC], s_rstd[kBwdMaxC], s_gamma[kBwdMaxC], s_beta[kBwdMaxC], s_m1[kBwdMaxC], s_m2[kBwdMaxC];
    const int b = blockIdx.y, C = p.C, nch = C >> 3, cpg = C / p.groups;
    const Geom& g = p.g;
    for (int c = threadIdx.x; c < C; c += kGbThreads) {
        const float2 mr = *reinterpret_cast<const float2*>(p.mr + ((long long)b * C + c) * 2);
        s_mean[c] = mr.x; s_rstd[c] = mr.y;
        s_gamma[c] = p.gamma[c]; s_beta[c] = p.beta[c];
        s_m1[c] = p.m12[((long long)b * p.groups + c / cpg) * 2];
        s_m2[c] = p.m12[((long long)b * p.groups + c / cpg) * 2 + 1];
    }
    __syncthreads();
    const uint32_t rem = blockIdx.x * kGbThreads + threadIdx.x;
    if (rem >= (uint32_t)g.rows_per_sample) return;
    const SubRow sr = decode_subrow(g, rem);
    const long long row = (long long)b * g.rows_per_sample + rem;
    const long long srow = row + sr.src_delta;
    for (int k = 0; k < nch; ++k) {
        float dx[8], dyv[8];
        if (sr.kind == 2) {
#pragma unroll
            for (int j = 0; j < 8; ++j) dx[j] = dyv[j] = 0.f;
        } else {
            float x[8];
            unpack8(load8(p.x + plcp_offset(g, k, srow)), x);
            unpack8(load8(p.dy + plcp_offset(g, k, srow)), dyv);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int c = k * 8 + j;
                const float xh = (x[j] - s_mean[c]) * s_rstd[c];
                float dz = dyv[j];
                if (p.act) {
                    const float z = fmaf(s_gamma[c], xh, s_beta[c]);
                    const float sg = sigmoid_f(z);
                    dz *= sg * (1.f + z * (1.f - sg));
                }
                dx[j] = s_rstd[c] * (s_gamma[c] * dz - s_m1[c] - xh * s_m2[c]);
            }
        }
        __half* o = p.dx + plcp_offset(g, k, row);
        if (p.dx_accumulate) {
            float old[8];
            unpack8(load8(o), old);
#pragma unroll
            for (int j = 0; j < 8; ++j) dx[j] += old[j];
        }
        store8(o, pack8(dx));
        // residual gradient: the residual may be one tensor or a concatenation of two (identity residual over a concat input)
        if (p.dres) {
            __half* r = (k < p.res_split) ? p.dres + plcp_offset(g, k, row) : p.dres2 + plcp_offset(g, k - p.res_split, row);
            const int accf = (k < p.res_split) ? p.dres_accumulate : p.dres2_accumulate;
            float rv[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) rv[j] = dyv[j];
            if (accf) {
                float old[8];
                unpack8(load8(r), old);
#pragma unroll
                for (int j = 0; j < 8; ++j) rv[j] += old[j];
            }
            store8(r, pack8(rv));
        }
    }
}

// =================================================================================================
// Weight gradient of a stride-1 CylindricalConv (3x3x3 or 1x1x1) on the tensor cores (mma.sync m16n8k16):
//   dW[tap][ci][co] = sum over interior rows m of X[m + off(tap)][ci] * dY[m][co]       (constant row offsets in PLCP)
// One CTA = one 32(ci) x 32(co) block of every tap; it walks a strided set of 128-row chunks of the whole batch, stages the
// X window (chunk + halo rows) and the dY chunk (non-interior rows zeroed) in shared memory in the global chunk-planar
// layout, and each of its 8 warps accumulates its taps' 32x32 blocks in registers (A = X^T and B = dY through
// ldmatrix.trans, the same operand scheme as attn_context_kernel).  Partial blocks are added to the fp32 gradient with atomics.
// =================================================================================================
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t* r) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

constexpr int kWgRows = 128;          // rows per chunk (K of the MMAs: 8 k16 steps)
constexpr int kWgWarps = 16;
constexpr int kWgMaxTapsPerWarp = 2;  // 27 taps over 16 warps (8 warps x 4 taps: 150 TFLOP/s, 16 x 2: 198, 28 x 1: 167 at DS2 L0)

__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gsrc, bool valid) {
    const uint32_t n = valid ? 16u : 0u;       // src-size 0: the 16 bytes are zero-filled, the source is not read
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(n) : "memory");
}

// The CTA walks a CONTIGUOUS range of 128-row chunks, so the X window (chunk + halo rows on both sides) slides: it lives in a
// ring of wrows + 128 rows per chunk plane and every step fetches only the 128 new rows (cp.async, overlapped with the MMAs
// of the current chunk) - X is read ~once instead of (128 + 2*halo)/128 times.  dY chunks are double-buffered the same way.
__global__ void __launch_bounds__(kWgWarps * 32) wgrad_mma_kernel(WgradParams p) {
    extern __shared__ __align__(16) unsigned char wsm[];
    const Geom& g = p.g;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cib = blockIdx.y / p.ncob, cob = blockIdx.y - cib * p.ncob;     // 32-channel blocks of ci / co
    const int halo = p.taps == 27 ? g.plane + g.Rp + 1 : 0;
    const int wrows = kWgRows + 2 * halo;                                       // X window rows
    const int ring = wrows + kWgRows;
    __half* xs = reinterpret_cast<__half*>(wsm);                               // [4 chunks][ring][8]
    __half* ys = xs + (size_t)4 * ring * 8;                                    // [2][4 chunks][kWgRows][8]
    // source tensor / chunk of this ci block (a conv over a concatenated input has two sources)
    const int c0 = cib * 4;                                                     // first 8-channel chunk of the block
    const __half* xsrc; int xk0;
    if (c0 < p.src_chunks0) { xsrc = p.x0; xk0 = c0; } else { xsrc = p.x1; xk0 = c0 - p.src_chunks0; }
    const int navail = (c0 < p.src_chunks0 ? p.src_chunks0 : p.Cin_pad / 8) - c0;   // chunk planes that exist (a 16-channel input has 2)
    // 27 taps: warps split the taps.  1 tap: every warp takes it and the warps split the k16 steps of each chunk instead.
    const bool ksplit = p.taps == 1;
    int ntap_w = 0, taps_w[kWgMaxTapsPerWarp], off_w[kWgMaxTapsPerWarp];
    if (ksplit) { taps_w[0] = 0; ntap_w = 1; }
    else for (int t = warp; t < p.taps; t += kWgWarps) taps_w[ntap_w++] = t;
#pragma unroll
    for (int a = 0; a < kWgMaxTapsPerWarp; ++a) {
        off_w[a] = 0;
        if (a < ntap_w && p.taps == 27) { const int t = taps_w[a]; off_w[a] = (t / 9 - 1) * g.plane + ((t / 3) % 3 - 1) * g.Rp + (t % 3 - 1); }
    }
    float acc[kWgMaxTapsPerWarp][2][4][4];       // [tap][ci 16-tile][co 8-tile][frag]
#pragma unroll
    for (int a = 0; a < kWgMaxTapsPerWarp; ++a)
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int e = 0; e < 4; ++e) acc[a][i][j][e] = 0.f;
    const long long nrows = (long long)p.B * g.rows_per_sample;
    const int nchunks = (int)((nrows + kWgRows - 1) / kWgRows);
    const int per = (nchunks + gridDim.x - 1) / gridDim.x;
    const int c_lo = blockIdx.x * per, c_hi = min(nchunks, c_lo + per);
    if (c_lo >= c_hi) return;
    const long long rbase = (long long)c_lo * kWgRows - halo;                  // global row of ring-relative row 0
    // rows [rel0, rel0 + count) of the window sequence into their ring slots
    auto load_x = [&](int rel0, int count) {
        for (int k = 0; k < 4; ++k)
            for (int r = threadIdx.x; r < count; r += kWgWarps * 32) {
                const int rel = rel0 + r;
                const long long row = rbase + rel;
                const bool ok = k < navail && row >= -(long long)g.guard && row < nrows + g.guard;
                cp_async16_zfill(xs + ((size_t)k * ring + rel % ring) * 8, ok ? (const void*)(xsrc + plcp_offset(g, xk0 + k, row)) : (const void*)xsrc, ok);
            }
    };
    auto load_y = [&](int buf, int ch) {
        for (int i = threadIdx.x; i < 4 * kWgRows; i += kWgWarps * 32) {
            const int k = i / kWgRows, r = i - k * kWgRows;
            const long long row = (long long)ch * kWgRows + r;
            bool ok = row < nrows;
            if (ok) ok = decode_subrow(g, (uint32_t)(row % g.rows_per_sample)).kind == 0;     // interior rows only
            cp_async16_zfill(ys + (((size_t)buf * 4 + k) * kWgRows + r) * 8, ok ? (const void*)(p.dy + plcp_offset(g, cob * 4 + k, row)) : (const void*)p.dy, ok);
        }
    };
    load_x(0, wrows);
    load_y(0, c_lo);
    asm volatile("cp.async.commit_group;" ::: "memory");
    const int mi = lane >> 3, mr = lane & 7;
    int wstart = 0;                                                             // ring slot of the window's first row
    for (int i = 0; c_lo + i < c_hi; ++i) {
        if (c_lo + i + 1 < c_hi) { load_x(wrows + i * kWgRows, kWgRows); load_y((i + 1) & 1, c_lo + i + 1); }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();
        const __half* yb = ys + (size_t)(i & 1) * 4 * kWgRows * 8;
        int abase[kWgMaxTapsPerWarp];
#pragma unroll
        for (int a = 0; a < kWgMaxTapsPerWarp; ++a) {
            int b = wstart + halo + off_w[a] + (mi >> 1) * 8 + mr;            // < ring + wrows < 2*ring
            if (b >= ring) b -= ring;
            abase[a] = b;
        }
#pragma unroll 2
        for (int k16 = ksplit ? warp : 0; k16 < kWgRows / 16; k16 += ksplit ? kWgWarps : 1) {
            // B fragments (dY): matrices (rows k 0-7 | 8-15) x (co chunk j): x4.trans -> b0,b1 of two co 8-tiles
            uint32_t bf[4][2];
#pragma unroll
            for (int jj = 0; jj < 2; ++jj) {
                uint32_t r[4];
                // m0: (k 0-7, chunk 2jj) m1: (k 8-15, chunk 2jj) m2: (k 0-7, chunk 2jj+1) m3: (k 8-15, chunk 2jj+1)
                const int chunk = 2 * jj + (mi >> 1), krow = k16 * 16 + (mi & 1) * 8 + mr;
                ldsm_x4_t(smem_u32(yb + ((size_t)chunk * kWgRows + krow) * 8), r);
                bf[2 * jj][0] = r[0]; bf[2 * jj][1] = r[1]; bf[2 * jj + 1][0] = r[2]; bf[2 * jj + 1][1] = r[3];
            }
#pragma unroll
            for (int a = 0; a < kWgMaxTapsPerWarp; ++a) {
                if (a >= ntap_w) break;
                int krow = abase[a] + k16 * 16;
                if (krow >= ring) krow -= ring;
#pragma unroll
                for (int it = 0; it < 2; ++it) {
                    // A fragment (X^T), ci 16-tile it: a0 (ci 0-7, k 0-7) a1 (ci 8-15, k 0-7) a2 (ci 0-7, k 8-15) a3 (ci 8-15, k 8-15)
                    uint32_t af[4];
                    const int chunk = 2 * it + (mi & 1);
                    ldsm_x4_t(smem_u32(xs + ((size_t)chunk * ring + krow) * 8), af);
#pragma unroll
                    for (int j = 0; j < 4; ++j) mma16816(acc[a][it][j], af, bf[j][0], bf[j][1]);
                }
            }
        }
        __syncthreads();
        wstart += kWgRows;
        if (wstart >= ring) wstart -= ring;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // D fragment: c0,c1 -> (row g, cols 2t,2t+1), c2,c3 -> (row g+8, ...)   row = ci within the 16-tile, col = co within the 8-tile
    const int gq = lane >> 2, tq = lane & 3;
#pragma unroll
    for (int a = 0; a < kWgMaxTapsPerWarp; ++a) {
        if (a >= ntap_w) break;
        const int tap = taps_w[a];
#pragma unroll
        for (int it = 0; it < 2; ++it)
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int ci = cib * 32 + it * 16 + gq + (e >> 1) * 8, co = cob * 32 + j * 8 + tq * 2 + (e & 1);
                    if (ci < p.Cin_real && acc[a][it][j][e] != 0.f)
                        atomicAdd(p.dw + ((long long)co * p.Cin_real + ci) * p.taps + tap, acc[a][it][j][e]);   // master layout [co][ci][tap]
                }
    }
}

// =================================================================================================
// Strided / transposed CylindricalConv backward (models.py:25-96 with stride 2), exact adjoints of conv_generic_kernel.
// Data gradient: thread = (input row, 8 input channels); for every tap and every circular copy of its phi position it finds
// the output voxel that read it (inverse of the forward index map) and gathers W[tap][ci][co] * dY[out][co].
// Weight gradient: block = (tap, 32 ci x 32 co block, row split); threads own (co, 4 ci) and walk the output voxels.
// =================================================================================================
__device__ __forceinline__ bool strided_out_pos(const ConvParams& p, int zi, int pi, int ri, int a, int bq, int c, int copy,
                                                int& zo, int& po, int& ro) {
    const Geom& go = p.gout;
    const Geom& gi = p.gin;
    const int q = pi + p.cp + copy * gi.P;                  // position on the circularly padded phi axis
    if (q < 0 || q >= gi.P + 2 * p.cp) return false;
    if (!p.transposed) {
        int t = zi + p.pz - a; if (t < 0 || t % p.sz) return false; zo = t / p.sz;
        t = q - bq;            if (t < 0 || t % p.sp) return false; po = t / p.sp;
        t = ri + p.pr - c;     if (t < 0 || t % p.sr) return false; ro = t / p.sr;
    } else {
        zo = zi * p.sz + a - p.pz;
        po = q * p.sp + bq - (p.kp - 1);
        ro = ri * p.sr + c - p.pr;
    }
    return zo >= 0 && zo < go.Z && po >= 0 && po < go.P && ro >= 0 && ro < go.R;
}

__global__ void conv_strided_dgrad_kernel(ConvParams p, const __half* __restrict__ dy, __half* __restrict__ dx, int dx_accumulate,
                                          long long nrows_in) {
    const long long row = (long long)blockIdx.x * 32 + threadIdx.x;
    const int cig = threadIdx.y;                            // input chunk
    if (row >= nrows_in) return;
    const Geom& gi = p.gin;
    const RowInfo ri = decode_row(gi, row);
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    if (ri.kind != 2) {
        int pi = ri.pp - 1;
        if (pi < 0) pi += gi.P;
        if (pi >= gi.P) pi -= gi.P;
        const int nco = p.Cout >> 3;
        for (int a = 0; a < p.kz; ++a)
            for (int bq = 0; bq < p.kp; ++bq)
                for (int c = 0; c < p.kr; ++c)
                    for (int copy = -1; copy <= 1; ++copy) {
                        int zo, po, ro;
                        if (!strided_out_pos(p, ri.z, pi, ri.rr, a, bq, c, copy, zo, po, ro)) continue;
                        const long long orow = interior_row(p.gout, ri.b, zo, po, ro);
                        const int tap = (a * p.kp + bq) * p.kr + c;
                        const float* wt = p.w + ((long long)tap * p.Cin_total + cig * 8) * p.Cout;
                        for (int k = 0; k < nco; ++k) {
                            float f[8];
                            unpack8(load8(dy + plcp_offset(p.gout, k, orow)), f);
#pragma unroll
                            for (int i = 0; i < 8; ++i)
#pragma unroll
                                for (int j = 0; j < 8; ++j) acc[i] = fmaf(__ldg(wt + (long long)i * p.Cout + k * 8 + j), f[j], acc[i]);
                        }
                    }
    }
    __half* o = dx + plcp_offset(gi, cig, row);
    if (dx_accumulate) {
        float old[8];
        unpack8(load8(o), old);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] += old[i];
    }
    store8(o, pack8(acc));
}

// Weight gradient of a strided / transposed conv on the tensor cores.  grid (taps, ci blocks * co blocks, splits); the block
// walks 128-voxel tiles of the COARSE grid (output voxels of a strided conv; circularly padded input voxels of a transposed
// conv, whose tap (a,b,c) lands on output (2z+a-pz, 2q+b-(kp-1), 2r+c-pr)), so every staged (voxel, tap) pair is a real
// product.  Thread pair = one voxel: one thread gathers its 4 dY chunks, the other its 4 X chunks (cp.async, zero fill where
// the pair does not exist) into fp16 chunk planes [chunk][voxel][8], double-buffered; then dW[32 ci][32 co] += X^T dY with
// mma.sync m16n8k16 (K = the 128 voxels, one k16 step per warp; operands through ldmatrix.trans as in wgrad_mma_kernel).
__global__ void __launch_bounds__(256) conv_strided_wgrad_kernel(ConvParams p, const __half* __restrict__ x, const __half* __restrict__ dy,
                                                                 float* __restrict__ dw, int Cin_real, int master_transposed) {
    __shared__ __align__(16) __half sx[2][4 * 128 * 8], sy[2][4 * 128 * 8];
    const int tap = blockIdx.x;
    const int ncob = p.Cout / 32;
    const int cib = blockIdx.y / ncob, cob = blockIdx.y - cib * ncob;
    const int a = tap / (p.kp * p.kr), bq = (tap / p.kr) % p.kp, c = tap % p.kr;
    const Geom& go = p.gout;
    const Geom& gi = p.gin;
    const int wZ = p.transposed ? gi.Z : go.Z, wP = p.transposed ? gi.P + 2 * p.cp : go.P, wR = p.transposed ? gi.R : go.R;
    const int vps = wZ * wP * wR;
    const long long nvox = (long long)p.B * vps;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, mi = lane >> 3, mr = lane & 7;
    const int sv = threadIdx.x >> 1, sh = threadIdx.x & 1;          // staging role: voxel of the tile, 0 = dY / 1 = X
    const int xavail = (p.Cin_total >> 3) - cib * 4;                // X chunk planes that exist in this ci block
    auto stage = [&](int buf, long long v0) {
        const long long v = v0 + sv;
        bool ok = v < nvox;
        long long row = 0;
        if (ok) {
            const int b = (int)(v / vps);
            int rem = (int)(v - (long long)b * vps);
            const int wz = rem / (wP * wR); rem -= wz * wP * wR;
            const int wp = rem / wR, wr = rem - wp * wR;
            int zo, po, ro, zi, pin, rin;
            if (!p.transposed) {
                zo = wz; po = wp; ro = wr;
                zi = zo * p.sz + a - p.pz; rin = ro * p.sr + c - p.pr; pin = po * p.sp + bq - p.cp;
            } else {
                zi = wz; rin = wr; pin = wp - p.cp;
                zo = zi * p.sz + a - p.pz; ro = rin * p.sr + c - p.pr; po = wp * p.sp + bq - (p.kp - 1);
            }
            pin %= gi.P; if (pin < 0) pin += gi.P;
            ok = zi >= 0 && zi < gi.Z && rin >= 0 && rin < gi.R && zo >= 0 && zo < go.Z && po >= 0 && po < go.P && ro >= 0 && ro < go.R;
            if (ok) row = sh ? interior_row(gi, b, zi, pin, rin) : interior_row(go, b, zo, po, ro);
        }
        __half* dstp = (sh ? sx[buf] : sy[buf]) + sv * 8;
        const __half* srcp = sh ? x : dy;
        const Geom& gs = sh ? gi : go;
        const int k0 = sh ? cib * 4 : cob * 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool okk = ok && (!sh || k < xavail);
            cp_async16_zfill(dstp + k * 128 * 8, okk ? (const void*)(srcp + plcp_offset(gs, k0 + k, row)) : (const void*)srcp, okk);
        }
    };
    float acc[2][4][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;
    const long long stride = (long long)gridDim.z * 128;
    long long v0 = (long long)blockIdx.z * 128;
    if (v0 < nvox) stage(0, v0);
    asm volatile("cp.async.commit_group;" ::: "memory");
    for (int it = 0; v0 < nvox; v0 += stride, ++it) {
        if (v0 + stride < nvox) stage((it + 1) & 1, v0 + stride);
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();
        const __half* xb = sx[it & 1];
        const __half* yb = sy[it & 1];
        uint32_t bf[4][2];
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
            uint32_t r[4];
            const int chunk = 2 * jj + (mi >> 1), krow = warp * 16 + (mi & 1) * 8 + mr;
            ldsm_x4_t(smem_u32(yb + ((size_t)chunk * 128 + krow) * 8), r);
            bf[2 * jj][0] = r[0]; bf[2 * jj][1] = r[1]; bf[2 * jj + 1][0] = r[2]; bf[2 * jj + 1][1] = r[3];
        }
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            uint32_t af[4];
            const int chunk = 2 * t + (mi & 1), krow = warp * 16 + (mi >> 1) * 8 + mr;
            ldsm_x4_t(smem_u32(xb + ((size_t)chunk * 128 + krow) * 8), af);
#pragma unroll
            for (int j = 0; j < 4; ++j) mma16816(acc[t][j], af, bf[j][0], bf[j][1]);
        }
        __syncthreads();
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    const int gq = lane >> 2, tq = lane & 3;
    const int taps = p.kz * p.kp * p.kr;
#pragma unroll
    for (int t = 0; t < 2; ++t)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int ci = cib * 32 + t * 16 + gq + (e >> 1) * 8, co = cob * 32 + j * 8 + tq * 2 + (e & 1);
                if (ci >= Cin_real || acc[t][j][e] == 0.f) continue;
                // master layouts: Conv3d [co][ci][tap]; ConvTranspose3d [ci][co][tap]
                const long long idx = master_transposed ? ((long long)ci * p.Cout + co) * taps + tap : ((long long)co * Cin_real + ci) * taps + tap;
                atomicAdd(dw + idx, acc[t][j][e]);
            }
}

// per-channel sum over interior rows of a PLCP tensor (bias gradient of convolutions that are not followed by a GroupNorm)
__global__ void __launch_bounds__(256) channel_sum_kernel(Geom g, const __half* __restrict__ t, int C, float* __restrict__ out, long long nrows) {
    __shared__ float s_acc[kBwdMaxC];
    for (int c = threadIdx.x; c < C; c += blockDim.x) s_acc[c] = 0.f;
    __syncthreads();
    const int nch = C >> 3;
    for (int k = 0; k < nch; ++k) {
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x; row < nrows; row += (long long)gridDim.x * blockDim.x) {
            if (decode_subrow(g, (uint32_t)(row % g.rows_per_sample)).kind != 0) continue;
            float f[8];
            unpack8(load8(t + plcp_offset(g, k, row)), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float v = acc[j];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if ((threadIdx.x & 31) == 0) atomicAdd(&s_acc[k * 8 + j], v);
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) atomicAdd(out + c, s_acc[c]);
}

// dst (+)= src over all rows of a C-channel PLCP tensor (gradient fan-in), optionally re-deriving the halo rows
__global__ void __launch_bounds__(256) plcp_accumulate_kernel(Geom g, const __half* __restrict__ src, __half* __restrict__ dst, int C,
                                                              int accumulate, long long nrows) {
    const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    const SubRow sr = decode_subrow(g, (uint32_t)(row % g.rows_per_sample));
    const long long srow = row + sr.src_delta;
    for (int k = 0; k < (C >> 3); ++k) {
        float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (sr.kind != 2) {
            unpack8(load8(src + plcp_offset(g, k, srow)), f);
            if (accumulate) {
                float o[8];
                unpack8(load8(dst + plcp_offset(g, k, srow)), o);
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] += o[j];
            }
        }
        store8(dst + plcp_offset(g, k, row), pack8(f));
    }
}

// =================================================================================================
// Linear attention backward (models.py:301-318, heads 1, d 32) on the raw 96-channel qkv tensor, per sample:
//   P = softmax_d(q), K = exp(k - kmax)/S, Cx = Ctx/S, att = s*Cx^T P, y = Wout att + b       (s = 32^-1/2)
// reduce pass (per voxel -> per sample):  dWout += dy att^T, db += dy, dCx[d][e] += s*P[d]*datt[e],  datt = Wout^T dy
// apply pass (per voxel):  dP = s*Cx datt, dq = P*(dP - <P,dP>);  dK = dCx v, dk = K*(dK - T[d]), T[d] = sum_e dCx*Cx;
//                          dv[e] = sum_d K[d] dCx[d][e]
// Thread = voxel (32-element vectors in registers, the 32x32 matrices in shared memory).
// =================================================================================================
__device__ __forceinline__ float ord2f_b(unsigned k) {
    unsigned u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    return __uint_as_float(u);
}

__device__ __forceinline__ void ldsm_x4_n(uint32_t addr, uint32_t* r) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ uint32_t pack2h(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack2h(uint32_t u) { return __half22float2(*reinterpret_cast<__half2*>(&u)); }

constexpr int kAbStr = 80;            // bytes per row of the 32-wide fp16 matrices (conflict-free ldmatrix)
constexpr float kDcxScale = 0.0625f;

// reduce pass: batches of 128 voxels; phase 1 (thread = voxel) leaves P, datt, att and dy in shared memory as fp16 chunk planes
// ([8-channel chunk][voxel][8], the layout wgrad_mma_kernel stages), phase 2 contracts the batch on the tensor cores
// (mma.sync m16n8k16, K = the 128 voxels): warps 0-1 dCx = P^T datt (16 d-rows each), warps 2-3 dWout = dy^T att (16-row
// c tiles), accumulators in registers across the block's batches; one set of global atomics per block at the end.
__global__ void __launch_bounds__(128) attn_bwd_reduce_kernel(AttnBwdParams p) {
    extern __shared__ __align__(16) unsigned char arsm_[];
    const int b = blockIdx.y, C = p.C, ncc = C >> 3, KS = C >> 4, wstr = C * 2 + 16;
    const Geom& g = p.g;
    unsigned char* s_cx = arsm_;                                   // Cx[d][e] fp16, rows of kAbStr bytes (B operand, k = d through .trans)
    unsigned char* s_wt = s_cx + 32 * kAbStr;                       // Wout^T[e][c] fp16, rows of wstr bytes
    __half* h_P = reinterpret_cast<__half*>(s_wt + 32 * wstr);      // [4][128][8]
    __half* h_datt = h_P + 4 * 128 * 8;
    __half* h_att = h_datt + 4 * 128 * 8;
    __half* h_dy = h_att + 4 * 128 * 8;                             // [C/8][128][8]
    const float* ctx = p.ctx + (long long)b * 33 * 32;
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) {
        const int d = i >> 5, e = i & 31;
        *reinterpret_cast<__half*>(s_cx + d * kAbStr + e * 2) = __float2half_rn(ctx[i] / ctx[32 * 32 + d]);
    }
    for (int i = threadIdx.x; i < C * 32; i += blockDim.x) {
        const int c = i >> 5, e = i & 31;
        *reinterpret_cast<__half*>(s_wt + e * wstr + c * 2) = __float2half_rn(p.wout[i]);
    }
    __syncthreads();
    const float scale = 0.17677669529663687f;   // 32^-0.5
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    const int mi = lane >> 3, mr = lane & 7, gq = lane >> 2, tq = lane & 3;
    float acc[4][4][4];                          // [row tile][e 8-tile][frag]: warps 0-1 use tile 0 (dCx), warps 2-3 up to 4 c tiles
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[a][j][e] = 0.f;
    float a_db = 0.f;
    const int nct = C >> 4;                      // 16-row c tiles of dWout
    const uint32_t r0 = blockIdx.x * p.rows_per_block;
    uint32_t r1 = r0 + p.rows_per_block;
    if (r1 > (uint32_t)g.rows_per_sample) r1 = g.rows_per_sample;
    const long long srow0 = (long long)b * g.rows_per_sample;
    const long long cs = g.chunk_stride * 8;
    const __half* qb = p.qkv + plcp_offset(g, 0, srow0);
    const __half* dyb = p.dy + plcp_offset(g, 0, srow0);
    for (uint32_t rem0 = r0; rem0 < r1; rem0 += 128) {
        __syncthreads();
        // phase 1: the warp's 32 rows as two m16 tiles; P, datt = s dy Wout, att = s P Cx in accumulator layout, written to the
        // chunk planes as half2 (rows that are not interior voxels contribute zeros)
        {
            const uint32_t base = rem0 + warp * 32;
            const uint32_t rem = base + lane;
            const unsigned vmask = __ballot_sync(0xffffffffu, rem < r1 && decode_subrow(g, rem).kind == 0);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                const int lA = warp * 32 + mt * 16 + gq, lB = lA + 8;               // rows within the 128-row batch
                const uint32_t rA = rem0 + lA, rB = rem0 + lB;
                const bool vA = (vmask >> (mt * 16 + gq)) & 1u, vB = (vmask >> (mt * 16 + gq + 8)) & 1u;
                const long long oA = (long long)rA * 8 + 2 * tq, oB = (long long)rB * 8 + 2 * tq;
                float P[4][4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 a = unpack2h(vA ? __ldg(reinterpret_cast<const unsigned*>(qb + oA + j * cs)) : 0u);
                    const float2 c = unpack2h(vB ? __ldg(reinterpret_cast<const unsigned*>(qb + oB + j * cs)) : 0u);
                    P[j][0] = a.x; P[j][1] = a.y; P[j][2] = c.x; P[j][3] = c.y;
                }
                float m0 = P[0][0], m1 = P[0][2];
#pragma unroll
                for (int j = 0; j < 4; ++j) { m0 = fmaxf(m0, fmaxf(P[j][0], P[j][1])); m1 = fmaxf(m1, fmaxf(P[j][2], P[j][3])); }
                m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
                m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
                float s0 = 0.f, s1 = 0.f;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    P[j][0] = __expf(P[j][0] - m0); P[j][1] = __expf(P[j][1] - m0); s0 += P[j][0] + P[j][1];
                    P[j][2] = __expf(P[j][2] - m1); P[j][3] = __expf(P[j][3] - m1); s1 += P[j][2] + P[j][3];
                }
                s0 += __shfl_xor_sync(0xffffffffu, s0, 1); s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
                s1 += __shfl_xor_sync(0xffffffffu, s1, 1); s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
                const float i0 = vA ? 1.f / s0 : 0.f, i1 = vB ? 1.f / s1 : 0.f;          // P = 0 on rows that do not count
                uint32_t pa[2][4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t hA = pack2h(P[j][0] * i0, P[j][1] * i0), hB = pack2h(P[j][2] * i1, P[j][3] * i1);
                    *reinterpret_cast<unsigned*>(h_P + ((size_t)j * 128 + lA) * 8 + 2 * tq) = hA;
                    *reinterpret_cast<unsigned*>(h_P + ((size_t)j * 128 + lB) * 8 + 2 * tq) = hB;
                    pa[j >> 1][(j & 1) * 2] = hA; pa[j >> 1][(j & 1) * 2 + 1] = hB;       // A fragment of P for att
                }
                // datt = s dy Wout  (dy fragments from global, copied to the dy chunk planes on the way)
                float dt[4][4];
#pragma unroll
                for (int j = 0; j < 4; ++j) dt[j][0] = dt[j][1] = dt[j][2] = dt[j][3] = 0.f;
                for (int sp = 0; sp < KS; sp += 2) {
                    uint32_t a[2][4];
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int s = sp + u;
                        const bool have = s < KS;
                        a[u][0] = (have && vA) ? __ldg(reinterpret_cast<const unsigned*>(dyb + oA + (2 * s) * cs)) : 0u;
                        a[u][1] = (have && vB) ? __ldg(reinterpret_cast<const unsigned*>(dyb + oB + (2 * s) * cs)) : 0u;
                        a[u][2] = (have && vA) ? __ldg(reinterpret_cast<const unsigned*>(dyb + oA + (2 * s + 1) * cs)) : 0u;
                        a[u][3] = (have && vB) ? __ldg(reinterpret_cast<const unsigned*>(dyb + oB + (2 * s + 1) * cs)) : 0u;
                        if (have) {
                            *reinterpret_cast<unsigned*>(h_dy + ((size_t)(2 * s) * 128 + lA) * 8 + 2 * tq) = a[u][0];
                            *reinterpret_cast<unsigned*>(h_dy + ((size_t)(2 * s) * 128 + lB) * 8 + 2 * tq) = a[u][1];
                            *reinterpret_cast<unsigned*>(h_dy + ((size_t)(2 * s + 1) * 128 + lA) * 8 + 2 * tq) = a[u][2];
                            *reinterpret_cast<unsigned*>(h_dy + ((size_t)(2 * s + 1) * 128 + lB) * 8 + 2 * tq) = a[u][3];
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        uint32_t bw[4];
                        ldsm_x4_n(smem_u32(s_wt) + (j * 8 + (lane & 7)) * wstr + sp * 32 + (lane >> 3) * 16, bw);
                        mma16816(dt[j], a[0], bw[0], bw[1]);
                        if (sp + 1 < KS) mma16816(dt[j], a[1], bw[2], bw[3]);
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    *reinterpret_cast<unsigned*>(h_datt + ((size_t)j * 128 + lA) * 8 + 2 * tq) = pack2h(dt[j][0] * scale, dt[j][1] * scale);
                    *reinterpret_cast<unsigned*>(h_datt + ((size_t)j * 128 + lB) * 8 + 2 * tq) = pack2h(dt[j][2] * scale, dt[j][3] * scale);
                }
                // att = s P Cx : A = P (k = d), B[k = d][n = e] = Cx[d][e] through ldmatrix.trans
#pragma unroll
                for (int j = 0; j < 4; ++j) dt[j][0] = dt[j][1] = dt[j][2] = dt[j][3] = 0.f;
#pragma unroll
                for (int ks = 0; ks < 2; ++ks)
#pragma unroll
                    for (int jj = 0; jj < 2; ++jj) {
                        uint32_t bt[4];
                        ldsm_x4_t(smem_u32(s_cx) + (ks * 16 + (mi & 1) * 8 + mr) * kAbStr + (2 * jj + (mi >> 1)) * 16, bt);
                        mma16816(dt[2 * jj], pa[ks], bt[0], bt[1]);
                        mma16816(dt[2 * jj + 1], pa[ks], bt[2], bt[3]);
                    }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    *reinterpret_cast<unsigned*>(h_att + ((size_t)j * 128 + lA) * 8 + 2 * tq) = pack2h(dt[j][0] * scale, dt[j][1] * scale);
                    *reinterpret_cast<unsigned*>(h_att + ((size_t)j * 128 + lB) * 8 + 2 * tq) = pack2h(dt[j][2] * scale, dt[j][3] * scale);
                }
            }
        }
        __syncthreads();
        // phase 2: A = (P | dy)^T through ldmatrix.trans on the chunk planes, B = (datt | att) likewise (as wgrad_mma_kernel)
        const __half* Am = warp < 2 ? h_P : h_dy;
        const __half* Bm = warp < 2 ? h_datt : h_att;
        const int ntile = warp < 2 ? 1 : (nct - (warp - 2) + 1) / 2;
#pragma unroll 2
        for (int k16 = 0; k16 < 8; ++k16) {
            uint32_t bf[4][2];
#pragma unroll
            for (int jj = 0; jj < 2; ++jj) {
                uint32_t r[4];
                const int chunk = 2 * jj + (mi >> 1), krow = k16 * 16 + (mi & 1) * 8 + mr;
                ldsm_x4_t(smem_u32(Bm + ((size_t)chunk * 128 + krow) * 8), r);
                bf[2 * jj][0] = r[0]; bf[2 * jj][1] = r[1]; bf[2 * jj + 1][0] = r[2]; bf[2 * jj + 1][1] = r[3];
            }
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                if (a >= ntile) break;
                const int tile = warp < 2 ? warp : (warp - 2) + 2 * a;       // 16-row tile of d (dCx) or c (dWout)
                uint32_t af[4];
                const int chunk = 2 * tile + (mi & 1), krow = k16 * 16 + (mi >> 1) * 8 + mr;
                ldsm_x4_t(smem_u32(Am + ((size_t)chunk * 128 + krow) * 8), af);
#pragma unroll
                for (int j = 0; j < 4; ++j) mma16816(acc[a][j], af, bf[j][0], bf[j][1]);
            }
        }
        if (t < C) {
            const __half* col = h_dy + ((size_t)(t >> 3) * 128) * 8 + (t & 7);
            float sdb = 0.f;
            for (int n = 0; n < 128; ++n) sdb += __half2float(col[n * 8]);
            a_db += sdb;
        }
    }
    const int ntile = warp < 2 ? 1 : (nct - (warp - 2) + 1) / 2;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        if (a >= ntile) break;
        const int tile = warp < 2 ? warp : (warp - 2) + 2 * a;
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int r = tile * 16 + gq + (e >> 1) * 8, c = j * 8 + tq * 2 + (e & 1);
                if (warp < 2) atomicAdd(p.dcx + (long long)b * 1024 + r * 32 + c, acc[a][j][e]);
                else atomicAdd(p.dwout + r * 32 + c, acc[a][j][e]);
            }
    }
    if (t < C) atomicAdd(p.dbout + t, a_db);
}

__global__ void __launch_bounds__(128) attn_bwd_apply_kernel(AttnBwdParams p) {
    // matrices every thread reads at the SAME address (broadcast): unpadded and 16-byte aligned so the e-loops read float4
    __shared__ __align__(16) float s_cx[32][32];      // Cx[d][e]
    __shared__ __align__(16) float s_dcx[32][32];     // dCx[d][e]
    __shared__ __align__(16) float s_wout[kBwdMaxC][32];
    __shared__ float s_km[32], s_sinv[32], s_T[32];
    const int b = blockIdx.y, C = p.C;
    const Geom& g = p.g;
    const float* ctx = p.ctx + (long long)b * 33 * 32;
    for (int i = threadIdx.x; i < 32; i += blockDim.x) { s_km[i] = ord2f_b(p.kmax[b * 32 + i]); s_sinv[i] = 1.f / ctx[32 * 32 + i]; }
    __syncthreads();
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) {
        const int d = i >> 5, e = i & 31;
        s_cx[d][e] = ctx[i] * s_sinv[d];
        s_dcx[d][e] = p.dcx[(long long)b * 1024 + i];
    }
    for (int i = threadIdx.x; i < C * 32; i += blockDim.x) s_wout[i >> 5][i & 31] = p.wout[i];
    __syncthreads();
    for (int d = threadIdx.x; d < 32; d += blockDim.x) {
        float t = 0.f;
        for (int e = 0; e < 32; ++e) t = fmaf(s_dcx[d][e], s_cx[d][e], t);
        s_T[d] = t;
    }
    __syncthreads();
    const float scale = 0.17677669529663687f;   // 32^-0.5
    const uint32_t r0 = blockIdx.x * p.rows_per_block;
    uint32_t r1 = r0 + p.rows_per_block;
    if (r1 > (uint32_t)g.rows_per_sample) r1 = g.rows_per_sample;
    for (uint32_t rem = r0 + threadIdx.x; rem < r1; rem += blockDim.x) {
        const bool interior = decode_subrow(g, rem).kind == 0;
        const long long row = (long long)b * g.rows_per_sample + rem;
        if (!interior) {
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);     // rows a 1x1x1 data-gradient conv reads only for itself
            for (int k = 0; k < 12; ++k) *reinterpret_cast<uint4*>(p.dqkv + plcp_offset(g, k, row)) = z;
            continue;
        }
        float P[32], datt[32], out[8];
        {
            float q[32];
            for (int k = 0; k < 4; ++k) unpack8(load8(p.qkv + plcp_offset(g, k, row)), q + k * 8);
            float mx = q[0];
#pragma unroll
            for (int d = 1; d < 32; ++d) mx = fmaxf(mx, q[d]);
            float sum = 0.f;
#pragma unroll
            for (int d = 0; d < 32; ++d) { P[d] = __expf(q[d] - mx); sum += P[d]; }
            const float inv = 1.f / sum;
#pragma unroll
            for (int d = 0; d < 32; ++d) P[d] *= inv;
        }
#pragma unroll
        for (int e = 0; e < 32; ++e) datt[e] = 0.f;
        for (int k = 0; k < (C >> 3); ++k) {
            float dy[8];
            unpack8(load8(p.dy + plcp_offset(g, k, row)), dy);
#pragma unroll
            for (int j = 0; j < 8; ++j)
#pragma unroll
                for (int e4 = 0; e4 < 8; ++e4) {
                    const float4 w = *reinterpret_cast<const float4*>(&s_wout[k * 8 + j][e4 * 4]);
                    datt[e4 * 4] = fmaf(w.x, dy[j], datt[e4 * 4]); datt[e4 * 4 + 1] = fmaf(w.y, dy[j], datt[e4 * 4 + 1]);
                    datt[e4 * 4 + 2] = fmaf(w.z, dy[j], datt[e4 * 4 + 2]); datt[e4 * 4 + 3] = fmaf(w.w, dy[j], datt[e4 * 4 + 3]);
                }
        }
        float kv[32], vv[32];
        for (int k = 0; k < 4; ++k) unpack8(load8(p.qkv + plcp_offset(g, 4 + k, row)), kv + k * 8);
        for (int k = 0; k < 4; ++k) unpack8(load8(p.qkv + plcp_offset(g, 8 + k, row)), vv + k * 8);
        // dq = P*(dP - <P,dP>),  dP = s * Cx datt
        float dP[32], dot = 0.f;
#pragma unroll
        for (int d = 0; d < 32; ++d) {
            float t = 0.f;
#pragma unroll
            for (int e4 = 0; e4 < 8; ++e4) {
                const float4 w = *reinterpret_cast<const float4*>(&s_cx[d][e4 * 4]);
                t = fmaf(w.x, datt[e4 * 4], t); t = fmaf(w.y, datt[e4 * 4 + 1], t); t = fmaf(w.z, datt[e4 * 4 + 2], t); t = fmaf(w.w, datt[e4 * 4 + 3], t);
            }
            dP[d] = scale * t;
            dot = fmaf(P[d], dP[d], dot);
        }
        for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int j = 0; j < 8; ++j) out[j] = P[k * 8 + j] * (dP[k * 8 + j] - dot);
            store8(p.dqkv + plcp_offset(g, k, row), pack8(out));
        }
        // dk = K*(dCx v - T)
        for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int d = k * 8 + j;
                float t = 0.f;
#pragma unroll
                for (int e4 = 0; e4 < 8; ++e4) {
                    const float4 w = *reinterpret_cast<const float4*>(&s_dcx[d][e4 * 4]);
                    t = fmaf(w.x, vv[e4 * 4], t); t = fmaf(w.y, vv[e4 * 4 + 1], t); t = fmaf(w.z, vv[e4 * 4 + 2], t); t = fmaf(w.w, vv[e4 * 4 + 3], t);
                }
                const float K = __expf(kv[d] - s_km[d]) * s_sinv[d];
                out[j] = K * (t - s_T[d]);
                kv[d] = K;                          // keep K for dv
            }
            store8(p.dqkv + plcp_offset(g, 4 + k, row), pack8(out));
        }
        // dv[e] = sum_d K[d] dCx[d][e]   (vv is dead: reuse it as the accumulator; d outer so the e-loop reads float4)
#pragma unroll
        for (int e = 0; e < 32; ++e) vv[e] = 0.f;
#pragma unroll
        for (int d = 0; d < 32; ++d)
#pragma unroll
            for (int e4 = 0; e4 < 8; ++e4) {
                const float4 w = *reinterpret_cast<const float4*>(&s_dcx[d][e4 * 4]);
                vv[e4 * 4] = fmaf(kv[d], w.x, vv[e4 * 4]); vv[e4 * 4 + 1] = fmaf(kv[d], w.y, vv[e4 * 4 + 1]);
                vv[e4 * 4 + 2] = fmaf(kv[d], w.z, vv[e4 * 4 + 2]); vv[e4 * 4 + 3] = fmaf(kv[d], w.w, vv[e4 * 4 + 3]);
            }
        for (int k = 0; k < 4; ++k) store8(p.dqkv + plcp_offset(g, 8 + k, row), pack8(vv + k * 8));
    }
}

// ---------------------------------------------------------------------------------------------------------
// apply pass on the tensor cores (C a multiple of 16).  A warp owns 32 consecutive rows (two m16 tiles); the four per-voxel
// 32x32 matvecs of the SIMT form become mma.sync m16n8k16 GEMMs chained through registers, the accumulator fragment of one
// being repacked (fp16) as the A fragment of the next, as fa_out_kernel does in the forward pass:
//   datt = dy Wout (K = C)  ->  dP = s datt Cx^T  ->  dq = P (dP - <P,dP>)          P = softmax_d(q), in accumulator layout
//   dKpre = v dCx^T  ->  dk = K (dKpre - T)   ;   dv = K dCx                          K = exp(k - kmax) / S
// q / k are read from global memory directly in the accumulator layout, dy / v directly as A fragments (in PLCP one 8-row x
// 8-channel fragment is 128 contiguous bytes); the small matrices sit in shared memory as fp16 B operands (dCx scaled by
// 1/16 to stay well inside fp16 range).  Halo / guard rows of dqkv are written as zeros (whole 32-byte sectors).
// ---------------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(128) attn_bwd_apply_mma_kernel(AttnBwdParams p) {
    extern __shared__ __align__(16) unsigned char abs_[];
    const int b = blockIdx.y, C = p.C, KS = C >> 4, wstr = C * 2 + 16;
    const Geom& g = p.g;
    unsigned char* s_cx = abs_;                          // Cx[d][e] fp16, rows of kAbStr bytes
    unsigned char* s_dcx = s_cx + 32 * kAbStr;           // dCx[d][e] / 16
    unsigned char* s_wt = s_dcx + 32 * kAbStr;           // Wout^T[e][c], rows of wstr bytes
    float* s_km = reinterpret_cast<float*>(s_wt + 32 * wstr);
    float* s_sinv = s_km + 32;
    float* s_T = s_sinv + 32;
    const float* ctx = p.ctx + (long long)b * 33 * 32;
    for (int i = threadIdx.x; i < 32; i += blockDim.x) { s_km[i] = ord2f_b(p.kmax[b * 32 + i]); s_sinv[i] = 1.f / ctx[32 * 32 + i]; }
    __syncthreads();
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) {
        const int d = i >> 5, e = i & 31;
        *reinterpret_cast<__half*>(s_cx + d * kAbStr + e * 2) = __float2half_rn(ctx[i] * s_sinv[d]);
        *reinterpret_cast<__half*>(s_dcx + d * kAbStr + e * 2) = __float2half_rn(p.dcx[(long long)b * 1024 + i] * kDcxScale);
    }
    for (int i = threadIdx.x; i < C * 32; i += blockDim.x) {
        const int c = i >> 5, e = i & 31;
        *reinterpret_cast<__half*>(s_wt + e * wstr + c * 2) = __float2half_rn(p.wout[i]);
    }
    if (threadIdx.x < 32) {                               // T[d] = sum_e dCx[d][e] Cx[d][e]  (fp32, from the fp32 sources)
        const int d = threadIdx.x;
        float t = 0.f;
        for (int e = 0; e < 32; ++e) t = fmaf(p.dcx[(long long)b * 1024 + d * 32 + e], ctx[d * 32 + e] * s_sinv[d], t);
        s_T[d] = t;
    }
    __syncthreads();
    const float scale = 0.17677669529663687f;   // 32^-0.5
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, gq = lane >> 2, tq = lane & 3;
    const uint32_t r0 = blockIdx.x * p.rows_per_block;
    uint32_t r1 = r0 + p.rows_per_block;
    if (r1 > (uint32_t)g.rows_per_sample) r1 = g.rows_per_sample;
    const long long srow0 = (long long)b * g.rows_per_sample;
    const long long cs = g.chunk_stride * 8;
    const __half* qb = p.qkv + plcp_offset(g, 0, srow0);
    const __half* dyb = p.dy + plcp_offset(g, 0, srow0);
    __half* ob = p.dqkv + plcp_offset(g, 0, srow0);
    for (uint32_t base = r0 + warp * 32; base < r1; base += 4 * 32) {
        const uint32_t rem = base + lane;
        const bool inb = rem < r1;
        const unsigned vmask = __ballot_sync(0xffffffffu, inb && decode_subrow(g, rem).kind == 0);
        const unsigned imask = __ballot_sync(0xffffffffu, inb);
        if (vmask == 0u) {          // nothing to differentiate in this group: zero rows
            if (inb) {
                const uint4 z = make_uint4(0u, 0u, 0u, 0u);
                for (int k = 0; k < 12; ++k) *reinterpret_cast<uint4*>(ob + (long long)rem * 8 + k * cs) = z;
            }
            continue;
        }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            const uint32_t rA = base + mt * 16 + gq, rB = rA + 8;         // this lane's two rows of the m tile
            const bool ldA = rA < r1, ldB = rB < r1;
            const bool vA = (vmask >> (mt * 16 + gq)) & 1u, vB = (vmask >> (mt * 16 + gq + 8)) & 1u;
            const bool sA = (imask >> (mt * 16 + gq)) & 1u, sB = (imask >> (mt * 16 + gq + 8)) & 1u;
            const long long oA = (long long)rA * 8 + 2 * tq, oB = (long long)rB * 8 + 2 * tq;
            // ---- P = softmax_d(q) in accumulator layout: P[j][0,1] = row A cols 8j+2t,+1 ; [2,3] = row B
            float P[4][4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 a = unpack2h(ldA ? __ldg(reinterpret_cast<const unsigned*>(qb + oA + j * cs)) : 0u);
                const float2 c = unpack2h(ldB ? __ldg(reinterpret_cast<const unsigned*>(qb + oB + j * cs)) : 0u);
                P[j][0] = a.x; P[j][1] = a.y; P[j][2] = c.x; P[j][3] = c.y;
            }
            float m0 = P[0][0], m1 = P[0][2];
#pragma unroll
            for (int j = 0; j < 4; ++j) { m0 = fmaxf(m0, fmaxf(P[j][0], P[j][1])); m1 = fmaxf(m1, fmaxf(P[j][2], P[j][3])); }
            m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
            m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                P[j][0] = __expf(P[j][0] - m0); P[j][1] = __expf(P[j][1] - m0); s0 += P[j][0] + P[j][1];
                P[j][2] = __expf(P[j][2] - m1); P[j][3] = __expf(P[j][3] - m1); s1 += P[j][2] + P[j][3];
            }
            s0 += __shfl_xor_sync(0xffffffffu, s0, 1); s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
            s1 += __shfl_xor_sync(0xffffffffu, s1, 1); s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
            const float i0 = 1.f / s0, i1 = 1.f / s1;
#pragma unroll
            for (int j = 0; j < 4; ++j) { P[j][0] *= i0; P[j][1] *= i0; P[j][2] *= i1; P[j][3] *= i1; }
            // ---- datt = dy Wout : A = dy fragments from global, B = Wout^T rows (e) x k (c)
            float acc[4][4];
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
            for (int sp = 0; sp < KS; sp += 2) {            // two k16 steps (32 channels of dy) per ldmatrix.x4
                uint32_t a[2][4];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int s = sp + u;
                    const bool have = s < KS;
                    a[u][0] = (have && ldA) ? __ldg(reinterpret_cast<const unsigned*>(dyb + oA + (2 * s) * cs)) : 0u;
                    a[u][1] = (have && ldB) ? __ldg(reinterpret_cast<const unsigned*>(dyb + oB + (2 * s) * cs)) : 0u;
                    a[u][2] = (have && ldA) ? __ldg(reinterpret_cast<const unsigned*>(dyb + oA + (2 * s + 1) * cs)) : 0u;
                    a[u][3] = (have && ldB) ? __ldg(reinterpret_cast<const unsigned*>(dyb + oB + (2 * s + 1) * cs)) : 0u;
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint32_t bw[4];
                    ldsm_x4_n(smem_u32(s_wt) + (j * 8 + (lane & 7)) * wstr + sp * 32 + (lane >> 3) * 16, bw);
                    mma16816(acc[j], a[0], bw[0], bw[1]);
                    if (sp + 1 < KS) mma16816(acc[j], a[1], bw[2], bw[3]);
                }
            }
            // ---- dP = s datt Cx^T ; dq = P (dP - <P, dP>)
            uint32_t af[2][4];
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
                af[ks][0] = pack2h(acc[2 * ks][0], acc[2 * ks][1]); af[ks][1] = pack2h(acc[2 * ks][2], acc[2 * ks][3]);
                af[ks][2] = pack2h(acc[2 * ks + 1][0], acc[2 * ks + 1][1]); af[ks][3] = pack2h(acc[2 * ks + 1][2], acc[2 * ks + 1][3]);
            }
            float d0 = 0.f, d1 = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint32_t bc[4];
                ldsm_x4_n(smem_u32(s_cx) + (j * 8 + (lane & 7)) * kAbStr + (lane >> 3) * 16, bc);
                acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
                mma16816(acc[j], af[0], bc[0], bc[1]);
                mma16816(acc[j], af[1], bc[2], bc[3]);
#pragma unroll
                for (int e = 0; e < 4; ++e) acc[j][e] *= scale;               // dP = 32^-1/2 Cx datt
                d0 = fmaf(P[j][0], acc[j][0], fmaf(P[j][1], acc[j][1], d0));
                d1 = fmaf(P[j][2], acc[j][2], fmaf(P[j][3], acc[j][3], d1));
            }
            d0 += __shfl_xor_sync(0xffffffffu, d0, 1); d0 += __shfl_xor_sync(0xffffffffu, d0, 2);
            d1 += __shfl_xor_sync(0xffffffffu, d1, 1); d1 += __shfl_xor_sync(0xffffffffu, d1, 2);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float q0 = vA ? P[j][0] * (acc[j][0] - d0) : 0.f, q1 = vA ? P[j][1] * (acc[j][1] - d0) : 0.f;
                const float q2 = vB ? P[j][2] * (acc[j][2] - d1) : 0.f, q3 = vB ? P[j][3] * (acc[j][3] - d1) : 0.f;
                if (sA) *reinterpret_cast<unsigned*>(ob + oA + j * cs) = pack2h(q0, q1);
                if (sB) *reinterpret_cast<unsigned*>(ob + oB + j * cs) = pack2h(q2, q3);
            }
            // ---- K = exp(k - kmax) / S (accumulator layout, reusing P) ; v as A fragments
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 a = unpack2h(ldA ? __ldg(reinterpret_cast<const unsigned*>(qb + oA + (4 + j) * cs)) : 0u);
                const float2 c = unpack2h(ldB ? __ldg(reinterpret_cast<const unsigned*>(qb + oB + (4 + j) * cs)) : 0u);
                const int dcol = j * 8 + 2 * tq;
                const float k0 = s_km[dcol], k1 = s_km[dcol + 1], v0 = s_sinv[dcol], v1 = s_sinv[dcol + 1];
                P[j][0] = __expf(a.x - k0) * v0; P[j][1] = __expf(a.y - k1) * v1;
                P[j][2] = __expf(c.x - k0) * v0; P[j][3] = __expf(c.y - k1) * v1;
            }
            uint32_t av[2][4];
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
                av[ks][0] = ldA ? __ldg(reinterpret_cast<const unsigned*>(qb + oA + (8 + 2 * ks) * cs)) : 0u;
                av[ks][1] = ldB ? __ldg(reinterpret_cast<const unsigned*>(qb + oB + (8 + 2 * ks) * cs)) : 0u;
                av[ks][2] = ldA ? __ldg(reinterpret_cast<const unsigned*>(qb + oA + (8 + 2 * ks + 1) * cs)) : 0u;
                av[ks][3] = ldB ? __ldg(reinterpret_cast<const unsigned*>(qb + oB + (8 + 2 * ks + 1) * cs)) : 0u;
            }
            // ---- dk = K (v dCx^T - T)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint32_t bd[4];
                ldsm_x4_n(smem_u32(s_dcx) + (j * 8 + (lane & 7)) * kAbStr + (lane >> 3) * 16, bd);
                acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
                mma16816(acc[j], av[0], bd[0], bd[1]);
                mma16816(acc[j], av[1], bd[2], bd[3]);
                const int dcol = j * 8 + 2 * tq;
                const float t0 = s_T[dcol], t1 = s_T[dcol + 1];
                const float inv = 1.f / kDcxScale;
                const float q0 = vA ? P[j][0] * (acc[j][0] * inv - t0) : 0.f, q1 = vA ? P[j][1] * (acc[j][1] * inv - t1) : 0.f;
                const float q2 = vB ? P[j][2] * (acc[j][2] * inv - t0) : 0.f, q3 = vB ? P[j][3] * (acc[j][3] * inv - t1) : 0.f;
                if (sA) *reinterpret_cast<unsigned*>(ob + oA + (4 + j) * cs) = pack2h(q0, q1);
                if (sB) *reinterpret_cast<unsigned*>(ob + oB + (4 + j) * cs) = pack2h(q2, q3);
            }
            // ---- dv = K dCx : A = K repacked, B[k = d][n = e] = dCx[d][e] through ldmatrix.trans
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
                af[ks][0] = pack2h(P[2 * ks][0], P[2 * ks][1]); af[ks][1] = pack2h(P[2 * ks][2], P[2 * ks][3]);
                af[ks][2] = pack2h(P[2 * ks + 1][0], P[2 * ks + 1][1]); af[ks][3] = pack2h(P[2 * ks + 1][2], P[2 * ks + 1][3]);
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
#pragma unroll
            for (int ks = 0; ks < 2; ++ks)
#pragma unroll
                for (int jj = 0; jj < 2; ++jj) {
                    uint32_t bt[4];
                    // matrices: (k 0-7, n tile 2jj) (k 8-15, n tile 2jj) (k 0-7, n tile 2jj+1) (k 8-15, n tile 2jj+1)
                    const int mi = lane >> 3, mr = lane & 7;
                    ldsm_x4_t(smem_u32(s_dcx) + (ks * 16 + (mi & 1) * 8 + mr) * kAbStr + (2 * jj + (mi >> 1)) * 16, bt);
                    mma16816(acc[2 * jj], af[ks], bt[0], bt[1]);
                    mma16816(acc[2 * jj + 1], af[ks], bt[2], bt[3]);
                }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float inv = 1.f / kDcxScale;
                const float q0 = vA ? acc[j][0] * inv : 0.f, q1 = vA ? acc[j][1] * inv : 0.f;
                const float q2 = vB ? acc[j][2] * inv : 0.f, q3 = vB ? acc[j][3] * inv : 0.f;
                if (sA) *reinterpret_cast<unsigned*>(ob + oA + (8 + j) * cs) = pack2h(q0, q1);
                if (sB) *reinterpret_cast<unsigned*>(ob + oB + (8 + j) * cs) = pack2h(q2, q3);
            }
        }
    }
}

// =================================================================================================
// Conditioning MLPs backward (models.py:575-608, 176-178): one block per sample recomputes the forward (as embed_kernel),
// back-propagates the gradient of the additive conditioning vectors through the per-block projections, the SiLU and both
// 3-layer MLPs, and leaves (input, output-gradient) pairs of every Linear in a per-sample record; linear_wgrad_kernel then
// contracts the batch:  dW[o][i] += sum_b dY[b][o] X[b][i],  db[o] += sum_b dY[b][o].
// Record layout (floats): see EmbedBwdLayout in cd_kernels.cuh.
// =================================================================================================
__device__ void dense_fwd(const float* __restrict__ W, const float* __restrict__ bias, const float* in, float* pre, float* out, int n_in,
                          int n_out, int act) {
    for (int o = threadIdx.x; o < n_out; o += blockDim.x) {
        float a = bias[o];
        for (int i = 0; i < n_in; ++i) a = fmaf(W[(long long)o * n_in + i], in[i], a);
        pre[o] = a;
        out[o] = act ? gelu_erf(a) : a;
    }
    __syncthreads();
}
// din[i] = sum_o W[o][i] * dpre[o]
__device__ void dense_bwd_in(const float* __restrict__ W, const float* dpre, float* din, int n_in, int n_out) {
    for (int i = threadIdx.x; i < n_in; i += blockDim.x) {
        float a = 0.f;
        for (int o = 0; o < n_out; ++o) a = fmaf(W[(long long)o * n_in + i], dpre[o], a);
        din[i] = a;
    }
    __syncthreads();
}
__device__ __forceinline__ float gelu_grad(float x) {
    const float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752f));
    return cdf + x * 0.3989422804014327f * __expf(-0.5f * x * x);
}

__global__ void __launch_bounds__(256) embed_backward_kernel(EmbedParams p, EmbedBwdLayout L, const float* __restrict__ sigma,
                                                             const float* __restrict__ cond, const float* __restrict__ daddvec,
                                                             float* __restrict__ rec) {
    __shared__ float t0[1], t1p[128], t1[128], t2p[128], t2[128], t3[128];
    __shared__ float c0[256], c1p[256], c1[256], c2p[128], c2[128], c3[128];
    __shared__ float cpre[256], cvec[256], dcv[256], da[256], db2[256];
    const int b = blockIdx.x, half = p.cond_dim / 2, h2 = half / 2;
    float* r = rec + (long long)b * L.stride;
    if (threadIdx.x == 0) {
        const float s = sigma[b];
        t0[0] = (p.time_mode == 0) ? s / sqrtf(1.f + s * s) : (p.time_mode == 1 ? 0.5f * logf(s) : s);
    }
    for (int i = threadIdx.x; i < p.cond_size; i += blockDim.x) c0[i] = cond[(long long)b * p.cond_size + i];
    __syncthreads();
    dense_fwd(p.tw[0], p.tb[0], t0, t1p, t1, 1, h2, 1);
    dense_fwd(p.tw[1], p.tb[1], t1, t2p, t2, h2, half, 1);
    dense_fwd(p.tw[2], p.tb[2], t2, cpre, t3, half, half, 0);
    dense_fwd(p.cw[0], p.cb[0], c0, c1p, c1, p.cond_size, p.cond_hidden, 1);
    dense_fwd(p.cw[1], p.cb[1], c1, c2p, c2, p.cond_hidden, half, 1);
    dense_fwd(p.cw[2], p.cb[2], c2, cpre + half, c3, half, half, 0);
    for (int i = threadIdx.x; i < p.cond_dim; i += blockDim.x) { cvec[i] = silu_f(cpre[i]); dcv[i] = 0.f; }
    __syncthreads();
    // projections: dY = daddvec slice, X = cvec;  d cvec += W^T dY
    for (int i = threadIdx.x; i < p.cond_dim; i += blockDim.x) {
        float a = 0.f;
        for (int k = 0; k < p.n_proj; ++k) {
            const float* W = p.pw[k];
            const float* dy = daddvec + (long long)b * p.addvec_stride + p.poff[k];
            for (int o = 0; o < p.pC[k]; ++o) a = fmaf(W[(long long)o * p.cond_dim + i], dy[o], a);
        }
        const float z = cpre[i], sg = sigmoid_f(z);
        dcv[i] = a * sg * (1.f + z * (1.f - sg));        // gradient wrt conditions (pre-SiLU) = outputs of the two MLPs
        r[L.cvec + i] = cvec[i];
    }
    __syncthreads();
    // time MLP backward: layer 3 (no activation) <- dcv[0:half]
    for (int i = threadIdx.x; i < half; i += blockDim.x) { r[L.dt3 + i] = dcv[i]; r[L.t2 + i] = t2[i]; r[L.dc3 + i] = dcv[half + i]; r[L.c2 + i] = c2[i]; }
    dense_bwd_in(p.tw[2], dcv, da, half, half);
    for (int i = threadIdx.x; i < half; i += blockDim.x) { da[i] *= gelu_grad(t2p[i]); r[L.dt2 + i] = da[i]; }
    for (int i = threadIdx.x; i < h2; i += blockDim.x) r[L.t1 + i] = t1[i];
    __syncthreads();
    dense_bwd_in(p.tw[1], da, db2, h2, half);
    for (int i = threadIdx.x; i < h2; i += blockDim.x) r[L.dt1 + i] = db2[i] * gelu_grad(t1p[i]);
    if (threadIdx.x == 0) r[L.t0] = t0[0];
    // cond MLP backward
    dense_bwd_in(p.cw[2], dcv + half, da, half, half);
    for (int i = threadIdx.x; i < half; i += blockDim.x) { da[i] *= gelu_grad(c2p[i]); r[L.dc2 + i] = da[i]; }
    for (int i = threadIdx.x; i < p.cond_hidden; i += blockDim.x) r[L.c1 + i] = c1[i];
    __syncthreads();
    dense_bwd_in(p.cw[1], da, db2, p.cond_hidden, half);
    for (int i = threadIdx.x; i < p.cond_hidden; i += blockDim.x) r[L.dc1 + i] = db2[i] * gelu_grad(c1p[i]);
    for (int i = threadIdx.x; i < p.cond_size; i += blockDim.x) r[L.c0 + i] = c0[i];
}

// dW[o][i] += sum_b dY[b*sy + o] * X[b*sx + i];  db[o] += sum_b dY   (thread = weight element; i == n_in -> bias)
__global__ void linear_wgrad_kernel(const float* __restrict__ dY, long long sy, const float* __restrict__ X, long long sx, int B, int n_out,
                                    int n_in, float* __restrict__ dW, float* __restrict__ dbias) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_out * (n_in + 1)) return;
    const int o = idx / (n_in + 1), i = idx - o * (n_in + 1);
    float a = 0.f;
    if (i < n_in) {
        for (int b = 0; b < B; ++b) a = fmaf(dY[b * sy + o], X[b * sx + i], a);
        dW[(long long)o * n_in + i] += a;
    } else {
        for (int b = 0; b < B; ++b) a += dY[b * sy + o];
        dbias[o] += a;
    }
}

// =================================================================================================
// Optimizer + weight refresh
// =================================================================================================
// torch.optim.Adam defaults (train/train.py:144): no weight decay, no amsgrad.  grad carries the loss scale.
__global__ void adam_kernel(float* __restrict__ w, float* __restrict__ m, float* __restrict__ v, float* __restrict__ gbuf, long long n,
                            float lr, float beta1, float beta2, float eps, float bc1, float bc2, float inv_scale) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float gr = gbuf[i] * inv_scale;
    const float mi = beta1 * m[i] + (1.f - beta1) * gr;
    const float vi = beta2 * v[i] + (1.f - beta2) * gr * gr;
    m[i] = mi; v[i] = vi;
    w[i] -= lr * (mi / bc1) / (sqrtf(vi / bc2) + eps);
}

// master conv weight -> the layouts the kernels read.  mode 0: fp32 [tap][ci_pad][co]; 1: fp16 R-stacked tcgen05 layout (see
// mma_pack_weights); 2 / 3: the same two layouts of the data-gradient convolution (taps flipped, channels transposed,
// output channels = input channels [ci_lo, ci_lo + n_out) of the forward conv)
//   4: gather-tcgen05 layout [tap][kc][co][8] (gather_pack_weights); 5: the merged output-pair layout of the stride-2 transposed
//   conv [(a*kp+b)*3+dr+1][kc][2*Cout][8] (gather_pack_weights_pairT); both read the mode-0 array `pk`, optionally with the
//   channel roles swapped.
__device__ __forceinline__ void repack_one(const RepackParams& p, long long i);
__global__ void repack_conv_kernel(RepackParams p) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p.n) repack_one(p, i);
}
// all jobs of a handle in one launch: grid (blocks of the largest job, jobs)
__global__ void repack_table_kernel(const RepackParams* __restrict__ table) {
    __shared__ RepackParams p;
    if (threadIdx.x == 0) p = table[blockIdx.y];
    __syncthreads();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p.n) repack_one(p, i);
}
__device__ __forceinline__ void repack_one(const RepackParams& p, long long i) {
    if (p.mode >= 4) {
        const int Cg_in = p.swap ? p.Cout : p.Cin_pad, Cg_out = p.swap ? p.Cin_pad : p.Cout;
        const int nkc = (Cg_in + 15) / 16 * 2;
        long long t = i;
        const int j = (int)(t % 8); t /= 8;
        int n, kc, tap, co;
        bool live = true;
        if (p.mode == 4) {
            n = (int)(t % Cg_out); t /= Cg_out;
            kc = (int)(t % nkc); t /= nkc;
            tap = (int)t; co = n;
        } else {
            const int N2 = 2 * Cg_out;
            n = (int)(t % N2); t /= N2;
            kc = (int)(t % nkc); t /= nkc;
            const int di = (int)(t % 3); t /= 3;          // t = a*kp + b
            const int er = n / Cg_out;
            co = n - er * Cg_out;
            const int c = er + 1 - 2 * (di - 1);
            live = c >= 0 && c < p.kr;
            tap = (int)t * p.kr + c;
        }
        const int ci = kc * 8 + j;
        float v = 0.f;
        if (live && ci < Cg_in) {
            // straight from the master (not from the mode-0 array: all jobs of a handle run in ONE launch, unordered)
            const int fci = p.swap ? co : ci, fco = p.swap ? ci : co;          // forward conv's input / output channel
            if (fci < p.Cin_real)
                v = p.transposed ? p.master[((long long)fci * p.Cout + fco) * p.taps + tap] : p.master[((long long)fco * p.Cin_real + fci) * p.taps + tap];
        }
        p.out_h[i] = __float2half_rn(v);
        return;
    }
    const bool dgrad = p.mode >= 2;
    const int Cin = dgrad ? p.Cout : p.Cin_pad, Cout = dgrad ? p.n_out : p.Cout;     // of the convolution being packed
    int tap, ci, co;
    if ((p.mode & 1) == 0) {
        co = (int)(i % Cout); ci = (int)((i / Cout) % Cin); tap = (int)(i / ((long long)Cout * Cin));
    } else {
        const int CB = p.CB, ng = p.taps == 27 ? 3 : 1, nshift = p.taps == 27 ? 9 : 1, N = ng * CB, nkc = Cin / 8;
        long long t = i;
        const int j = (int)(t % 8); t /= 8;
        const int n = (int)(t % N); t /= N;
        const int kc = (int)(t % nkc); t /= nkc;
        const int s = (int)(t % nshift); t /= nshift;
        const int cb = (int)t;
        tap = s * ng + n / CB; co = cb * CB + n % CB; ci = kc * 8 + j;
    }
    float v = 0.f;
    if (!dgrad) {
        if (ci < p.Cin_real) v = p.transposed ? p.master[((long long)ci * p.Cout + co) * p.taps + tap] : p.master[((long long)co * p.Cin_real + ci) * p.taps + tap];
    } else {
        // packed conv: in-channel = forward co (ci here), out-channel = forward ci (co here + ci_lo), tap flipped
        const int fci = co + p.ci_lo, fco = ci, ftap = p.taps - 1 - tap;
        if (fci < p.Cin_real) v = p.master[((long long)fco * p.Cin_real + fci) * p.taps + ftap];
    }
    if (p.mode & 1) p.out_h[i] = __float2half_rn(v);
    else p.out_f[i] = v;
}

__global__ void fill_zero_kernel(float* p, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = 0.f;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
void launch_loss_backward(const LossBwdParams& p, int B, cudaStream_t st) {
    const long long nvox = (long long)B * p.g.Z * p.g.P * p.g.R;
    loss_backward_kernel<<<(unsigned)((nvox + 255) / 256), 256, 0, st>>>(p, nvox);
}

void launch_gn_backward(GnBwdParams p, int B, cudaStream_t st) {
    const Geom& g = p.g;
    const int nch = p.C >> 3;
    if (nch < 1 || nch > kGbThreads / 8 || p.C > kBwdMaxC) return;     // checked by the caller
    gn_bwd_prep_kernel<<<B, 128, 0, st>>>(p);
    p.rows_per_block = 1024;     // 2048 left the last of 5 blocks per DS2 sample with 88 rows and one ragged wave (ncu: 30 % warps active)
    dim3 grid1((g.rows_per_sample + p.rows_per_block - 1) / p.rows_per_block, B);
    gn_bwd_reduce_kernel<<<grid1, kGbThreads, 0, st>>>(p);
    gn_bwd_finalize_kernel<<<B, 256, 0, st>>>(p);
    dim3 grid2((g.rows_per_sample + kGbThreads - 1) / kGbThreads, B);
    gn_bwd_apply_kernel<<<grid2, kGbThreads, 0, st>>>(p);
}

int launch_wgrad_mma(const WgradParams& p, cudaStream_t st) {
    if (p.Cout % 32 || p.Cin_pad % 16 || (p.taps != 27 && p.taps != 1)) return 1;
    if (p.x1 && (p.src_chunks0 % 4 || p.Cin_pad % 32)) return 1;          // a ci block must not straddle the two sources
    const Geom& g = p.g;
    const int halo = p.taps == 27 ? g.plane + g.Rp + 1 : 0;
    if (halo > g.guard) return 1;
    const size_t smem = ((size_t)4 * (2 * kWgRows + 2 * halo) + 2 * 4 * kWgRows) * 16;      // X ring + two dY buffers
    if (smem > 227 * 1024) return 1;
    static size_t configured = 0;
    if (smem > configured) {
        if (cudaFuncSetAttribute(wgrad_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 1;
        configured = smem;
    }
    WgradParams q = p;
    q.ncob = p.Cout / 32;
    const int nblk = ((p.Cin_pad + 31) / 32) * q.ncob;
    const long long nrows = (long long)p.B * g.rows_per_sample;
    const int nchunks = (int)((nrows + kWgRows - 1) / kWgRows);
    // one CTA per SM (the accumulators take the register file); contiguous chunk ranges, at least 4 chunks each so the window
    // start-up (wrows rows) is amortised
    int gx = 148 / nblk > 0 ? 148 / nblk : 1;
    if (gx > (nchunks + 3) / 4) gx = (nchunks + 3) / 4;
    if (gx < 1) gx = 1;
    gx = (nchunks + ((nchunks + gx - 1) / gx) - 1) / ((nchunks + gx - 1) / gx);        // no empty CTAs
    wgrad_mma_kernel<<<dim3(gx, nblk), kWgWarps * 32, smem, st>>>(q);
    return 0;
}

void launch_conv_strided_backward(const ConvParams& p, const __half* x, const __half* dy, __half* dx, int dx_accumulate, float* dw,
                                  int Cin_real, int master_transposed, cudaStream_t st) {
    if (dx) {
        const long long nrows = (long long)p.B * p.gin.rows_per_sample;
        dim3 block(32, p.Cin_total / 8);
        conv_strided_dgrad_kernel<<<(unsigned)((nrows + 31) / 32), block, 0, st>>>(p, dy, dx, dx_accumulate, nrows);
    }
    if (dw) {
        const int taps = p.kz * p.kp * p.kr;
        const int nblk = ((p.Cin_total + 31) / 32) * (p.Cout / 32);
        int split = (148 * 8) / (taps * nblk);
        if (split < 1) split = 1;
        conv_strided_wgrad_kernel<<<dim3(taps, nblk, split), 256, 0, st>>>(p, x, dy, dw, Cin_real, master_transposed);
    }
}

void launch_channel_sum(const Geom& g, const __half* t, int C, float* out, int B, cudaStream_t st) {
    const long long nrows = (long long)B * g.rows_per_sample;
    channel_sum_kernel<<<148 * 2, 256, 0, st>>>(g, t, C, out, nrows);
}

void launch_plcp_accumulate(const Geom& g, const __half* src, __half* dst, int C, int accumulate, int B, cudaStream_t st) {
    const long long nrows = (long long)B * g.rows_per_sample;
    plcp_accumulate_kernel<<<(unsigned)((nrows + 255) / 256), 256, 0, st>>>(g, src, dst, C, accumulate, nrows);
}

int launch_attn_backward(AttnBwdParams p, int B, cudaStream_t st) {
    if (p.C > kBwdMaxC || p.C % 8) return 1;
    fill_zero_kernel<<<(unsigned)(((long long)B * 1024 + 255) / 256), 256, 0, st>>>(p.dcx, (long long)B * 1024);
    p.rows_per_block = 2048;
    dim3 grid((p.g.rows_per_sample + p.rows_per_block - 1) / p.rows_per_block, B);
    if (p.C % 16) return 1;
    const size_t smem = (size_t)32 * kAbStr + (size_t)32 * (p.C * 2 + 16) + ((size_t)3 * 4 + p.C / 8) * 128 * 8 * sizeof(__half);
    static size_t configured = 0;
    if (smem > configured) {
        if (cudaFuncSetAttribute(attn_bwd_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 1;
        configured = smem;
    }
    attn_bwd_reduce_kernel<<<grid, 128, smem, st>>>(p);
    static const bool simt_apply = getenv("CD_ATTN_BWD_SIMT") != nullptr;       // diagnostics: the per-voxel SIMT form
    if (simt_apply) { attn_bwd_apply_kernel<<<grid, 128, 0, st>>>(p); return 0; }
    const size_t asmem = (size_t)2 * 32 * kAbStr + (size_t)32 * (p.C * 2 + 16) + 3 * 32 * sizeof(float);
    p.rows_per_block = 1024;
    dim3 agrid((p.g.rows_per_sample + p.rows_per_block - 1) / p.rows_per_block, B);
    attn_bwd_apply_mma_kernel<<<agrid, 128, asmem, st>>>(p);
    return 0;
}

void launch_embed_backward(const EmbedParams& p, const EmbedBwdLayout& L, const float* sigma, const float* cond, const float* daddvec,
                           float* rec, int B, cudaStream_t st) {
    embed_backward_kernel<<<B, 256, 0, st>>>(p, L, sigma, cond, daddvec, rec);
}

// all Linear weight gradients of a step in one launch: grid (blocks of the largest job, jobs), B passed per launch
__global__ void linear_wgrad_table_kernel(const LinearJob* __restrict__ table, int B) {
    const LinearJob j = table[blockIdx.y];
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= j.n_out * (j.n_in + 1)) return;
    const int o = idx / (j.n_in + 1), i = idx - o * (j.n_in + 1);
    float a0 = 0.f, a1 = 0.f;
    if (i < j.n_in) {
        int b = 0;
        for (; b + 1 < B; b += 2) {           // two independent chains: the loop is load-latency bound
            a0 = fmaf(j.dY[b * j.sy + o], j.X[b * j.sx + i], a0);
            a1 = fmaf(j.dY[(b + 1) * j.sy + o], j.X[(b + 1) * j.sx + i], a1);
        }
        if (b < B) a0 = fmaf(j.dY[b * j.sy + o], j.X[b * j.sx + i], a0);
        j.dW[(long long)o * j.n_in + i] += a0 + a1;
    } else {
        for (int b = 0; b < B; ++b) a0 += j.dY[b * j.sy + o];
        j.dbias[o] += a0;
    }
}
void launch_linear_wgrad_table(const LinearJob* table_dev, int njobs, int max_n, int B, cudaStream_t st) {
    if (njobs < 1) return;
    linear_wgrad_table_kernel<<<dim3((max_n + 127) / 128, njobs), 128, 0, st>>>(table_dev, B);
}
void launch_linear_wgrad(const float* dY, long long sy, const float* X, long long sx, int B, int n_out, int n_in, float* dW, float* dbias,
                         cudaStream_t st) {
    const int n = n_out * (n_in + 1);
    linear_wgrad_kernel<<<(n + 127) / 128, 128, 0, st>>>(dY, sy, X, sx, B, n_out, n_in, dW, dbias);
}

void launch_adam(float* w, float* m, float* v, float* g, long long n, float lr, float beta1, float beta2, float eps, int step,
                 float inv_scale, cudaStream_t st) {
    const float bc1 = 1.f - powf(beta1, (float)step), bc2 = 1.f - powf(beta2, (float)step);
    adam_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(w, m, v, g, n, lr, beta1, beta2, eps, bc1, bc2, inv_scale);
}

void launch_repack_table(const RepackParams* table_dev, int njobs, long long max_n, cudaStream_t st) {
    if (njobs < 1) return;
    repack_table_kernel<<<dim3((unsigned)((max_n + 255) / 256), njobs), 256, 0, st>>>(table_dev);
}
void launch_repack_conv(const RepackParams& p, cudaStream_t st) {
    repack_conv_kernel<<<(unsigned)((p.n + 255) / 256), 256, 0, st>>>(p);
}

void launch_fill_zero(float* p, long long n, cudaStream_t st) {
    fill_zero_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p, n);
}

}  // namespace cd
