// tcgen05 / TMEM implicit-GEMM convolution for stride-1 3x3x3 CylindricalConv with the three Z taps stacked in GEMM-N, sm_100a
// (reference semantics: models/models.py:65-96, circular phi + zero Z/R padding = the halo rows of the PLCP tensor).
//
// cd_conv_mma.cu stacks the three R taps in N and pays for it in the epilogue: the partial sums of an output row sit on three
// neighbouring TMEM lanes (rows m-1, m, m+1), so every output element costs two shuffles plus a cross-warp exchange, and that
// epilogue - not the tensor pipe - bounds the kernel.  Here the GEMM is
//     E[r, a*CB + co] = sum over (dp,dr) in {-1,0,1}^2, ci:  X[r + dp*Rp + dr, ci] * W[dz=a-1, dp, dr][ci, co]
//     out[r, co]      = E[r - plane, 0*CB + co] + E[r, 1*CB + co] + E[r + plane, 2*CB + co] + bias[co]
// and the M tiles are STRIPS: 128 rows starting at p*plane + 128*c (p = z-plane index over the whole batch, c = strip of the
// plane).  Row r - plane of strip (p, c) is the same lane of strip (p-1, c): the three partial sums of an output row are on the
// SAME TMEM lane of three consecutive planes' accumulators, and the recombination is three TMEM loads and two additions - no
// shuffles, no shared memory, no barrier.  A CTA owns one strip index c and walks a contiguous range of planes with a rolling
// window of accumulator slots (E(p-1), E(p), E(p+1) being read while E(p+2) is computed).  Row decode disappears too: a
// thread's (phi row, r) is fixed by its strip and lane for the whole launch, only z changes, uniformly.
// Cost: strips have a period of one plane, so a plane of `plane` rows costs ceil(plane/128) M = 128 MMAs (an M = 64 MMA is no
// cheaper): 96 % row efficiency for Dataset 3's 988-row plane, 70 % for Dataset 2's 180 rows - the launcher takes this form
// only where that efficiency is high, cd_conv_mma.cu otherwise.
//
//   * A operand: one cp.async.bulk window of 128 + 2*(Rp+1) rows per 8-channel chunk and strip; the nine (dp,dr) shifts are row
//     offsets into it through the UMMA shared-memory descriptor (no-swizzle K-major, as in cd_conv_mma.cu).
//   * B operand: packed weights [shift (dp,dr)][kchunk][n = a*CB + co][8] fp16, resident in shared memory.
//   * D: 5 TMEM slots of 128 lanes x 96 fp32 columns, slot of plane-strip k = k % 5.  Each slot is read by the epilogues of three
//     consecutive output planes; its empty barrier counts all three (12 warp arrivals), which is what lets the epilogue groups run
//     out of step with one another.  Two MMA-issuing warps alternate plane-strips.
#include <cuda.h>
#include <cstdio>
#include <cstdlib>

#include "cd_kernels.cuh"

namespace cd {
namespace {

constexpr int kCB = 32;              // output channels (one block)
constexpr int kN = 3 * kCB;          // GEMM N
constexpr int kSlots = 5;            // 5 x 96 = 480 of the 512 TMEM columns
constexpr int kMaxStages = 8;
// epilogue groups (output planes in flight): three; two in the fused-GroupNorm form, whose eight transform warps take their place
#ifndef ZS_G_FUSED
#define ZS_G_FUSED 2
#endif
#ifndef ZS_G_PLAIN
#define ZS_G_PLAIN 3
#endif
__host__ __device__ constexpr int zs_groups(bool fuse) { return fuse ? ZS_G_FUSED : ZS_G_PLAIN; }
constexpr int kIssuers = 2;          // MMA-issuing warps, alternating plane-strips (one warp issues an MMA every ~80 clk, the pipe takes one every 56)
constexpr int kFirstEpi = 1 + kIssuers;
#ifndef ZS_XF_WARPS
#define ZS_XF_WARPS 8
#endif
constexpr int kXfWarps = ZS_XF_WARPS;          // transform warps of the fused-GroupNorm form
__host__ __device__ constexpr int zs_first_xf(bool fuse) { return kFirstEpi + 4 * zs_groups(fuse); }
__host__ __device__ constexpr int zs_threads(bool fuse) { return 32 * (zs_first_xf(fuse) + (fuse ? kXfWarps : 0)); }

struct ZsParams {
    Geom g;
    const __half* src[2];
    int src_chunks[2];
    int nchunks;            // Cin / 8
    const __half* w;
    const float* bias;
    __half* dst;
    double* stats;
    int stats_C;
    int S;                  // strips per plane
    int ppc;                // output planes per CTA
    int NP;                 // planes in the batch: B * (Z + 1)
    int wrows;              // rows per shared-memory window: 128 + 2 * (Rp + 1)
    int a_stage_bytes;
    int nstages;
    int w_bytes;
    uint32_t idesc;
    int write_halo;
    // fused GroupNorm + SiLU + conditioning on the input (ConvParams::gn_*)
    const double* gn_stats; const float* gn_gamma; const float* gn_beta; int gn_cpg; double gn_n; const float* gn_addvec; int gn_addvec_stride;
    int B;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// bounded wait: a protocol bug must trap, not hang the GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try(bar, parity)) return;
    long long t0 = clock64();
    while (!mbar_try(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) { printf("cd_conv_zstack: mbarrier timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x); __trap(); }
    }
}
__device__ __forceinline__ void mbar_wait_relaxed(uint32_t bar, uint32_t parity) {
    if (mbar_try(bar, parity)) return;
    long long t0 = clock64();
    while (!mbar_try(bar, parity)) {
        __nanosleep(40);
        if (clock64() - t0 > 4000000000LL) { printf("cd_conv_zstack: mbarrier timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x); __trap(); }
    }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;  // descriptor version (sm_100)
    return d;                // base_offset 0, lbo_mode 0, layout SWIZZLE_NONE
}
// warp-convergent forms: every lane executes the call, one elected lane performs the operation (see cd_conv_mma.cu)
__device__ __forceinline__ void umma_f16_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n.reg .pred p, e;\nelect.sync _|e, 0xffffffff;\nsetp.ne.b32 p, %4, 0;\n@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit_elect(uint32_t bar) {
    asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\n@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}"
                 ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// SiLU through one MUFU, as gn_apply_kernel: x*sigmoid(x) = 0.5*x*(1 + tanh(x/2))
__device__ __forceinline__ float zs_silu_half(float h) {      // h = x/2
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
    return fmaf(h, t, h);
}

struct ZsCtl {
    uint64_t a_full[kMaxStages], a_empty[kMaxStages], a_ready[kMaxStages];
    uint64_t t_full[kSlots], t_empty[kSlots];
    uint64_t w_res;
    uint32_t tmem_base;
    uint32_t pad;
};
constexpr int kZsCtlBytes = 512;
constexpr int kZsBiasBytes = 128;
static_assert(sizeof(ZsCtl) <= kZsCtlBytes, "ctl block too small");

// CPG: channels per GroupNorm group of the output statistics; NQ: K16 steps per shift (C_in / 16); FUSE: GroupNorm + SiLU +
// conditioning applied to the staged input windows by four more warps before the MMAs read them
template <int CPG, int NQ, bool FUSE>
__global__ void __launch_bounds__(zs_threads(FUSE), 1) conv_zs_kernel(const ZsParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    ZsCtl* ctl = reinterpret_cast<ZsCtl*>(smem);
    float* sbias = reinterpret_cast<float*>(smem + kZsCtlBytes);
    unsigned char* wbuf = smem + kZsCtlBytes + kZsBiasBytes;
    unsigned char* abuf = wbuf + ((p.w_bytes + 127) / 128) * 128;
    const uint32_t wbuf_s = smem_u32(wbuf), abuf_s = smem_u32(abuf);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const Geom& g = p.g;
    constexpr int kGroups = zs_groups(FUSE), kFirstXf = zs_first_xf(FUSE);

    if (threadIdx.x == 0) {
        for (int s = 0; s < kMaxStages; ++s) {
            mbar_init(smem_u32(&ctl->a_full[s]), 1); mbar_init(smem_u32(&ctl->a_empty[s]), 1); mbar_init(smem_u32(&ctl->a_ready[s]), kXfWarps);
        }
        for (int s = 0; s < kSlots; ++s) { mbar_init(smem_u32(&ctl->t_full[s]), 1); mbar_init(smem_u32(&ctl->t_empty[s]), 12); }
        mbar_init(smem_u32(&ctl->w_res), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < kCB) sbias[threadIdx.x] = p.bias ? __ldg(p.bias + threadIdx.x) : 0.f;
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&ctl->tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = ctl->tmem_base;

    // this CTA: strip c of the output planes [p0, p1); E is computed for the planes p0-1 .. p1 (K = J + 2 plane-strips, index k)
    const int c = blockIdx.x % p.S;
    const int p0 = (blockIdx.x / p.S) * p.ppc;
    const int p1 = p0 + p.ppc < p.NP ? p0 + p.ppc : p.NP;
    const int J = p1 > p0 ? p1 - p0 : 0;
    const int K = J > 0 ? J + 2 : 0;
    const uint32_t chunk_bytes = (uint32_t)p.wrows * 16u;

    if (warp == 0) {
        // ===================== producer: resident weights + one activation window per plane-strip =====================
        if (lane == 0 && K > 0) {
            mbar_expect_tx(smem_u32(&ctl->w_res), p.w_bytes);
            for (int off = 0; off < p.w_bytes; off += 32768) {
                const int n = p.w_bytes - off < 32768 ? p.w_bytes - off : 32768;
                bulk_g2s(wbuf_s + off, reinterpret_cast<const unsigned char*>(p.w) + off, n, smem_u32(&ctl->w_res));
            }
            int st = 0;
            uint32_t ph = 0;
            long long row = (long long)(p0 - 1) * g.plane + 128 * c - (g.Rp + 1);     // first window row of plane-strip k = 0
            for (int k = 0; k < K; ++k, row += g.plane, st = (st + 1 == p.nstages) ? 0 : st + 1, ph ^= (uint32_t)(st == 0)) {
                mbar_wait_relaxed(smem_u32(&ctl->a_empty[st]), ph ^ 1);
                mbar_expect_tx(smem_u32(&ctl->a_full[st]), p.a_stage_bytes);
                for (int ch = 0; ch < p.nchunks; ++ch) {
                    const int sidx = ch < p.src_chunks[0] ? 0 : 1;
                    const int kk = sidx ? ch - p.src_chunks[0] : ch;
                    bulk_g2s(abuf_s + st * p.a_stage_bytes + ch * chunk_bytes, p.src[sidx] + plcp_offset(g, kk, row), chunk_bytes,
                             smem_u32(&ctl->a_full[st]));
                }
            }
        }
    } else if (warp < kFirstEpi) {
        // ===================== MMA issuers (warp-convergent, one elected lane each; issuer i takes the plane-strips k = i mod 2) =====
        // Slot ring safety with two producers: issuer A reaches strip k only after its strip k-2 got its slot, i.e. after the reader
        // of plane k-5 finished, i.e. after the OTHER issuer's strip k-5 - the previous use of A's slot k % 5 - was issued and read.
        if (K > 0) {
            const int iw = warp - 1;
            const uint64_t adesc_hi = make_desc(0, chunk_bytes, 128u);
            const uint64_t bdesc_hi = make_desc(0, (uint32_t)kN * 16u, 128u);
            const uint32_t ahi32 = (uint32_t)(adesc_hi >> 32), bhi32 = (uint32_t)(bdesc_hi >> 32);
            const uint32_t qstep = (2u * chunk_bytes) >> 4, bstep = (2u * kN * 16u) >> 4;
            const uint32_t bshift = (uint32_t)p.nchunks * kN;     // weight offset (16-byte units) between shifts
            uint32_t a_s[9], b_s[9];                              // descriptor low words per shift (dp,dr): row offset / weight block
#pragma unroll
            for (int s = 0; s < 9; ++s) {
                a_s[s] = (uint32_t)adesc_hi + (uint32_t)((g.Rp + 1) + (s / 3 - 1) * g.Rp + (s % 3 - 1));
                b_s[s] = (uint32_t)bdesc_hi + (wbuf_s >> 4) + (uint32_t)s * bshift;
            }
            mbar_wait(smem_u32(&ctl->w_res), 0);
            int st = 0;
            uint32_t aph = 0, slot = 0, tph = 0;
            for (int k = 0; k < K; ++k) {
                if ((k & 1) == iw) {
                    mbar_wait(smem_u32(FUSE ? &ctl->a_ready[st] : &ctl->a_full[st]), aph);
                    mbar_wait(smem_u32(&ctl->t_empty[slot]), tph ^ 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a_lo0 = (abuf_s + st * p.a_stage_bytes) >> 4;
                    const uint32_t dcol = tmem_base + slot * kN;
#pragma unroll
                    for (int s = 0; s < 9; ++s)
#pragma unroll
                        for (int q = 0; q < NQ; ++q)
                            umma_f16_elect(dcol, ((uint64_t)ahi32 << 32) | (a_s[s] + a_lo0 + (uint32_t)q * qstep),
                                           ((uint64_t)bhi32 << 32) | (b_s[s] + (uint32_t)q * bstep), p.idesc, (s | q) != 0);
                    umma_commit_elect(smem_u32(&ctl->t_full[slot]));
                    umma_commit_elect(smem_u32(&ctl->a_empty[st]));
                }
                st = (st + 1 == p.nstages) ? 0 : st + 1; aph ^= (uint32_t)(st == 0);
                slot = (slot + 1 == kSlots) ? 0u : slot + 1u; tph ^= (uint32_t)(slot == 0u);
            }
        }
    } else if (FUSE && warp >= kFirstXf) {
        // ===================== transform: GroupNorm + SiLU + conditioning on the staged windows, in place =====================
        // A thread keeps one 8-channel chunk (its scale / shift / conditioning live in registers, re-derived when the sample
        // changes) and walks that chunk's rows of the window.  Rows outside the plane, rows of the zero column (rr = R) and whole
        // zero planes (z = Z, planes outside the batch) are written as zeros - what gn_apply_kernel leaves there.  (The only
        // out-of-plane row an interior output reads is row -1, a zero-column row.)
        if (K > 0) {
            constexpr int NCH = 2 * NQ, RSTEP = (32 * kXfWarps) / NCH;
            const int tid = (warp - kFirstXf) * 32 + lane;
            const int ch = tid / RSTEP, r0 = tid % RSTEP;      // a warp's lanes take consecutive rows of ONE chunk: 512 contiguous bytes, no bank conflicts
            uint32_t zmask = 0;                       // bit n: window row r0 + RSTEP*n is always zero
            int npieces = 0;
            for (int w = r0; w < p.wrows; w += RSTEP, ++npieces) {
                const int prow = 128 * c - (g.Rp + 1) + w;
                const bool zero = prow < 0 || prow >= g.plane || (prow % g.Rp) == g.R;
                zmask |= (uint32_t)zero << npieces;
            }
            float sc[8], sh[8], ad[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) { sc[j] = 0.f; sh[j] = 0.f; ad[j] = 0.f; }
            const int Zp1 = g.Z + 1, C = 8 * NCH;
            int q = p0 - 1;
            int b = q >= 0 ? q / Zp1 : -1, z = q - b * Zp1, cur_b = -2;
            int st = 0;
            uint32_t ph = 0;
            for (int k = 0; k < K; ++k, ++q, ++z) {
                if (z >= Zp1) { z -= Zp1; ++b; }
                const bool plane_ok = b >= 0 && b < p.B && z < g.Z;
                if (plane_ok && b != cur_b) {
                    cur_b = b;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int cc = 8 * ch + j, g0 = (cc / p.gn_cpg) * p.gn_cpg;
                        double s = 0.0, qq = 0.0;
                        for (int i = 0; i < p.gn_cpg; ++i) {
                            const double* d = p.gn_stats + ((long long)b * C + g0 + i) * 2;
                            s += d[0]; qq += d[1];
                        }
                        const double n = p.gn_n;
                        const float mean = (float)(s / n);
                        float var = (float)(qq / n - (s / n) * (s / n));
                        if (var < 0.f) var = 0.f;
                        // (halved: SiLU's x/2 is taken here - scaling by a power of two commutes with the rounding of the FMA)
                        const float scj = __ldg(p.gn_gamma + cc) * rsqrtf(var + 1e-5f);
                        sc[j] = 0.5f * scj;
                        sh[j] = 0.5f * (__ldg(p.gn_beta + cc) - mean * scj);
                        ad[j] = p.gn_addvec ? __ldg(p.gn_addvec + (long long)b * p.gn_addvec_stride + cc) : 0.f;
                    }
                }
                mbar_wait_relaxed(smem_u32(&ctl->a_full[st]), ph);
                unsigned char* base = abuf + (size_t)st * p.a_stage_bytes + (size_t)ch * chunk_bytes + (size_t)r0 * 16;
                // three pieces per batch: their loads, MUFU chains and stores overlap
                for (int n0 = 0; n0 < npieces; n0 += 3) {
                    uint4 raw[3];
                    bool live[3];
#pragma unroll
                    for (int u = 0; u < 3; ++u) {
                        const int n = n0 + u;
                        live[u] = n < npieces && plane_ok && !((zmask >> n) & 1u);
                        raw[u] = make_uint4(0u, 0u, 0u, 0u);
                        if (live[u]) raw[u] = *reinterpret_cast<const uint4*>(base + (size_t)n * RSTEP * 16);
                    }
#pragma unroll
                    for (int u = 0; u < 3; ++u) {
                        if (live[u]) {
                            float f[8];
                            unpack8(*reinterpret_cast<const Half8*>(&raw[u]), f);
#pragma unroll
                            for (int j = 0; j < 8; ++j) f[j] = zs_silu_half(fmaf(f[j], sc[j], sh[j])) + ad[j];
                            const Half8 o = pack8(f);
                            raw[u] = *reinterpret_cast<const uint4*>(&o);
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 3; ++u)
                        if (n0 + u < npieces) *reinterpret_cast<uint4*>(base + (size_t)(n0 + u) * RSTEP * 16) = raw[u];
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> visible to the UMMA operand reads
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(&ctl->a_ready[st]));
                st = (st + 1 == p.nstages) ? 0 : st + 1; ph ^= (uint32_t)(st == 0);
            }
        }
    } else {
        // ===================== epilogue: three TMEM loads per output element, bias, statistics, stores =====================
        const int q = warp & 3;                      // TMEM lane quarter this warp may read (warp id % 4)
        const int group = (warp - kFirstEpi) >> 2;   // output planes j = group, group + G, ...
        const int prow = 128 * c + q * 32 + lane;    // this thread's row within every plane: fixed for the launch
        const int ppi = prow / g.Rp, rri = prow - ppi * g.Rp;
        const bool lane_ok = prow < g.plane && rri < g.R && ppi >= 1 && ppi <= g.P;
        const bool halo_up = lane_ok && p.write_halo && ppi == 1;        // copy to pp = P+1
        const bool halo_dn = lane_ok && p.write_halo && ppi == g.P;      // copy to pp = 0
        const long long halo_delta8 = (long long)g.P * g.Rp * 8;
        const long long cs8 = g.chunk_stride * 8;
        constexpr int NG = kCB / CPG;
        float gs[NG], gq[NG];
#pragma unroll
        for (int i = 0; i < NG; ++i) { gs[i] = 0.f; gq[i] = 0.f; }
        int cur_b = -1;
        auto flush = [&]() {   // warp-uniform call sites only
            if (cur_b >= 0) {
#pragma unroll
                for (int i = 0; i < NG; ++i) {
                    float s_tot = gs[i], q_tot = gq[i];
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) { s_tot += __shfl_xor_sync(0xffffffffu, s_tot, o); q_tot += __shfl_xor_sync(0xffffffffu, q_tot, o); }
                    if (lane == 0) {
                        double* d = p.stats + ((long long)cur_b * p.stats_C + i * CPG) * 2;
                        atomicAdd(d, (double)s_tot);
                        atomicAdd(d + 1, (double)q_tot);
                    }
                    gs[i] = 0.f; gq[i] = 0.f;
                }
            }
        };
        // iteration jj = j + 2 of this group: output plane j = jj - 2 reads the plane-strips {jj-2, jj-1, jj} that exist (0 <= k < K);
        // jj = 0, 1 and jj = J+2, J+3 are the virtual planes before / after the range: they only release the slots a real plane
        // would have read there, so that every slot use sees exactly three readers.
        const int Zp1 = g.Z + 1;
        int pl = p0 + group - 2;                                         // global plane of iteration jj = group
        int b = pl >= 0 ? pl / Zp1 : -1, z = pl - b * Zp1;               // (floor semantics for the virtual planes before plane 0)
        __half* optr = p.dst + ((long long)g.guard + (long long)pl * g.plane + prow) * 8;
        const long long ostep = (long long)kGroups * g.plane * 8;
        for (int jj = group; jj < J + 4 && K > 0; jj += kGroups, z += kGroups, optr += ostep) {
            while (z >= Zp1) { z -= Zp1; ++b; }
            // every plane-strip this iteration reads (or only releases) must be complete: consecutive strips come from different
            // issuing warps, and a commit only covers the MMAs of the warp that made it
#pragma unroll
            for (int d = 2; d >= 0; --d) {
                const int k = jj - d;
                if (k >= 0 && k < K) {
                    const uint32_t lap = (uint32_t)k / (uint32_t)kSlots;
                    mbar_wait_relaxed(smem_u32(&ctl->t_full[(uint32_t)k - lap * kSlots]), lap & 1u);
                }
            }
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const bool real = jj >= 2 && jj < J + 2;
            // release: this iteration is done with the plane-strips {jj-2, jj-1, jj} that exist (as soon as its last TMEM load is back)
            auto release = [&]() {
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
#pragma unroll
                    for (int d = 0; d < 3; ++d) {
                        const int k = jj - d;
                        if (k >= 0 && k < K) mbar_arrive(smem_u32(&ctl->t_empty[(uint32_t)k % kSlots]));
                    }
                }
            };
            if (!(real && z < g.Z)) release();
            else {
                const uint32_t s0 = (uint32_t)(jj - 2) % kSlots, s1 = (uint32_t)(jj - 1) % kSlots, s2 = (uint32_t)jj % kSlots;
                const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
                if (p.stats && b != cur_b) { flush(); cur_b = b; }
#pragma unroll
                for (int c0 = 0; c0 < kCB; c0 += 16) {
                    uint32_t r0[16], r1[16], r2[16];
                    tmem_ld16_issue(lane_addr + s1 * kN + kCB + c0, r1);        // dz =  0 block of this plane
                    tmem_ld16_issue(lane_addr + s0 * kN + c0, r0);              // dz = -1 block of the previous plane
                    tmem_ld16_issue(lane_addr + s2 * kN + 2 * kCB + c0, r2);    // dz = +1 block of the next plane
                    tmem_ld_wait();
                    if (c0 + 16 >= kCB) release();
                    float acc[16];
#pragma unroll
                    for (int k = 0; k < 16; k += 4) {
                        const float4 bv = *reinterpret_cast<const float4*>(sbias + c0 + k);
                        acc[k] = __uint_as_float(r1[k]) + bv.x + (__uint_as_float(r0[k]) + __uint_as_float(r2[k]));
                        acc[k + 1] = __uint_as_float(r1[k + 1]) + bv.y + (__uint_as_float(r0[k + 1]) + __uint_as_float(r2[k + 1]));
                        acc[k + 2] = __uint_as_float(r1[k + 2]) + bv.z + (__uint_as_float(r0[k + 2]) + __uint_as_float(r2[k + 2]));
                        acc[k + 3] = __uint_as_float(r1[k + 3]) + bv.w + (__uint_as_float(r0[k + 3]) + __uint_as_float(r2[k + 3]));
                    }
                    if (lane_ok) {
#pragma unroll
                        for (int k = 0; k < 16; k += 8) {
                            const Half8 h = pack8(acc + k);
                            __half* o = optr + ((c0 + k) >> 3) * cs8;
                            store8(o, h);
                            if (p.write_halo) {
                                if (halo_up) store8(o + halo_delta8, h);
                                if (halo_dn) store8(o - halo_delta8, h);
                            }
                        }
                        if (p.stats) {
#pragma unroll
                            for (int k = 0; k < 16; ++k) { gs[(c0 + k) / CPG] += acc[k]; gq[(c0 + k) / CPG] = fmaf(acc[k], acc[k], gq[(c0 + k) / CPG]); }
                        }
                    }
                }
            }
        }
        if (p.stats) flush();
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
bool zstack_conv_supported(int Cin, int Cout, int taps) { return taps == 27 && Cout == kCB && (Cin == 16 || Cin == 32 || Cin == 64); }

size_t zstack_packed_elems(int Cin, int Cout, int taps) { return (size_t)taps * Cin * Cout; }

// w: fp32 [tap][ci][co], tap = (a*3+b)*3+c over (dz,dp,dr).  out: fp16 [shift = b*3+c][kc][n = a*CB + co][8]
void zstack_pack_weights(const float* w, int Cin, int Cout, __half* out) {
    const int nkc = Cin / 8;
    for (int s = 0; s < 9; ++s)
        for (int kc = 0; kc < nkc; ++kc)
            for (int n = 0; n < 3 * Cout; ++n)
                for (int j = 0; j < 8; ++j) {
                    const int a = n / Cout, co = n % Cout;
                    const int tap = a * 9 + s;
                    out[(((size_t)s * nkc + kc) * (3 * Cout) + n) * 8 + j] = __float2half_rn(w[((size_t)tap * Cin + kc * 8 + j) * Cout + co]);
                }
}

// Row efficiency of the strip decomposition: rows of a plane / rows the MMAs cover.
static double zstack_efficiency(const Geom& g) { const int S = (g.plane + 127) / 128; return (double)g.plane / (128.0 * S); }

// Break-even against the R-stacked kernel is near 0.78 (section 4 of DESIGN.md); CD_ZS_MIN_EFF overrides (diagnostics).
static double zstack_min_efficiency() {
    static const double v = [] { const char* e = getenv("CD_ZS_MIN_EFF"); return e ? atof(e) : 0.85; }();
    return v;
}
bool zstack_geom_efficient(const Geom& g) { return zstack_efficiency(g) >= zstack_min_efficiency(); }

// force = 1: take this form whatever its row efficiency (tests).  Returns 0 when launched.
int launch_conv_zstack(const ConvParams& cp, const __half* packed_w, int force, cudaStream_t st) {
    const int taps = cp.kz * cp.kp * cp.kr;
    if (cp.transposed || cp.sz != 1 || cp.sp != 1 || cp.sr != 1) return 1;
    if (!(cp.kz == 3 && cp.kp == 3 && cp.kr == 3 && cp.pz == 1 && cp.cp == 1 && cp.pr == 1)) return 1;
    if (!zstack_conv_supported(cp.Cin_total, cp.Cout, taps)) return 1;
    if (cp.gin.Z != cp.gout.Z || cp.gin.P != cp.gout.P || cp.gin.R != cp.gout.R) return 1;
    for (int s = 0; s < cp.nsrc; ++s) if (cp.src[s].C % 8 != 0) return 1;
    if (cp.addsrc || cp.pn_stats) return 1;
    const bool fuse = cp.gn_stats != nullptr;
    if (fuse && (cp.nsrc != 1 || cp.Cin_total != 32 || !cp.gn_gamma || !cp.gn_beta || cp.gn_groups <= 0 || cp.Cin_total % cp.gn_groups)) return 1;
    if (cp.write_halo && cp.gout.P < 2) return 1;
    const Geom& g = cp.gout;
    // the row windows reach Rp + 1 rows past either end of a strip and one plane past either end of the batch: inside the guard
    if (g.guard < g.plane + 128 + g.Rp + 1) return 1;
    static int num_sms = 0, max_smem = 0;
    if (!num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    }
    ZsParams p;
    p.g = g;
    p.S = (g.plane + 127) / 128;
    p.NP = cp.B * (g.Z + 1);
    if (p.S > num_sms) return 1;
    const int ctas_per_strip = num_sms / p.S;
    // worth it only where most MMA rows are real rows and every CTA walks enough planes to amortise its two extra plane-strips
    // (a 16-channel input - the init conv - has half the MMA work per strip and is epilogue-bound in the R-stacked form: there this
    //  form wins down to Dataset 2's 70 % row efficiency, 0.30 against 0.37 ms)
    const double min_eff = cp.Cin_total == 16 && zstack_min_efficiency() > 0.6 ? 0.6 : zstack_min_efficiency();
    if (!force && (zstack_efficiency(g) < min_eff || p.NP < 16 * ctas_per_strip || (long long)p.S * ctas_per_strip * 10 < 9LL * num_sms)) return 1;
    if ((long long)cp.B * g.rows_per_sample + 4096 >= (1LL << 31)) return 1;
    p.src[0] = cp.src[0].base; p.src_chunks[0] = cp.src[0].C / 8;
    p.src[1] = cp.nsrc > 1 ? cp.src[1].base : nullptr; p.src_chunks[1] = cp.nsrc > 1 ? cp.src[1].C / 8 : 0;
    p.nchunks = cp.Cin_total / 8;
    if (p.src_chunks[0] + p.src_chunks[1] != p.nchunks) return 1;
    p.w = packed_w; p.bias = cp.bias; p.dst = cp.dst; p.stats = cp.stats;
    p.stats_C = cp.stats_C > 0 ? cp.stats_C : cp.Cout;
    p.write_halo = cp.write_halo;
    p.gn_stats = cp.gn_stats; p.gn_gamma = cp.gn_gamma; p.gn_beta = cp.gn_beta; p.gn_addvec = cp.gn_addvec; p.gn_addvec_stride = cp.gn_addvec_stride;
    p.gn_cpg = fuse ? cp.Cin_total / cp.gn_groups : 1;
    p.gn_n = cp.gn_count * p.gn_cpg;
    p.B = cp.B;
    int cpg = 4;
    if (cp.stats) { cpg = cp.stats_cpg; if (cpg != 4 && cpg != 8) return 1; }
    p.ppc = (p.NP + ctas_per_strip - 1) / ctas_per_strip;
    const int ctas_used = (p.NP + p.ppc - 1) / p.ppc;
    p.wrows = 128 + 2 * (g.Rp + 1);
    if (const char* e = getenv("CD_ZS_WPAD")) p.wrows += atoi(e);      // diagnostics: chunk stride not a multiple of 128 bytes
    p.a_stage_bytes = p.nchunks * p.wrows * 16;
    p.w_bytes = 9 * cp.Cin_total * kN * 2;
    const size_t fixed = kZsCtlBytes + kZsBiasBytes + 256 + ((size_t)p.w_bytes + 127) / 128 * 128;
    if (fixed + 2 * (size_t)p.a_stage_bytes > (size_t)max_smem) return 1;
    p.nstages = (int)(((size_t)max_smem - fixed) / p.a_stage_bytes);
    if (p.nstages > kMaxStages) p.nstages = kMaxStages;
    const size_t smem_total = fixed + (size_t)p.nstages * p.a_stage_bytes;
    p.idesc = (1u << 4) | ((uint32_t)(kN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);  // f16 x f16 -> f32, K-major A and B, M = 128
    const int grid = p.S * ctas_used;
    cudaError_t e = cudaSuccess;
    auto go = [&](auto kern, bool& configured, bool fused) {
        if (!configured) { e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem); configured = e == cudaSuccess; }
        if (e == cudaSuccess) kern<<<grid, zs_threads(fused), smem_total, st>>>(p);
    };
    static bool cfg[8] = {false, false, false, false, false, false, false, false};
    const int nq = p.nchunks / 2;
    if (fuse && cpg == 4 && nq == 2) go(conv_zs_kernel<4, 2, true>, cfg[6], true);
    else if (fuse && cpg == 8 && nq == 2) go(conv_zs_kernel<8, 2, true>, cfg[7], true);
    else if (fuse) return 1;
    else if (cpg == 4 && nq == 1) go(conv_zs_kernel<4, 1, false>, cfg[0], false);
    else if (cpg == 4 && nq == 2) go(conv_zs_kernel<4, 2, false>, cfg[1], false);
    else if (cpg == 4 && nq == 4) go(conv_zs_kernel<4, 4, false>, cfg[2], false);
    else if (cpg == 8 && nq == 1) go(conv_zs_kernel<8, 1, false>, cfg[3], false);
    else if (cpg == 8 && nq == 2) go(conv_zs_kernel<8, 2, false>, cfg[4], false);
    else if (cpg == 8 && nq == 4) go(conv_zs_kernel<8, 4, false>, cfg[5], false);
    else return 1;
    return e == cudaSuccess ? 0 : 1;
}

}  // namespace cd
