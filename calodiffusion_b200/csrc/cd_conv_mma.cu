// tcgen05 / TMEM implicit-GEMM convolution for stride-1 CylindricalConv (3x3x3 and 1x1x1), sm_100a.
//
// GEMM view: D[M = padded-linear rows, N = C_out] += A_tap[M, C_in] * W_tap[C_in, C_out]
// summed over the 27 (or 1) taps.  Thanks to the PLCP layout (cd_common.cuh) the A operand of tap
// (dz,dp,dr) is the SAME activation rows shifted by dz*plane + dp*Rp + dr, so a CTA
//   * bulk-copies (cp.async.bulk, the 1-D form of TMA; chunk planes are contiguous) three z-windows of
//     (128*MT + 2*(Rp+1)) rows per 8-channel chunk into shared memory ONCE per tile,
//   * addresses every tap as a shifted window of that smem through the UMMA shared-memory descriptor
//     (no-swizzle K-major canonical layout: 8 rows x 16 B core matrices, SBO = 128 B, LBO = chunk stride),
//   * streams W_tap ([kchunk][n][8] fp16, N x K K-major) through an mbarrier ring,
//   * accumulates in TMEM (MT accumulators of 128 lanes x C_out fp32 columns) with tcgen05.mma issued by
//     one thread, and
//   * drains TMEM with tcgen05.ld in four epilogue warps that add the bias, round to fp16, store the
//     interior rows and reduce the GroupNorm statistics (sum / sum of squares per sample and channel).
// Circular phi padding and zero Z/R padding therefore cost no instructions here: they are the halo rows
// of the PLCP tensor (models/models.py:90-96 semantics).
#include <cuda.h>
#include <cstdio>

#include "cd_kernels.cuh"

namespace cd {

namespace {

constexpr int kThreads = 192;  // warp 0: loads, warp 1: MMA issue + TMEM alloc, warps 2-5: epilogue
constexpr int kMaxStages = 4;

struct MmaParams {
    Geom g;
    const __half* src[2];
    int src_chunks[2];      // 8-channel chunks per source
    int nchunks;            // total
    const __half* w;        // [tap][kc][n][8]
    const float* bias;
    __half* dst;
    double* stats;
    int Cout;
    int ntaps;              // 27 or 1
    int MT;                 // 128-row sub-tiles per CTA tile
    int wrows;              // rows per smem window
    int nwin;               // 3 (3x3x3) or 1
    int halo;               // Rp+1 or 0
    int nstages;
    int stage_bytes;        // one tap of weights
    long long nrows;        // B * rows_per_sample
    int ntiles;
    int tmem_cols;
    uint32_t idesc;
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
        if (clock64() - t0 > 4000000000LL) { printf("cd_conv_mma: mbarrier timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x); __trap(); }
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
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct SmemCtl {
    uint64_t a_full, a_empty, tmem_full, tmem_empty;
    uint64_t w_full[kMaxStages], w_empty[kMaxStages];
    uint32_t tmem_base;
    int bmin, bmax;
    int pad;
};

__global__ void __launch_bounds__(kThreads) conv_mma_kernel(const MmaParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    SmemCtl* ctl = reinterpret_cast<SmemCtl*>(smem);
    // layout: [ctl 256 B][stats scratch][weight ring][A windows]
    float* red = reinterpret_cast<float*>(smem + 256);                 // [128][Cout+1] + [4][Cout][2]
    const int red_bytes = ((128 * (p.Cout + 1) + 8 * p.Cout) * 4 + 127) / 128 * 128;
    unsigned char* wring = smem + 256 + red_bytes;
    unsigned char* abuf = wring + (size_t)p.nstages * p.stage_bytes;
    const uint32_t wring_s = smem_u32(wring), abuf_s = smem_u32(abuf);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const Geom& g = p.g;

    if (threadIdx.x == 0) {
        mbar_init(smem_u32(&ctl->a_full), 1);
        mbar_init(smem_u32(&ctl->a_empty), 1);
        mbar_init(smem_u32(&ctl->tmem_full), 1);
        mbar_init(smem_u32(&ctl->tmem_empty), 4);
        for (int s = 0; s < p.nstages; ++s) { mbar_init(smem_u32(&ctl->w_full[s]), 1); mbar_init(smem_u32(&ctl->w_empty[s]), 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&ctl->tmem_base)), "r"(p.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = ctl->tmem_base;

    const int tile_rows = 128 * p.MT;
    const uint32_t win_bytes = (uint32_t)p.wrows * 16u;
    const uint32_t chunk_bytes = win_bytes * p.nwin;   // smem bytes per 8-channel chunk

    if (warp == 0) {
        // ===================== producer: activation windows + weight ring =====================
        if (lane == 0) {
            uint32_t it = 0, wcount = 0;
            for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
                const long long m0 = (long long)tile * tile_rows;
                mbar_wait(smem_u32(&ctl->a_empty), (it & 1) ^ 1);
                mbar_expect_tx(smem_u32(&ctl->a_full), chunk_bytes * p.nchunks);
                int ch = 0;
                for (int s = 0; s < 2; ++s) {
                    for (int k = 0; k < p.src_chunks[s]; ++k, ++ch) {
                        const __half* plane0 = p.src[s] + plcp_offset(g, k, 0);
                        for (int w = 0; w < p.nwin; ++w) {
                            long long r0 = m0 + (long long)(w - (p.nwin >> 1)) * g.plane - p.halo;
                            bulk_g2s(abuf_s + ch * chunk_bytes + w * win_bytes, plane0 + r0 * 8, win_bytes, smem_u32(&ctl->a_full));
                        }
                    }
                }
                for (int t = 0; t < p.ntaps; ++t, ++wcount) {
                    const uint32_t s = wcount % p.nstages, ph = (wcount / p.nstages) & 1;
                    mbar_wait(smem_u32(&ctl->w_empty[s]), ph ^ 1);
                    mbar_expect_tx(smem_u32(&ctl->w_full[s]), p.stage_bytes);
                    bulk_g2s(wring_s + s * p.stage_bytes, reinterpret_cast<const unsigned char*>(p.w) + (size_t)t * p.stage_bytes,
                             p.stage_bytes, smem_u32(&ctl->w_full[s]));
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (one thread) =====================
        if (lane == 0) {
            uint32_t it = 0, wcount = 0;
            const int k16 = p.nchunks >> 1;
            for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
                mbar_wait(smem_u32(&ctl->tmem_empty), (it & 1) ^ 1);
                mbar_wait(smem_u32(&ctl->a_full), it & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int t = 0; t < p.ntaps; ++t, ++wcount) {
                    const uint32_t s = wcount % p.nstages, ph = (wcount / p.nstages) & 1;
                    mbar_wait(smem_u32(&ctl->w_full[s]), ph);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    int win = 0, shift = 0;
                    if (p.ntaps == 27) {
                        const int a = t / 9, b = (t / 3) % 3, c = t % 3;
                        win = a;
                        shift = (b - 1) * g.Rp + (c - 1);
                    }
                    const uint32_t a_row0 = abuf_s + win * win_bytes + (uint32_t)(p.halo + shift) * 16u;
                    const uint32_t b0 = wring_s + s * p.stage_bytes;
                    for (int j = 0; j < p.MT; ++j) {
                        for (int q = 0; q < k16; ++q) {
                            const uint64_t ad = make_desc(a_row0 + (uint32_t)(2 * q) * chunk_bytes + (uint32_t)j * 2048u, chunk_bytes, 128u);
                            const uint64_t bd = make_desc(b0 + (uint32_t)(2 * q) * p.Cout * 16u, (uint32_t)p.Cout * 16u, 128u);
                            umma_f16(tmem_base + (uint32_t)(j * p.Cout), ad, bd, p.idesc, (t | q) != 0);
                        }
                    }
                    umma_commit(smem_u32(&ctl->w_empty[s]));
                }
                umma_commit(smem_u32(&ctl->a_empty));
                umma_commit(smem_u32(&ctl->tmem_full));
            }
        }
    } else {
        // ===================== epilogue: TMEM -> registers -> global (+ GroupNorm statistics) =====================
        const int ew = warp & 3;                 // TMEM lane quarter this warp may read
        const int et = ew * 32 + lane;           // row within the 128-row sub-tile
        const int epi_tid = threadIdx.x - 64;    // 0..127
        const int pitch = p.Cout + 1;
        float* part = red + 128 * pitch;         // [4][Cout][2]
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
            const long long m0 = (long long)tile * tile_rows;
            mbar_wait(smem_u32(&ctl->tmem_full), it & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            for (int j = 0; j < p.MT; ++j) {
                const long long row = m0 + (long long)j * 128 + et;
                RowInfo ri;
                ri.kind = 2; ri.b = 0;
                if (row < p.nrows) ri = decode_row(g, row);
                const bool valid = ri.kind == 0;
                if (p.stats) {
                    if (epi_tid == 0) { ctl->bmin = 0x7fffffff; ctl->bmax = -1; }
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    if (valid) { atomicMin(&ctl->bmin, ri.b); atomicMax(&ctl->bmax, ri.b); }
                }
                for (int c0 = 0; c0 < p.Cout; c0 += 32) {
                    uint32_t r[32];
                    tmem_ld32(tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)(j * p.Cout + c0), r);
                    float v[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) + (p.bias ? __ldg(p.bias + c0 + i) : 0.f);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        Half8 h = pack8(v + 8 * k);
                        if (valid) *reinterpret_cast<Half8*>(p.dst + plcp_offset(g, (c0 >> 3) + k, row)) = h;
                        if (p.stats) unpack8(h, v + 8 * k);   // statistics of the values as stored
                    }
                    if (p.stats) {
#pragma unroll
                        for (int i = 0; i < 32; ++i) red[et * pitch + c0 + i] = valid ? v[i] : 0.f;
                    }
                }
                if (p.stats) {
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    const int bmin = ctl->bmin, bmax = ctl->bmax;
                    const bool uniform = bmax >= 0 && bmin == bmax;
                    if (uniform) {
                        // one sample in this sub-tile: column sums by 4 row-quarters
                        for (int c = lane; c < p.Cout; c += 32) {
                            float s = 0.f, q = 0.f;
                            for (int rr = ew * 32; rr < ew * 32 + 32; ++rr) { float x = red[rr * pitch + c]; s += x; q += x * x; }
                            part[(ew * p.Cout + c) * 2] = s;
                            part[(ew * p.Cout + c) * 2 + 1] = q;
                        }
                    }
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    if (uniform) {
                        // ... then one atomic pair per channel
                        for (int c = epi_tid; c < p.Cout; c += 128) {
                            float s = 0.f, q = 0.f;
                            for (int w4 = 0; w4 < 4; ++w4) { s += part[(w4 * p.Cout + c) * 2]; q += part[(w4 * p.Cout + c) * 2 + 1]; }
                            double* d = p.stats + ((long long)bmin * p.Cout + c) * 2;
                            atomicAdd(d, (double)s);
                            atomicAdd(d + 1, (double)q);
                        }
                    } else if (bmax >= 0 && valid) {
                        // sub-tile straddles two samples (rare): per-row atomics
                        for (int c = 0; c < p.Cout; ++c) {
                            float x = red[et * pitch + c];
                            double* d = p.stats + ((long long)ri.b * p.Cout + c) * 2;
                            atomicAdd(d, (double)x);
                            atomicAdd(d + 1, (double)(x * x));
                        }
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&ctl->tmem_empty));
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
    }
}

}  // namespace

int launch_conv_mma(const ConvParams& cp, const __half* packed_w, cudaStream_t st) {
    const int taps = cp.kz * cp.kp * cp.kr;
    if (cp.transposed || cp.sz != 1 || cp.sp != 1 || cp.sr != 1) return 1;
    if (!(taps == 27 && cp.kz == 3 && cp.kp == 3 && cp.kr == 3 && cp.pz == 1 && cp.cp == 1 && cp.pr == 1) && taps != 1) return 1;
    if (cp.Cout % 32 != 0 || cp.Cout > 256 || cp.Cin_total % 16 != 0) return 1;
    if (cp.gin.Z != cp.gout.Z || cp.gin.P != cp.gout.P || cp.gin.R != cp.gout.R) return 1;
    for (int s = 0; s < cp.nsrc; ++s) if (cp.src[s].C % 8 != 0) return 1;

    static int num_sms = 0, max_smem = 0;
    if (!num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        cudaFuncSetAttribute(conv_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
    }
    MmaParams p;
    p.g = cp.gout;
    p.src[0] = cp.src[0].base; p.src_chunks[0] = cp.src[0].C / 8;
    p.src[1] = cp.nsrc > 1 ? cp.src[1].base : nullptr; p.src_chunks[1] = cp.nsrc > 1 ? cp.src[1].C / 8 : 0;
    p.nchunks = cp.Cin_total / 8;
    if (p.src_chunks[0] + p.src_chunks[1] != p.nchunks) return 1;
    p.w = packed_w; p.bias = cp.bias; p.dst = cp.dst; p.stats = cp.stats;
    p.Cout = cp.Cout; p.ntaps = taps;
    p.nwin = taps == 27 ? 3 : 1;
    p.halo = taps == 27 ? cp.gout.Rp + 1 : 0;
    p.stage_bytes = cp.Cin_total * cp.Cout * 2;
    p.nrows = (long long)cp.B * cp.gout.rows_per_sample;
    const int red_bytes = ((128 * (cp.Cout + 1) + 8 * cp.Cout) * 4 + 127) / 128 * 128;
    // pick MT (sub-tiles per CTA) and ring depth: prefer >= 2 CTAs per SM, then the largest tile
    int best_mt = 0, best_stages = 0;
    size_t best_smem = 0;
    for (int pass = 0; pass < 2 && !best_mt; ++pass) {
        const size_t budget = pass == 0 ? (size_t)(max_smem / 2 - 2048) : (size_t)max_smem;
        for (int mt = 4; mt >= 1; mt >>= 1) {
            if (mt * cp.Cout > 256) continue;
            const size_t a_bytes = (size_t)p.nchunks * p.nwin * (128 * mt + 2 * p.halo) * 16;
            const size_t fixed = 256 + red_bytes + a_bytes;
            if (fixed + 2 * (size_t)p.stage_bytes > budget) continue;
            int stages = (int)((budget - fixed) / p.stage_bytes);
            if (stages > kMaxStages) stages = kMaxStages;
            if (stages > taps) stages = taps;
            if (stages < 1) continue;
            best_mt = mt; best_stages = stages; best_smem = fixed + (size_t)stages * p.stage_bytes;
            break;
        }
    }
    if (!best_mt) return 1;
    p.MT = best_mt; p.nstages = best_stages;
    p.wrows = 128 * p.MT + 2 * p.halo;
    const long long tile_rows = 128LL * p.MT;
    p.ntiles = (int)((p.nrows + tile_rows - 1) / tile_rows);
    int cols = p.MT * cp.Cout, tc = 32;
    while (tc < cols) tc <<= 1;
    p.tmem_cols = tc;
    p.idesc = (1u << 4) | ((uint32_t)(cp.Cout >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);  // f16 x f16 -> f32, K-major A and B
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, conv_mma_kernel, kThreads, best_smem);
    if (occ < 1) return 1;
    if (occ * tc > 512) occ = 512 / tc;   // TMEM columns per SM
    int grid = num_sms * occ;
    if (grid > p.ntiles) grid = p.ntiles;
    conv_mma_kernel<<<grid, kThreads, best_smem, st>>>(p);
    return 0;
}

}  // namespace cd
