// tcgen05 / TMEM gather implicit-GEMM convolution for the CylindricalConv variants whose taps are NOT constant
// row shifts of one PLCP tensor: the strided Downsample conv (k=(3,4,4), s=2, models/models.py:360-369), the
// transposed Upsample conv (CylindricalConvTrans, k=(3|4,4,4), s=2, padding (1,3,1), models/models.py:25-62,
// 335-357) and the thin-input init conv (C_in = 3|4 padded to 16).
//
// GEMM view: outputs are enumerated densely per "class" (one class for forward convs; the 8 output-parity classes
// of a stride-2 transposed conv, each with its own <= 2x2x2 tap subset):  D[m, co] = sum_t sum_ci A_t[m, ci] W_t[ci, co],
// where row m of A_t is the input voxel that tap t of output m reads (zero if it falls in the Z/R zero padding;
// phi wraps circularly).  Producer warps gather those 16-byte (8-channel) pieces with vector loads and build the
// canonical no-swizzle K-major operand tile in shared memory (a ring of stages of G taps), one elected thread
// streams the matching weight taps with cp.async.bulk, one thread issues tcgen05.mma into a double-buffered TMEM
// accumulator, and four epilogue warps add the bias, round to fp16 and scatter the rows (and their circular phi
// halo copies) into the output PLCP tensor.
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "cd_kernels.cuh"

namespace cd {

namespace {

constexpr int kGThreads = 14 * 32;  // warp 0: weights, warp 1: MMA + TMEM alloc, warps 2-5: epilogue, warps 6-13: gather
constexpr int kGMaxStages = 6;

struct GTap { signed char dz, dq, dr, pad; int widx; int delta; };   // delta: input row offset (dz*Pp + dq)*Rp + dr of the tap
struct GClass {
    int Zc, Pc, Rc;       // class-local output extents
    int ez, ep, er;       // output coordinate = local * om + e
    int tap0, ntaps;
    int tile0;            // first tile of this class
    long long rows;       // B * Zc * Pc * Rc
};

struct GatherParams {
    Geom gin, gout;
    const __half* src;
    int src_chunks;       // real 8-channel chunks of the source (chunks beyond are zero-filled)
    const __half* w;      // [tap][kc][n][8] fp16
    const float* bias;
    __half* dst;
    int Cout;
    int N;                // MMA width: Cout, or 2*Cout when a tile row produces the output pair (r, r+1) (transposed form)
    int pair_r;           // 1: columns [Cout, 2*Cout) are the output voxel r+1 (stored to the next PLCP row if r+1 < R)
    int mz, mp, mr;       // input index = local * m + tap offset
    int oz, op, orr;      // output index = local * o + class offset
    int nclass;
    GClass cls[8];
    GTap taps[64];
    int G;                // taps per stage
    int nstages;
    int a_stage_bytes, w_tap_bytes, w_stage_bytes;
    int ntiles;
    int write_halo;
    uint32_t idesc;
};

__device__ __forceinline__ uint32_t g_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void g_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void g_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void g_mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool g_mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void g_mbar_wait(uint32_t bar, uint32_t parity) {
    if (g_mbar_try(bar, parity)) return;
    long long t0 = clock64();
    while (!g_mbar_try(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) { printf("cd_conv_gather: mbarrier timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x); __trap(); }
    }
}
// waits of the warps that are not on the MMA issue path (weights, gather producers, epilogue): back off between polls.  The
// tight form costs ~10 instructions per poll on every lane; ncu showed a third of this kernel's issue slots going to such loops.
__device__ __forceinline__ void g_mbar_wait_relaxed(uint32_t bar, uint32_t parity) {
    if (g_mbar_try(bar, parity)) return;
    long long t0 = clock64();
    while (!g_mbar_try(bar, parity)) {
        __nanosleep(64);
        if (clock64() - t0 > 4000000000LL) { printf("cd_conv_gather: mbarrier timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x); __trap(); }
    }
}
__device__ __forceinline__ void g_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// 16-byte asynchronous copy global -> shared; src_bytes = 0 writes zeros.  Completion is signalled on an mbarrier by
// cp.async.mbarrier.arrive.noinc (counts as one of the barrier's expected arrivals once this thread's copies have landed).
__device__ __forceinline__ void g_cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void g_cp_async_arrive(uint32_t bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint64_t g_make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
__device__ __forceinline__ void g_umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
// warp-convergent forms (every lane executes the call, one elected lane performs the operation): with the issue loop walked by
// all 32 lanes the descriptor arithmetic stays in the uniform datapath; issued from divergent single-lane code every MMA costs a
// ~20-instruction R2UR waterfall (measured in cd_conv_mma.cu: 164 clk per MMA against a 44-56 clk tensor-pipe slot)
__device__ __forceinline__ void g_umma_f16_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n.reg .pred p, e;\nelect.sync _|e, 0xffffffff;\nsetp.ne.b32 p, %4, 0;\n@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void g_umma_commit_elect(uint32_t bar) {
    asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\n@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}"
                 ::"r"(bar) : "memory");
}
__device__ __forceinline__ void g_umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void g_tmem_ld16(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

struct GCtl {
    uint64_t full[kGMaxStages], empty[kGMaxStages];
    uint64_t t_full[2], t_empty[2];
    uint32_t tmem_base;
    uint32_t pad;
};
constexpr int kGCtlBytes = 256;
static_assert(sizeof(GCtl) <= kGCtlBytes, "ctl block too small");

struct RowCoord { int b, z, p, r; bool ok; };

__device__ __forceinline__ int find_class(const GatherParams& p, int tile) {
    int c = 0;
#pragma unroll
    for (int k = 1; k < 8; ++k) if (k < p.nclass && tile >= p.cls[k].tile0) c = k;
    return c;
}
__device__ __forceinline__ RowCoord class_coord(const GClass& c, int tile, int i) {
    RowCoord rc;
    const long long L = (long long)(tile - c.tile0) * 128 + i;
    rc.ok = L < c.rows;
    const long long per = (long long)c.Zc * c.Pc * c.Rc;
    const long long Lc = rc.ok ? L : 0;
    rc.b = (int)(Lc / per);
    int rem = (int)(Lc - (long long)rc.b * per);
    rc.z = rem / (c.Pc * c.Rc);
    rem -= rc.z * c.Pc * c.Rc;
    rc.p = rem / c.Rc;
    rc.r = rem - rc.p * c.Rc;
    return rc;
}

template <int NCH>   // padded 8-channel chunks of the input (even); G = 16 / NCH taps per stage
__global__ void __launch_bounds__(kGThreads, 1) conv_gather_kernel(const __grid_constant__ GatherParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int G = 16 / NCH;
    GCtl* ctl = reinterpret_cast<GCtl*>(smem);
    unsigned char* abuf = smem + kGCtlBytes;
    unsigned char* wbuf = abuf + (size_t)p.nstages * p.a_stage_bytes;
    const uint32_t abuf_s = g_smem_u32(abuf), wbuf_s = g_smem_u32(wbuf);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int N = p.N;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kGMaxStages; ++s) { g_mbar_init(g_smem_u32(&ctl->full[s]), 129); g_mbar_init(g_smem_u32(&ctl->empty[s]), 1); }
        for (int s = 0; s < 2; ++s) { g_mbar_init(g_smem_u32(&ctl->t_full[s]), 1); g_mbar_init(g_smem_u32(&ctl->t_empty[s]), 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(g_smem_u32(&ctl->tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp == 0) {
        // ===================== weight taps: cp.async.bulk into the stage ring =====================
        if (lane == 0) {
            uint32_t slot = 0, ph = 0;
            for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
                const GClass& c = p.cls[find_class(p, tile)];
                const int nst = (c.ntaps + G - 1) / G;
                for (int st = 0; st < nst; ++st, slot = (slot + 1 == (uint32_t)p.nstages) ? 0u : slot + 1, ph ^= (uint32_t)(slot == 0)) {
                    const int t0 = st * G;
                    const int nt = c.ntaps - t0 < G ? c.ntaps - t0 : G;
                    g_mbar_wait_relaxed(g_smem_u32(&ctl->empty[slot]), ph ^ 1);
                    g_mbar_expect_tx(g_smem_u32(&ctl->full[slot]), (uint32_t)nt * p.w_tap_bytes);
                    for (int gi = 0; gi < nt; ++gi)
                        g_bulk_g2s(wbuf_s + slot * p.w_stage_bytes + gi * p.w_tap_bytes,
                                   reinterpret_cast<const unsigned char*>(p.w) + (size_t)p.taps[c.tap0 + t0 + gi].widx * p.w_tap_bytes,
                                   p.w_tap_bytes, g_smem_u32(&ctl->full[slot]));
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (all lanes walk the loop, the elected lane issues) =====================
        {
            uint32_t slot = 0, ph = 0, tcount = 0;              // stage ring position as counters (no runtime % and /)
            const uint64_t adesc_hi = g_make_desc(0, 2048u, 128u);
            const uint64_t bdesc_hi = g_make_desc(0, (uint32_t)N * 16u, 128u);
            const uint32_t bstep = (2u * N * 16u) >> 4;
            for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++tcount) {
                const GClass& c = p.cls[find_class(p, tile)];
                const int nst = (c.ntaps + G - 1) / G;
                const uint32_t acc = tcount & 1, aph = (tcount >> 1) & 1;
                const uint32_t dcol = tmem_base + acc * 256;
                g_mbar_wait(g_smem_u32(&ctl->t_empty[acc]), aph ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                uint32_t accum = 0;
                for (int st = 0; st < nst; ++st) {
                    const int t0 = st * G;
                    const int nt = c.ntaps - t0 < G ? c.ntaps - t0 : G;
                    g_mbar_wait(g_smem_u32(&ctl->full[slot]), ph);
                    // the tile was written through the generic proxy (cp.async); the MMA reads it through the async proxy
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    uint32_t al = (abuf_s + slot * p.a_stage_bytes) >> 4, bl = (wbuf_s + slot * p.w_stage_bytes) >> 4;
                    for (int e = 0; e < nt * (NCH / 2); ++e, al += 256u, bl += bstep) {   // stage layouts are contiguous in (tap, k16)
                        g_umma_f16_elect(dcol, adesc_hi | al, bdesc_hi | bl, p.idesc, accum);
                        accum = 1;
                    }
                    g_umma_commit_elect(g_smem_u32(&ctl->empty[slot]));
                    slot = (slot + 1 == (uint32_t)p.nstages) ? 0u : slot + 1;
                    ph ^= (uint32_t)(slot == 0);
                }
                g_umma_commit_elect(g_smem_u32(&ctl->t_full[acc]));
            }
        }
    } else if (warp < 6) {
        // ===================== epilogue: TMEM -> fp16 rows scattered into the output PLCP tensor =====================
        const int q = warp & 3;
        const int i = q * 32 + lane;
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++tcount) {
            const GClass& c = p.cls[find_class(p, tile)];
            const RowCoord rc = class_coord(c, tile, i);
            const int zo = rc.z * p.oz + c.ez, po = rc.p * p.op + c.ep, ro = rc.r * p.orr + c.er;
            const long long row = interior_row(p.gout, rc.b, zo, po, ro);
            long long mirror = -1;
            if (p.write_halo) {
                if (po == 0) mirror = row + (long long)p.gout.P * p.gout.Rp;
                else if (po == p.gout.P - 1) mirror = row - (long long)p.gout.P * p.gout.Rp;
            }
            // P == 1: p = 0 is both edges; the single phi row mirrors to both halo rows
            const long long mirror2 = (p.write_halo && p.gout.P == 1) ? row - (long long)p.gout.Rp : -1;
            const uint32_t acc = tcount & 1, aph = (tcount >> 1) & 1;
            g_mbar_wait_relaxed(g_smem_u32(&ctl->t_full[acc]), aph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * 256;
            const bool ok2 = rc.ok && p.pair_r && (ro + 1 < p.gout.R);     // second voxel of the pair exists (odd R: last one does not)
            for (int c0 = 0; c0 < N; c0 += 16) {
                float v[16];
                g_tmem_ld16(taddr + c0, v);
                const int half = c0 >= p.Cout ? 1 : 0;              // which voxel of the pair these columns belong to
                const int cc = c0 - half * p.Cout;
                if (p.bias) {
#pragma unroll
                    for (int k = 0; k < 16; ++k) v[k] += __ldg(p.bias + cc + k);
                }
                if (half ? ok2 : rc.ok) {
#pragma unroll
                    for (int k = 0; k < 2; ++k) {
                        const Half8 h = pack8(v + 8 * k);
                        __half* o = p.dst + plcp_offset(p.gout, (cc >> 3) + k, row + half);
                        store8(o, h);
                        if (mirror >= 0) store8(o + (mirror - row) * 8, h);
                        if (mirror2 >= 0) store8(o + (mirror2 - row) * 8, h);
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) g_mbar_arrive(g_smem_u32(&ctl->t_empty[acc]));
        }
    } else {
        // ===================== gather producers: two groups of 128 threads alternate stages =====================
        const int pg = (warp - 6) >> 2;
        const int i = ((warp - 6) & 3) * 32 + lane;
        uint32_t n = 0, rslot = 0, rph = 0;          // stage counter and its ring position (counters: no runtime % and /)
        for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
            const GClass& c = p.cls[find_class(p, tile)];
            const int nst = (c.ntaps + G - 1) / G;
            const RowCoord rc = class_coord(c, tile, i);
            const int zb = rc.z * p.mz, qb = rc.p * p.mp, rb = rc.r * p.mr;
            // 16-byte piece of chunk plane 0 at the tile row's base input voxel (padded phi index qb); taps add a row delta
            const uint4* bp = reinterpret_cast<const uint4*>(
                p.src + plcp_offset(p.gin, 0, (long long)rc.b * p.gin.rows_per_sample + ((long long)zb * p.gin.Pp + qb) * p.gin.Rp + rb));
            const uint4* bp0 = reinterpret_cast<const uint4*>(p.src + plcp_offset(p.gin, 0, 0));   // always-valid address for zero fills
            const int wrap = p.gin.P * p.gin.Rp;
            const long long cs = p.gin.chunk_stride;     // uint4 units between chunk planes
            for (int st = 0; st < nst; ++st, ++n, rslot = (rslot + 1 == (uint32_t)p.nstages) ? 0u : rslot + 1, rph ^= (uint32_t)(rslot == 0)) {
                if ((int)(n & 1) != pg) continue;
                const uint32_t slot = rslot, ph = rph;
                const int t0 = st * G;
                // asynchronous 16-byte gathers straight into the operand tile (zero-filled when the tap falls in the zero
                // padding): nothing is staged in registers, so a thread keeps as many stages in flight as the ring is deep
                g_mbar_wait_relaxed(g_smem_u32(&ctl->empty[slot]), ph ^ 1);
                const uint32_t dst_s = abuf_s + slot * p.a_stage_bytes + (uint32_t)i * 16u;
#pragma unroll
                for (int gi = 0; gi < G; ++gi) {
                    const bool tv = (t0 + gi) < c.ntaps;
                    const GTap tp = p.taps[c.tap0 + (tv ? t0 + gi : 0)];
                    const int qq = qb + tp.dq;
                    const bool ok = rc.ok && tv && (unsigned)(zb + tp.dz) < (unsigned)p.gin.Z && (unsigned)qq < (unsigned)(p.gin.P + 2) &&
                                    (unsigned)(rb + tp.dr) < (unsigned)p.gin.R;
                    // circular phi: the halo positions 0 and P+1 read the interior copies P and 1
                    const int delta = tp.delta + (qq == 0 ? wrap : 0) - (qq == p.gin.P + 1 ? wrap : 0);
                    const uint4* q = ok ? bp + delta : bp0;
#pragma unroll
                    for (int k = 0; k < NCH; ++k) {
                        const bool ld = ok && k < p.src_chunks;
                        g_cp_async16(dst_s + (uint32_t)(gi * NCH + k) * 2048u, ld ? q + k * cs : bp0, ld ? 16u : 0u);
                    }
                }
                g_cp_async_arrive(g_smem_u32(&ctl->full[slot]));
            }
        }
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
bool gather_conv_supported(int Cin, int Cout) {
    const int nch = (Cin + 15) / 16 * 2;   // padded chunks
    return (nch == 2 || nch == 4 || nch == 8 || nch == 16) && Cout % 16 == 0 && Cout <= 256 && Cin % 8 == 0;
}

size_t gather_packed_elems(int Cin, int Cout, int taps) { return (size_t)taps * ((Cin + 15) / 16 * 16) * Cout; }

// w: fp32 [tap][ci][co] -> fp16 [tap][kc][n][8] with the input channels zero-padded to a multiple of 16
void gather_pack_weights(const float* w, int Cin, int Cout, int taps, __half* out) {
    const int Cp = (Cin + 15) / 16 * 16, nkc = Cp / 8;
    for (int t = 0; t < taps; ++t)
        for (int kc = 0; kc < nkc; ++kc)
            for (int n = 0; n < Cout; ++n)
                for (int j = 0; j < 8; ++j) {
                    const int ci = kc * 8 + j;
                    out[(((size_t)t * nkc + kc) * Cout + n) * 8 + j] =
                        __float2half_rn(ci < Cin ? w[((size_t)t * Cin + ci) * Cout + n] : 0.f);
                }
}

// Transposed stride-2 conv with k_r = 4, R padding 1 (CylindricalConvTrans, models.py:25-62): one tile row produces the
// output pair r = 2r', 2r'+1 from the input offsets dr = -1, 0, +1; merged tap (a, b, dr) has the [ci][2*Cout] matrix
//   columns [0,Cout):      original tap c = 1 - 2dr  (output r = 2r',   dr in {-1, 0})
//   columns [Cout,2Cout):  original tap c = 2 - 2dr  (output r = 2r'+1, dr in { 0, 1})
// packed as [(a*kp + b)*3 + dr+1][kc][n2][8] fp16.
size_t gather_packed_elems_pairT(int Cin, int Cout, int kz, int kp) { return (size_t)kz * kp * 3 * ((Cin + 15) / 16 * 16) * 2 * Cout; }
void gather_pack_weights_pairT(const float* w, int Cin, int Cout, int kz, int kp, int kr, __half* out) {
    const int Cp = (Cin + 15) / 16 * 16, nkc = Cp / 8, N2 = 2 * Cout;
    for (int a = 0; a < kz; ++a)
        for (int b = 0; b < kp; ++b)
            for (int di = 0; di < 3; ++di) {
                const int dr = di - 1;
                const size_t t = ((size_t)a * kp + b) * 3 + di;
                for (int kc = 0; kc < nkc; ++kc)
                    for (int n2 = 0; n2 < N2; ++n2) {
                        const int er = n2 / Cout, co = n2 - er * Cout;
                        const int c = er + 1 - 2 * dr;
                        for (int j = 0; j < 8; ++j) {
                            const int ci = kc * 8 + j;
                            float v = 0.f;
                            if (c >= 0 && c < kr && ci < Cin) v = w[((size_t)((a * kp + b) * kr + c) * Cin + ci) * Cout + co];
                            out[((t * nkc + kc) * N2 + n2) * 8 + j] = __float2half_rn(v);
                        }
                    }
            }
}

template <int NCH>
static cudaError_t launch_ginst(const GatherParams& p, int grid, size_t smem, int max_smem, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(conv_gather_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    conv_gather_kernel<NCH><<<grid, kGThreads, smem, st>>>(p);
    return cudaSuccess;
}

int launch_conv_gather(const ConvParams& cp, const __half* packed_w, cudaStream_t st) {
    if (cp.nsrc != 1) return 1;
    if (!gather_conv_supported(cp.Cin_total, cp.Cout)) return 1;
    const int taps = cp.kz * cp.kp * cp.kr;
    if (!(cp.sz == cp.sp && cp.sp == cp.sr && (cp.sz == 1 || cp.sz == 2))) return 1;
    if (cp.transposed && cp.sz != 2) return 1;
    if (taps > 64) return 1;
    static int num_sms = 0, max_smem = 0;
    if (!num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    }
    GatherParams p;
    memset(&p, 0, sizeof(p));
    p.gin = cp.gin; p.gout = cp.gout;
    p.src = cp.src[0].base; p.src_chunks = cp.src[0].C / 8;
    p.w = packed_w; p.bias = cp.bias; p.dst = cp.dst; p.Cout = cp.Cout;
    p.write_halo = cp.write_halo;
    const int nch = (cp.Cin_total + 15) / 16 * 2;
    const int s = cp.sz;
    int ntap = 0, tile0 = 0;
    if (!cp.transposed) {
        // zi = zo*s + a - pz ; q = po*s + b + (1 - cp) (index into the 1-padded circular axis) ; ri = ro*s + c - pr
        p.mz = p.mp = p.mr = s; p.oz = p.op = p.orr = 1;
        p.nclass = 1;
        GClass& c = p.cls[0];
        c.Zc = cp.gout.Z; c.Pc = cp.gout.P; c.Rc = cp.gout.R; c.ez = c.ep = c.er = 0;
        c.tap0 = 0; c.tile0 = 0;
        for (int a = 0; a < cp.kz; ++a)
            for (int b = 0; b < cp.kp; ++b)
                for (int cc = 0; cc < cp.kr; ++cc) {
                    GTap& t = p.taps[ntap];
                    t.dz = (signed char)(a - cp.pz); t.dq = (signed char)(b + 1 - cp.cp); t.dr = (signed char)(cc - cp.pr); t.pad = 0;
                    t.widx = (a * cp.kp + b) * cp.kr + cc;
                    ++ntap;
                }
        c.ntaps = ntap;
        c.rows = (long long)cp.B * c.Zc * c.Pc * c.Rc;
        tile0 = (int)((c.rows + 127) / 128);
    } else {
        // stride-2 transposed conv: output parity classes; zi = z' + (ez + pz - a)/2, q = p' + (ep + kp-1 - b)/2 (padded
        // axis, must stay inside [0, Pin+2)), ri = r' + (er + pr - c)/2, taps restricted to matching parity
        if (cp.kr != 4 || cp.pr != 1 || 2 * cp.Cout > 256) return 1;
        p.mz = p.mp = p.mr = 1; p.oz = p.op = p.orr = 2;
        p.nclass = 0;
        p.pair_r = 1;
        for (int ez = 0; ez < 2; ++ez)
            for (int ep = 0; ep < 2; ++ep) {
                GClass c;
                c.Zc = (cp.gout.Z - ez + 1) / 2; c.Pc = (cp.gout.P - ep + 1) / 2; c.Rc = (cp.gout.R + 1) / 2;
                c.ez = ez; c.ep = ep; c.er = 0; c.tap0 = ntap; c.ntaps = 0; c.tile0 = tile0;
                c.rows = (long long)cp.B * c.Zc * c.Pc * c.Rc;
                if (c.rows <= 0) continue;
                for (int a = 0; a < cp.kz; ++a) {
                    if ((ez + cp.pz - a) & 1) continue;
                    for (int b = 0; b < cp.kp; ++b) {
                        if ((ep + cp.kp - 1 - b) & 1) continue;
                        for (int di = 0; di < 3; ++di) {
                            if (ntap >= 64) return 1;
                            GTap& t = p.taps[ntap];
                            // arithmetic shift keeps floor semantics for the (even) negative numerators
                            t.dz = (signed char)((ez + cp.pz - a) / 2); t.dq = (signed char)((ep + cp.kp - 1 - b) / 2);
                            t.dr = (signed char)(di - 1); t.pad = 0;
                            t.widx = (a * cp.kp + b) * 3 + di;
                            ++ntap; ++c.ntaps;
                        }
                    }
                }
                if (c.ntaps == 0) return 1;   // an output class without taps would need a bias-only path
                tile0 += (int)((c.rows + 127) / 128);
                p.cls[p.nclass++] = c;
            }
    }
    for (int t = 0; t < ntap; ++t)
        p.taps[t].delta = ((int)p.taps[t].dz * cp.gin.Pp + (int)p.taps[t].dq) * cp.gin.Rp + (int)p.taps[t].dr;
    p.ntiles = tile0;
    if (p.ntiles < 1) return 1;
    p.N = p.pair_r ? 2 * cp.Cout : cp.Cout;
    p.G = 16 / nch;
    p.a_stage_bytes = 16 * 2048;
    p.w_tap_bytes = nch * 8 * p.N * 2;
    p.w_stage_bytes = p.G * p.w_tap_bytes;
    int stages = (max_smem - kGCtlBytes - 256) / (p.a_stage_bytes + p.w_stage_bytes);
    if (stages > kGMaxStages) stages = kGMaxStages;
    // The strided conv fetches every input piece ~6 times (once per (output, tap) pair that reads it) with cp.async.ca: a ring of
    // 4 stages instead of 6 leaves ~70 KB more of the SM's L1 for those repeats (DS2 L0->L1: 0.489 -> 0.296 ms, DS3 0.644 -> 0.552;
    // 5 stages: no gain - the shared-memory carve-out moves in steps; the transposed form is neutral to slightly faster)
    if (stages > 4) stages = 4;
    if (const char* e = getenv("CD_GATHER_STAGES")) { const int v = atoi(e); if (v >= 2 && v < stages) stages = v; }   // diagnostics
    if (stages < 2) return 1;
    p.nstages = stages;
    const size_t smem = kGCtlBytes + (size_t)stages * (p.a_stage_bytes + p.w_stage_bytes) + 128;
    p.idesc = (1u << 4) | ((uint32_t)(p.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    int grid = num_sms < p.ntiles ? num_sms : p.ntiles;
    cudaError_t e;
    switch (nch) {
        case 2: e = launch_ginst<2>(p, grid, smem, max_smem, st); break;
        case 4: e = launch_ginst<4>(p, grid, smem, max_smem, st); break;
        case 8: e = launch_ginst<8>(p, grid, smem, max_smem, st); break;
        case 16: e = launch_ginst<16>(p, grid, smem, max_smem, st); break;
        default: return 1;
    }
    return e == cudaSuccess ? 0 : 1;
}

}  // namespace cd
