// tcgen05 / TMEM implicit-GEMM convolution for stride-1 CylindricalConv (3x3x3 and 1x1x1), sm_100a.
//
// GEMM view for the 3x3x3 case ("R-stacked"): for every padded-linear row m of the PLCP tensor (cd_common.cuh)
//     D[m, g*CB + co] = sum over (dz,dp) in {-1,0,1}^2, ci:  X[m + dz*plane + dp*Rp, ci] * W[dz,dp,dr=g-1][ci, co]
//     out[m, co]      = D[m-1, 0*CB+co] + D[m, 1*CB+co] + D[m+1, 2*CB+co] + bias[co]
// i.e. the three R-taps are stacked in the GEMM N dimension (N = 3*CB, CB = block of output channels) and
// recombined in the epilogue with two warp shuffles.  Why: with C_out = 32 a plain K-concatenated implicit GEMM
// issues M128xN32xK16 MMAs whose A operand (4 KB) must be re-read from shared memory for each of the 27 taps;
// measured on B200 (scripts/mma_microbench.cu) such an MMA takes 44 clk instead of its 16 clk tensor floor
// (128 B/clk shared-memory read limit).  Stacking R gives M128xN96xK16 MMAs: 56 clk for 3x the work (85 % of
// the tensor floor), and N=192 for C_out = 64 runs at the floor.
//
//   * A operand: cp.async.bulk (1-D TMA) copies contiguous row windows of each 8-channel chunk plane into shared
//     memory once per tile; every (dz,dp) shift is the same window at a row offset, addressed through the UMMA
//     shared-memory descriptor (no-swizzle K-major: 8 rows x 16 B core matrices, SBO 128 B, LBO = chunk stride).
//     Circular phi padding and zero Z/R padding are the halo rows of the PLCP tensor (models/models.py:90-96).
//   * B operand: packed weights [shift][kchunk][n][8] fp16, resident in shared memory for the whole persistent
//     CTA when they fit (the dominant 32->32 shape: 55 KB), else streamed through an mbarrier ring.
//   * D: TMEM accumulator slots of 128 lanes x N fp32 columns, ring-buffered so the epilogue of one 128-row
//     sub-tile overlaps the MMAs of the next.  Sub-tiles advance by 126 rows (rows 0 and 127 of a sub-tile lack
//     their m-1 / m+1 neighbour).
//   * Sub-tiles are dealt round-robin to G classes (3 for the narrow 3x3x3 forms, else 2): class c = MMA-issuing warp 1 + c, the
//     epilogue group of 4 warps c (one warp per TMEM lane quarter, all channels of the block) and the accumulator slots it owns.
//   * Epilogue: tcgen05.ld of D_0/D_1/D_2, R-recombination by two rotate-shuffles per channel:
//     lane 31 / lane 0 first publish their D_0 / D_2 rows in shared memory and, after one 128-thread named barrier,
//     replace them by the neighbouring warp-quarter's values, so the rotation delivers the right row to every lane
//     without per-element selects or boundary fix-ups; then bias, fp16 rounding, coalesced PLCP stores of the interior
//     rows (optionally also of their circular phi halo copies) and the GroupNorm statistics: per-thread fp32 sum /
//     sum of squares per (sample, group of CPG channels), kept in registers across the CTA's consecutive tiles and
//     flushed with a warp reduction + fp64 atomics on sample change (into the group's first channel slot).
//     The 1x1x1 form has no recombination and no statistics: it streams 16 columns at a time.
#include <cuda.h>
#include <cstdio>
#include <cstdlib>

#include "cd_kernels.cuh"

namespace cd {

#ifdef CD_MMA_DEBUG
__device__ long long g_mma_dbg[16];
#define DBG_T() clock64()
#define DBG_ADD(i, v) dbg_local[i] += (v)
#define DBG_DECL() long long dbg_local[16] = {0}
#define DBG_FLUSH(cond) do { if (cond) for (int i_ = 0; i_ < 16; ++i_) if (dbg_local[i_]) atomicAdd((unsigned long long*)&g_mma_dbg[i_], (unsigned long long)dbg_local[i_]); } while (0)
#define DBG_MODE(bit) (p.dbg_mode & (bit))     // 4: no exchange/shuffles, 8: no global stores, 16: no statistics, 32: no TMEM loads, 64: one issuer, 128: no MMAs
#else
#define DBG_MODE(bit) 0
#define DBG_T() 0LL
#define DBG_ADD(i, v)
#define DBG_DECL()
#define DBG_FLUSH(cond)
#endif

namespace {

#ifndef CD_EPI_SPLIT
#define CD_EPI_SPLIT 0                          // 1: split a block's channels over two sets of epilogue warps (measured slower than a third group, see below)
#endif
#ifndef CD_EPI_GROUPS
#define CD_EPI_GROUPS 3                         // epilogue groups (sub-tiles in flight) of the narrow 3x3x3 forms: 2, 3 or 4
#endif
#ifndef CD_EPI_CW3
#define CD_EPI_CW3 16                           // channels per epilogue pass of the three-group form (8: fewer live registers)
#endif
// Epilogue groups take sub-tiles round-robin.  A sub-tile's epilogue is a ~4000 clk dependent chain against ~1000 clk of MMAs
// (N = 96), so the narrow 3x3x3 forms run three groups (27 warps, 72 registers per thread); N = 192 has two accumulator slots
// and the 1x1x1 form is HBM-bound: two groups.
template <int CB, bool K3> __host__ __device__ constexpr int kGroups() { return (K3 && CB <= 48) ? CD_EPI_GROUPS : 2; }
template <int CB, bool K3> __host__ __device__ constexpr int kEpiWarps() { return kGroups<CB, K3>() * (CD_EPI_SPLIT ? 8 : 4); }   // groups x (1 or 2 channel halves) x 4 lane quarters
// Sub-tile n belongs to class n % G: issued by MMA warp 1 + n % G (one elected thread each), finished by epilogue group n % G,
// in TMEM accumulator slots that class owns.  One producer and one consumer per slot, each walking its uses in order, is what
// makes the parity waits on t_full / t_empty safe (a shared ring lets a fast issuer lap a slow group by a whole phase).
template <int CB, bool K3> __host__ __device__ constexpr int kFirstEpiWarp() { return 1 + kGroups<CB, K3>(); }
template <int CB, bool K3> __host__ __device__ constexpr int kThreads() { return 32 * (kFirstEpiWarp<CB, K3>() + kEpiWarps<CB, K3>()); }   // warp 0: loads, warps 1..G: MMA issue (warp 1 owns TMEM), then the epilogue warps
constexpr int kMaxGroups = CD_EPI_GROUPS > 2 ? CD_EPI_GROUPS : 2;
constexpr int kMaxWStages = 4;
constexpr int kAStages = 4;
constexpr int kMaxAcc = 8;

struct MmaParams {
    Geom g;
    const __half* src[2];
    int src_chunks[2];      // 8-channel chunks per source
    int nchunks;            // Cin / 8
    const __half* w;        // this launch's packed weights: [cb][shift][kc][n][8]
    const float* bias;
    __half* dst;
    double* stats;
    int stats_C;            // channels per sample of the statistics array (>= Cout: this launch may produce a channel slice)
    int Cout, CB, ncb;      // output channels, channels per block, blocks
    int N;                  // GEMM N = 3*CB (3x3x3) or CB (1x1x1)
    int nshift;             // 9 or 1
    int sub_stride;         // 126 or 128
    int MT;                 // sub-tiles per tile
    int tile_rows;          // output rows per tile = sub_stride * MT
    int d_lead;             // D row of sub-tile 0 relative to the tile's first output row: -1 or 0
    int wrows, nwin;        // rows per shared-memory window, windows per chunk (1 = union, 3 = one per dz)
    int row0;               // first window row relative to the tile's first D row (<= 0)
    int win_dz;             // window offset per dz in rows (3-window mode: plane; union mode: 0)
    int off_dz, off_dp;     // row offset inside a window per dz / dp step
    int nacc;               // TMEM accumulator slots
    int resident;           // 1: weights resident in smem, 0: streamed ring
    int w_bytes_cb;         // bytes of one channel block's weights
    int wstage_bytes;       // ring stage = one shift x one K group
    int kgroup_chunks;      // chunks per ring stage
    int nwstages;
    int a_stage_bytes;
    int astages;            // activation window stages (up to 4; 1 when shared memory is short)
    int ksplit, gchunks;    // resident form: a tile's K (input chunks) is staged in ksplit groups of gchunks chunks
    long long nrows;        // B * rows_per_sample
    int ntiles;
    int tiles_per_cta;
    uint32_t idesc;
    int write_halo;         // 1: interior rows with pp == 1 / pp == P are also stored to their phi halo copies
    const __half* addsrc;   // 3x3x3 form: tensor added to the result (interior rows) before rounding and statistics, or nullptr
    const double* pn_stats; const float* pn_wsum; const float* pn_wb; int pn_C; float pn_inv_count;   // folded PreNorm (1x1x1)
    int dbg_mode;           // CD_MMA_DEBUG builds only: 1 = epilogue does nothing but release slots, 2 = load A only once
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
// waits of the warps that are NOT on the MMA issue path (epilogue, producer): back off between polls so that a spinning warp
// does not take issue slots from the MMA-issuing warp sharing its scheduler
#ifndef CD_RELAX_WAIT
#define CD_RELAX_WAIT 1
#endif
__device__ __forceinline__ void mbar_wait_relaxed(uint32_t bar, uint32_t parity) {
#if CD_RELAX_WAIT
    if (mbar_try(bar, parity)) return;
    long long t0 = clock64();
    while (!mbar_try(bar, parity)) {
        __nanosleep(40);
        if (clock64() - t0 > 4000000000LL) { printf("cd_conv_mma: mbarrier timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x); __trap(); }
    }
#else
    mbar_wait(bar, parity);
#endif
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
// warp-convergent forms: every lane executes the call, one elected lane performs the operation.  Keeping the C++ control
// flow convergent lets ptxas hold descriptors / addresses in uniform registers (no per-MMA R2UR waterfall loop).
__device__ __forceinline__ void umma_f16_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n.reg .pred p, e;\nelect.sync _|e, 0xffffffff;\nsetp.ne.b32 p, %4, 0;\n@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit_elect(uint32_t bar) {
    asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\n@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}"
                 ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_elect(uint32_t bar) {
    asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\n@e mbarrier.arrive.shared::cta.b64 _, [%0];\n}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
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

__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld8_issue(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
}
template <int W> __device__ __forceinline__ void tmem_ldw_issue(uint32_t taddr, uint32_t* r) {
    static_assert(W == 8 || W % 16 == 0, "pass width");
    if constexpr (W == 8) tmem_ld8_issue(taddr, r);
    else {
#pragma unroll
        for (int k = 0; k < W; k += 16) tmem_ld16_issue(taddr + k, r + k);
    }
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// predicated 128-bit shared-memory accesses (no divergent branch: `if (lane == 31) {...}` compiles to BSSY / BSYNC regions)
__device__ __forceinline__ void sts128_if(bool pred, uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("{\n.reg .pred q;\nsetp.ne.b32 q, %5, 0;\n@q st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n}"
                 ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d), "r"((uint32_t)pred) : "memory");
}
__device__ __forceinline__ void lds128_if(bool pred, uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {   // a..d kept when !pred
    asm volatile("{\n.reg .pred q;\nsetp.ne.b32 q, %5, 0;\n@q ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];\n}"
                 : "+r"(a), "+r"(b), "+r"(c), "+r"(d) : "r"(addr), "r"((uint32_t)pred) : "memory");
}

struct SmemCtl {
    uint64_t a_full[kAStages], a_empty[kAStages];
    uint64_t t_full[kMaxAcc], t_empty[kMaxAcc];
    uint64_t w_full[kMaxWStages], w_empty[kMaxWStages];
    uint64_t w_res;
    uint32_t tmem_base;
    uint32_t pad;
};
constexpr int kCtlBytes = 512;
constexpr int kXchFloats = kMaxGroups /*group*/ * 2 /*parity*/ * 2 /*up,down*/ * 4 /*quarter*/ * 64;
constexpr int kXchBytes = kXchFloats * 4 + 1024;  // 16 KB exchange + 128 floats of bias + 128 floats of folded-PreNorm weight sums
static_assert(sizeof(SmemCtl) <= kCtlBytes, "ctl block too small");

// A sub-tile's epilogue is one long dependent instruction stream per warp (TMEM loads, exchange, 2 shuffles + 5 flops per
// output element, stores): ~4300 clk against ~1000 clk of MMAs with one group of 4 warps.  Round 1 split the channels of a block
// over two sets of 4 warps ("halves", CD_EPI_SPLIT=1: both sets read the same TMEM lane quarters at different columns); round 2
// measured the kernel bound by the epilogue's instruction COUNT, and the split doubles the per-row work (row decode, masks, barrier
// waits): three unsplit groups are faster (DS2 32->32 L0: 0.442 vs 0.468 ms) and are the default.
template <int CB, bool K3> __host__ __device__ constexpr int kHalves() { return (CD_EPI_SPLIT && CB % 32 == 0) ? 2 : 1; }

template <int CB, bool K3, int CPG>   // CPG: channels per GroupNorm group for the statistics (divides CB)
__global__ void __launch_bounds__((kThreads<CB, K3>()), 1) conv_mma_kernel(const MmaParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    SmemCtl* ctl = reinterpret_cast<SmemCtl*>(smem);
    float* xch = reinterpret_cast<float*>(smem + kCtlBytes);
    unsigned char* wbuf = smem + kCtlBytes + kXchBytes;
    const int w_smem_bytes = p.resident ? p.w_bytes_cb : p.nwstages * p.wstage_bytes;
    unsigned char* abuf = wbuf + ((w_smem_bytes + 127) / 128) * 128;
    const uint32_t wbuf_s = smem_u32(wbuf), abuf_s = smem_u32(abuf);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // broadcast: provably warp-uniform (role branches stay convergent)
    const int lane = threadIdx.x & 31;
    const Geom& g = p.g;
    constexpr int N = K3 ? 3 * CB : CB;
    constexpr int G = kGroups<CB, K3>();
    DBG_DECL();

    if (threadIdx.x == 0) {
        for (int s = 0; s < kAStages; ++s) { mbar_init(smem_u32(&ctl->a_full[s]), 1); mbar_init(smem_u32(&ctl->a_empty[s]), G); }
        for (int s = 0; s < kMaxAcc; ++s) { mbar_init(smem_u32(&ctl->t_full[s]), 1); mbar_init(smem_u32(&ctl->t_empty[s]), 4 * kHalves<CB, K3>()); }
        for (int s = 0; s < kMaxWStages; ++s) { mbar_init(smem_u32(&ctl->w_full[s]), 1); mbar_init(smem_u32(&ctl->w_empty[s]), G); }
        mbar_init(smem_u32(&ctl->w_res), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&ctl->tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = ctl->tmem_base;
    // accumulator slots of class c (sub-tiles n = c, c + G, ...; h = n / G counts the class's sub-tiles): the nacc slots are dealt
    // out as evenly as they go (N = 96, G = 3: 2 + 2 + 1); use h of a class is slot base + h % cnt with phase (h / cnt) & 1.
    // (h / cnt by multiply-high: exact for h < 2^29, cnt <= 8.)
    struct SlotRing { uint32_t base, cnt, magic; };
    auto ring_of = [&](uint32_t c) {
        const uint32_t lo = (uint32_t)p.nacc / (uint32_t)G, extra = (uint32_t)p.nacc - lo * (uint32_t)G;
        SlotRing r;
        r.cnt = lo + (c < extra ? 1u : 0u);
        r.base = c * lo + (c < extra ? c : extra);
        r.magic = 0xffffffffu / r.cnt + 1u;
        return r;
    };
    auto lap_of = [](const SlotRing& r, uint32_t h) { return r.cnt == 1u ? h : __umulhi(h, r.magic); };   // (the magic number wraps to 0 for cnt = 1)
    auto slot_of = [&](const SlotRing& r, uint32_t h) { return r.base + (h - lap_of(r, h) * r.cnt); };
    auto phase_of = [&](const SlotRing& r, uint32_t h) { return lap_of(r, h) & 1u; };

    const int cb = blockIdx.x % p.ncb;
    const int cta = blockIdx.x / p.ncb;
    const int tile_begin = cta * p.tiles_per_cta;
    int tile_end = tile_begin + p.tiles_per_cta;
    if (tile_end > p.ntiles) tile_end = p.ntiles;
    const int my_tiles = tile_end > tile_begin ? tile_end - tile_begin : 0;
    const uint32_t win_bytes = (uint32_t)p.wrows * 16u;
    const uint32_t chunk_bytes = win_bytes * p.nwin;   // smem bytes per 8-channel chunk
    const unsigned char* wsrc = reinterpret_cast<const unsigned char*>(p.w) + (size_t)cb * p.w_bytes_cb;
    const int nkg = (p.nchunks + p.kgroup_chunks - 1) / p.kgroup_chunks;

    if (warp == 0) {
        // ===================== producer: resident weights / weight ring + activation windows =====================
        if (lane == 0) {
            if (p.resident) {
                mbar_expect_tx(smem_u32(&ctl->w_res), p.w_bytes_cb);
                for (int off = 0; off < p.w_bytes_cb; off += 32768) {
                    int n = p.w_bytes_cb - off < 32768 ? p.w_bytes_cb - off : 32768;
                    bulk_g2s(wbuf_s + off, wsrc + off, n, smem_u32(&ctl->w_res));
                }
            }
            uint32_t ws = 0, wph = 0;                 // weight-ring stage / phase (counters instead of % and /)
            int it = 0, kgi = 0, stage = 0;
            uint32_t ph = 0;
            for (int item = 0; item < my_tiles * p.ksplit; ++item,
                     kgi = (kgi + 1 == p.ksplit) ? 0 : kgi + 1, it += (kgi == 0), stage = (stage + 1 == p.astages) ? 0 : stage + 1,
                     ph ^= (uint32_t)(stage == 0)) {
                const int tile = tile_begin + it;
                const long long d0 = (long long)tile * p.tile_rows + p.d_lead;
                mbar_wait_relaxed(smem_u32(&ctl->a_empty[stage]), ph ^ 1);
                const bool stale = DBG_MODE(2) && item >= p.astages;      // debug: reuse the windows already in shared memory
                if (stale) mbar_arrive(smem_u32(&ctl->a_full[stage]));
                else mbar_expect_tx(smem_u32(&ctl->a_full[stage]), p.a_stage_bytes);
                for (int ch = 0; ch < (stale ? 0 : p.gchunks); ++ch) {
                    const int c = kgi * p.gchunks + ch;                     // chunk of the (concatenated) input
                    const int sidx = c < p.src_chunks[0] ? 0 : 1;
                    const int k = sidx ? c - p.src_chunks[0] : c;
                    const __half* plane0 = p.src[sidx] + plcp_offset(g, k, 0);
                    for (int w = 0; w < p.nwin; ++w) {
                        const long long r0 = d0 + p.row0 + (long long)(w - (p.nwin >> 1)) * p.win_dz;
                        bulk_g2s(abuf_s + stage * p.a_stage_bytes + ch * chunk_bytes + w * win_bytes, plane0 + r0 * 8, win_bytes,
                                 smem_u32(&ctl->a_full[stage]));
                    }
                }
                if (!p.resident) {
                    for (int s = 0; s < p.nshift; ++s) {
                        for (int kg = 0; kg < nkg; ++kg, ws = (ws + 1 == (uint32_t)p.nwstages) ? 0u : ws + 1, wph ^= (uint32_t)(ws == 0)) {
                            const int kc0 = kg * p.kgroup_chunks;
                            const int kcn = (p.nchunks - kc0) < p.kgroup_chunks ? (p.nchunks - kc0) : p.kgroup_chunks;
                            const uint32_t bytes = (uint32_t)kcn * N * 16u;
                            mbar_wait(smem_u32(&ctl->w_empty[ws]), wph ^ 1);
                            mbar_expect_tx(smem_u32(&ctl->w_full[ws]), bytes);
                            bulk_g2s(wbuf_s + ws * p.wstage_bytes, wsrc + ((size_t)s * p.nchunks + kc0) * N * 16u, bytes,
                                     smem_u32(&ctl->w_full[ws]));
                        }
                    }
                }
            }
        }
    } else if (warp < kFirstEpiWarp<CB, K3>()) {
        // ===================== MMA issuers (one thread of each of two warps) =====================
        // A single thread cannot issue M128xN96xK16 MMAs with changing descriptors faster than one per ~105 clk
        // (measured, scripts/mma_microbench.cu: descriptor arithmetic -> R2UR -> UTCHMMA is a serial chain), while the
        // tensor pipe executes one in 56 clk.  Two issuing warps, each owning every other sub-tile (its own TMEM
        // accumulator slot), reach the 56 clk execution bound.  All offsets are precomputed in 16-byte descriptor units.
        // The issue loop runs warp-convergent (all 32 lanes walk it, only the elected lane executes the tcgen05 / arrive
        // instructions): descriptor arithmetic then stays in the uniform datapath, and the per-MMA descriptor low words are
        // taken from tables built once per launch, so an MMA costs a handful of instructions instead of a ~20-instruction
        // dependent chain (measured: 164 clk per MMA per issuing thread before, against a 56 clk tensor-pipe slot).
        {
            const uint32_t iw = (uint32_t)warp - 1u;
            const bool leader = lane == 0;
            constexpr int NS = K3 ? 9 : 1;
            uint32_t soff[NS];
#pragma unroll
            for (int s = 0; s < NS; ++s) {
                const int a = K3 ? s / 3 : 1, b = K3 ? s % 3 : 1;
                const int win = p.nwin == 3 ? a : 0;
                const int rowoff = -p.row0 + (a - 1) * p.off_dz + (b - 1) * p.off_dp;
                soff[s] = (uint32_t)(win * (int)win_bytes + rowoff * 16) >> 4;
            }
            const uint64_t adesc_hi = make_desc(0, chunk_bytes, 128u);
            const uint64_t bdesc_hi = make_desc(0, (uint32_t)N * 16u, 128u);
            const uint32_t qstep = (2u * chunk_bytes) >> 4, bstep = (2u * N * 16u) >> 4;
            const uint32_t jstep = ((uint32_t)p.sub_stride * 16u) >> 4;
            const int nq = p.nchunks >> 1;
            // sub-tiles are numbered n = cnt + j; (tc, th) = (n % G, n / G) of the current tile's first sub-tile, kept as counters
            uint32_t tc = 0, th = 0;
            const SlotRing ring = ring_of(iw);
            auto next_tile = [&]() { tc += (uint32_t)p.MT; while (tc >= (uint32_t)G) { tc -= (uint32_t)G; ++th; } };
            // is sub-tile j of the current tile mine, and which use of my class is it
            auto mine = [&](int j, uint32_t& h) { uint32_t c = tc + (uint32_t)j; h = th; while (c >= (uint32_t)G) { c -= (uint32_t)G; ++h; } return c == iw; };
            uint32_t ws = 0, wph = 0;                 // weight-ring stage / phase
            int astage = 0;                           // activation-ring stage / phase of the next item
            uint32_t aph = 0;
            auto next_astage = [&]() { astage = (astage + 1 == p.astages) ? 0 : astage + 1; aph ^= (uint32_t)(astage == 0); };
            if (p.resident) mbar_wait(smem_u32(&ctl->w_res), 0);
            const int gq = p.gchunks >> 1;                       // K16 steps per staged K group
            const uint32_t bshift = (uint32_t)p.nchunks * N;     // weight offset (16-byte units) between (dz,dp) shifts
            // descriptor low words of the NS x GQ MMAs of one sub-tile and K group (GQ = 2: every 32-wide K group)
            constexpr int GQF = 2;
            const uint32_t ahi32 = (uint32_t)(adesc_hi >> 32), bhi32 = (uint32_t)(bdesc_hi >> 32);
            uint32_t aoff[NS * GQF], boff[NS * GQF];
#pragma unroll
            for (int s = 0; s < NS; ++s)
#pragma unroll
                for (int q = 0; q < GQF; ++q) {
                    aoff[s * GQF + q] = (uint32_t)adesc_hi + soff[s] + (uint32_t)q * qstep;
                    boff[s * GQF + q] = (uint32_t)bdesc_hi + (wbuf_s >> 4) + (uint32_t)s * bshift + (uint32_t)q * bstep;
                }
            uint32_t boff1[NS];      // gq == 1: the shift stride of the weights is one K16 step
#pragma unroll
            for (int s = 0; s < NS; ++s) boff1[s] = (uint32_t)bdesc_hi + (wbuf_s >> 4) + (uint32_t)s * bshift;
            for (int it = 0; it < my_tiles; ++it) {
                if (p.resident) {
                    // a tile's K range arrives in ksplit stages; every sub-tile accumulator stays open until the last one
                    for (int kgi = 0; kgi < p.ksplit; ++kgi, next_astage()) {
                        const int stage = astage;
                        const uint32_t ph = aph;
                        const uint32_t a_lo0 = (abuf_s + stage * p.a_stage_bytes) >> 4;
                        const uint32_t b_kg = (uint32_t)(kgi * p.gchunks) * N;
                        const uint32_t b_lo0 = (wbuf_s >> 4) + b_kg;
                        const long long tA = DBG_T();
                        mbar_wait(smem_u32(&ctl->a_full[stage]), ph);
                        if (blockIdx.x == 0 && iw == 0 && leader) { DBG_ADD(0, DBG_T() - tA); DBG_ADD(3, 1); }
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        bool issued = false;
                        for (int j = 0; j < p.MT; ++j) {
                            uint32_t h;
                            if (!mine(j, h)) continue;
                            const uint32_t a_lo = a_lo0 + (uint32_t)j * jstep;
                            const uint32_t slot = slot_of(ring, h), tph = phase_of(ring, h);
                            const uint32_t dcol = tmem_base + slot * N;
                            const long long tB = DBG_T();
                            if (kgi == 0) {
                                mbar_wait(smem_u32(&ctl->t_empty[slot]), tph ^ 1);
                                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                            }
                            const long long tC = DBG_T();
                            if (gq == GQF) {
                                if (!DBG_MODE(128)) {
#pragma unroll
                                    for (int i = 0; i < NS * GQF; ++i)
                                        umma_f16_elect(dcol, ((uint64_t)ahi32 << 32) | (aoff[i] + a_lo),
                                                       ((uint64_t)bhi32 << 32) | (boff[i] + b_kg), p.idesc, (kgi | i) != 0);
                                }
                            } else if (gq == 1) {        // 16-channel K groups (the init conv): every other table entry
#pragma unroll
                                for (int s = 0; s < NS; ++s)
                                    umma_f16_elect(dcol, ((uint64_t)ahi32 << 32) | (aoff[s * GQF] + a_lo),
                                                   ((uint64_t)bhi32 << 32) | (boff1[s] + b_kg), p.idesc, (kgi | s) != 0);
                            } else {
#pragma unroll
                                for (int s = 0; s < NS; ++s) {
                                    uint32_t al = a_lo + soff[s], b_lo = b_lo0 + (uint32_t)s * bshift;
                                    for (int q = 0; q < gq; ++q, al += qstep, b_lo += bstep)
                                        if (!DBG_MODE(128)) umma_f16_elect(dcol, adesc_hi | al, bdesc_hi | b_lo, p.idesc, (kgi | s | q) != 0);
                                }
                            }
                            if (kgi == p.ksplit - 1) umma_commit_elect(smem_u32(&ctl->t_full[slot]));
                            issued = true;
                            if (blockIdx.x == 0 && iw == 0 && leader) { DBG_ADD(1, tC - tB); DBG_ADD(2, DBG_T() - tC); }
                        }
                        if (issued) umma_commit_elect(smem_u32(&ctl->a_empty[stage]));
                        else mbar_arrive_elect(smem_u32(&ctl->a_empty[stage]));
                    }
                    next_tile();
                    continue;
                }
                if (!leader) { next_tile(); continue; }      // streamed-weight form: the elected lane walks alone
                const int stage = astage;
                const uint32_t ph = aph;
                next_astage();
                const uint32_t a_lo0 = (abuf_s + stage * p.a_stage_bytes) >> 4;
                mbar_wait(smem_u32(&ctl->a_full[stage]), ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                bool issued = false;
                {
                    for (int j = 0; j < p.MT; ++j) {
                        uint32_t h;
                        if (mine(j, h)) mbar_wait(smem_u32(&ctl->t_empty[slot_of(ring, h)]), phase_of(ring, h) ^ 1);
                    }
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                    for (int s = 0; s < NS; ++s) {
                        for (int kg = 0; kg < nkg; ++kg, ws = (ws + 1 == (uint32_t)p.nwstages) ? 0u : ws + 1, wph ^= (uint32_t)(ws == 0)) {
                            const int kc0 = kg * p.kgroup_chunks;
                            const int kcn = (p.nchunks - kc0) < p.kgroup_chunks ? (p.nchunks - kc0) : p.kgroup_chunks;
                            mbar_wait(smem_u32(&ctl->w_full[ws]), wph);
                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                            const uint32_t b_lo0 = (wbuf_s + ws * p.wstage_bytes) >> 4;
                            const uint32_t a_lo1 = a_lo0 + soff[s] + (uint32_t)kc0 * (chunk_bytes >> 4);
                            bool used = false;
                            for (int j = 0; j < p.MT; ++j) {
                                uint32_t h;
                                if (!mine(j, h)) continue;
                                const uint32_t dcol = tmem_base + slot_of(ring, h) * N;
                                uint32_t al = a_lo1 + (uint32_t)j * jstep, b_lo = b_lo0;
                                for (int q = 0; q < (kcn >> 1); ++q, al += qstep, b_lo += bstep)
                                    umma_f16(dcol, adesc_hi | al, bdesc_hi | b_lo, p.idesc, (s | kg | q) != 0);
                                used = true;
                            }
                            if (used) umma_commit(smem_u32(&ctl->w_empty[ws]));
                            else mbar_arrive(smem_u32(&ctl->w_empty[ws]));
                            issued = issued || used;
                        }
                    }
                    for (int j = 0; j < p.MT; ++j) {
                        uint32_t h;
                        if (mine(j, h)) umma_commit(smem_u32(&ctl->t_full[slot_of(ring, h)]));
                    }
                    next_tile();
                }
                if (issued) umma_commit(smem_u32(&ctl->a_empty[stage]));
                else mbar_arrive(smem_u32(&ctl->a_empty[stage]));
            }
        }
    } else {
        // ===================== epilogue: TMEM -> registers -> global (+ GroupNorm statistics) =====================
        const int q = warp & 3;                   // TMEM lane quarter this warp may read (warp id % 4)
        constexpr int EW = kEpiWarps<CB, K3>();
        const int eset = (warp - kFirstEpiWarp<CB, K3>()) >> 2;    // sets of 4 warps; each set covers the four lane quarters
        const int group = eset % G;               // the groups take sub-tiles round-robin
        const int half = eset / G;                // which half of the block's channels this warp handles (0 when not split)
        constexpr int HALVES = kHalves<CB, K3>();
        constexpr int CH = CB / HALVES;           // channels per half
        const int i_row = q * 32 + lane;          // row within the 128-row sub-tile
        float* sbias = xch + kXchFloats;          // bias of this channel block, broadcast-read from shared memory
        float* swsum = sbias + 128;              // folded PreNorm: sum_c W'[o][c]; sbias then holds sum_c W[o][c]*beta[c]
        for (int c = threadIdx.x - 32 * kFirstEpiWarp<CB, K3>(); c < CB; c += 32 * EW) {
            sbias[c] = p.pn_stats ? __ldg(p.pn_wb + cb * CB + c) : (p.bias ? __ldg(p.bias + cb * CB + c) : 0.f);
            swsum[c] = p.pn_stats ? __ldg(p.pn_wsum + cb * CB + c) : 0.f;
        }
        asm volatile("bar.sync 15, %0;" ::"n"(32 * EW) : "memory");
        constexpr int NG = K3 ? CH / CPG : 1;     // statistics groups per half of the channel block
        static_assert(!K3 || CH % CPG == 0, "a GroupNorm group must not straddle the channel halves");
        float gs[NG], gq[NG];                     // this thread's running sum / sum of squares per group (sample cur_b)
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
                        double* d = p.stats + ((long long)cur_b * p.stats_C + cb * CB + half * CH + i * CPG) * 2;
                        atomicAdd(d, (double)s_tot);
                        atomicAdd(d + 1, (double)q_tot);
                    }
                    gs[i] = 0.f; gq[i] = 0.f;
                }
            }
        };
        const int halo_delta = g.P * g.Rp;
        const int total_sub = half < HALVES ? my_tiles * p.MT : 0;     // CB not a multiple of 32: the second half idles
        // Row of this thread in its group's sub-tile n: (tile_begin * MT + n) * sub_stride + d_lead + i_row, i.e. a constant stride of
        // G * sub_stride rows per iteration.  Its decomposition (sample, z, pp, rr) is kept incrementally (add the decomposed
        // stride, carry once per digit) instead of three divisions per sub-tile, and the store offset advances with it.
        const long long cs8 = g.chunk_stride * 8;                        // halves between consecutive chunk planes
        const long long obase = ((long long)(cb * (CB / 8)) * g.chunk_stride + g.guard) * 8;   // this channel block's first plane, row 0
        int e_stride = G * p.sub_stride;
        // loop invariants pinned in registers (ptxas otherwise re-loads them from the parameter bank every sub-tile and the
        // epilogue stalls on the constant-cache latency: 6 % of the kernel's stall samples)
        int gRp = g.Rp, gPp = g.Pp, gZ = g.Z, gR = g.R, gP = g.P, nrows_i = (int)p.nrows;
        asm volatile("" : "+r"(e_stride), "+r"(gRp), "+r"(gPp), "+r"(gZ), "+r"(gR), "+r"(gP), "+r"(nrows_i));
        int s_b, s_z, s_pp, s_rr;                                        // the stride, digit by digit
        {
            int t = e_stride;
            s_b = t / g.rows_per_sample; t -= s_b * g.rows_per_sample;
            s_z = t / g.plane; t -= s_z * g.plane;
            s_pp = t / g.Rp; s_rr = t - s_pp * g.Rp;
        }
        int row = (tile_begin * p.MT + group) * p.sub_stride + p.d_lead + i_row - e_stride;   // advanced at the top of the loop
        int c_b, c_z, c_pp, c_rr;                                        // digits of `row` (floor semantics for the rows before row 0)
        {
            int t = row;
            c_b = t >= 0 ? t / g.rows_per_sample : -((-t + g.rows_per_sample - 1) / g.rows_per_sample);
            t -= c_b * g.rows_per_sample;
            c_z = t / g.plane; t -= c_z * g.plane;
            c_pp = t / g.Rp; c_rr = t - c_pp * g.Rp;
        }
        int par = 1;                                                     // exchange-buffer parity of this group's current sub-tile
        const SlotRing ring = ring_of((uint32_t)group);
        uint32_t e_s = ring.cnt - 1u, e_ph = 1u;                         // slot index within the group's ring / phase (counters)
        for (int n = group; n < total_sub; n += G) {
            par ^= 1;
            if (++e_s == ring.cnt) { e_s = 0u; e_ph ^= 1u; }
            const uint32_t slot = ring.base + e_s, tph = e_ph;
            row += e_stride;
            c_rr += s_rr; c_pp += s_pp; c_z += s_z; c_b += s_b;
            if (c_rr >= gRp) { c_rr -= gRp; ++c_pp; }
            if (c_pp >= gPp) { c_pp -= gPp; ++c_z; }
            if (c_z > gZ) { c_z -= gZ + 1; ++c_b; }
            // which output row is this?
            // 1x1x1 stores cover every row of the sub-tile (whole 32-byte sectors, no read-modify-write in L2/HBM): interior rows
            // get the result, zero rows get zeros (the invariant every consumer relies on), phi halo rows get a don't-care value
            // (re-derived from the interior by whoever needs them) unless this launch maintains them (write_halo).
            const bool inrange = row >= 0 && row < nrows_i && (!K3 || (i_row >= 1 && i_row <= 126));
            const bool nonzero_row = c_z < gZ && c_rr < gR;
            const bool valid = inrange && nonzero_row && c_pp >= 1 && c_pp <= gP;
            const bool phi_halo = inrange && nonzero_row && !valid;
            const int sample = inrange ? c_b : -1;
            const uint32_t pp = (uint32_t)c_pp;
            // 1x1x1 (HBM-bound): whole-sector stores of every row.  3x3x3 (epilogue-bound): interior rows only, measured faster.
            const bool do_store = K3 ? valid : (inrange && !(p.write_halo && phi_halo));
            const bool keep = valid || (!K3 && phi_halo && !p.write_halo);
            const bool halo_up = valid && p.write_halo && pp == 1u;               // copy to pp = P+1
            const bool halo_dn = valid && p.write_halo && pp == (uint32_t)gP;     // copy to pp = 0
            const long long tE0 = DBG_T();
            mbar_wait_relaxed(smem_u32(&ctl->t_full[slot]), tph);
            const long long tE1 = DBG_T();
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#ifdef CD_MMA_DEBUG
            if (p.dbg_mode & 1) {
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(&ctl->t_empty[slot]));
                continue;
            }
#endif
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + slot * N;
            if constexpr (K3) {
                constexpr int CW = (G >= 3) ? CD_EPI_CW3 : 16;   // channels recombined per pass
                static_assert(CH % CW == 0, "channel half must be a multiple of the pass width");
                float* xup = xch + (size_t)(((group * 2 + par) * 2 + 0) * 4) * 64;
                float* xdn = xch + (size_t)(((group * 2 + par) * 2 + 1) * 4) * 64;
                // sample bookkeeping for the statistics (warp-uniform decisions)
                bool uniform = true;
                if (p.stats) {
                    const unsigned vmask = __ballot_sync(0xffffffffu, valid);
                    if (vmask) {
                        const int b_first = __shfl_sync(0xffffffffu, sample, __ffs(vmask) - 1);
                        uniform = __all_sync(0xffffffffu, !valid || sample == b_first);
                        if (uniform) { if (b_first != cur_b) { flush(); cur_b = b_first; } }
                        else { flush(); cur_b = -1; }
                    }
                }
#pragma unroll
                for (int cr = 0; cr < CH; cr += CW) {
                    constexpr int W = CW;
                    constexpr int w = W;
                    const int c0 = half * CH + cr;                 // first channel of this pass within the block
                    uint32_t r0[W], r1[W], r2[W];
#pragma unroll
                    for (int k = 0; k < W; ++k) { r0[k] = 0; r1[k] = 0; r2[k] = 0; }
                    if (!DBG_MODE(32)) {
                        tmem_ldw_issue<W>(taddr + CB + c0, r1);       // D_1 of row m
                        tmem_ldw_issue<W>(taddr + c0, r0);            // D_0 of row m   (wanted by lane + 1)
                        tmem_ldw_issue<W>(taddr + 2 * CB + c0, r2);   // D_2 of row m   (wanted by lane - 1)
                    }
                    tmem_ld_wait();
                    if (blockIdx.x == 0 && warp == kFirstEpiWarp<CB, K3>() && lane == 0) DBG_ADD(9, DBG_T() - tE1);
                    if (cr + CW >= CH) {
                        // this warp's last TMEM read of the sub-tile: the slot is free once all 4*HALVES warps have arrived
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) mbar_arrive(smem_u32(&ctl->t_empty[slot]));
                    }
                    // Neighbour rows: lanes 1..31 take D_0 of row m-1 from lane-1, lanes 0..30 take D_2 of row m+1 from lane+1
                    // (rotate shuffles).  The two rows that live in the neighbouring lane quarters (another warp) go through shared
                    // memory: published right after the TMEM loads, read back by lane 0 / lane 31 only AFTER the 32 shuffles, so
                    // the store -> barrier -> load round trip (slow while the MMAs saturate the shared-memory pipe: it cost 0.19 ms
                    // of 0.68 when it sat in front of the shuffles) overlaps with them instead of stalling the warp.
                    // (A split-phase mbarrier with lanes 0 / 31 finished in a second pass measured slower: 0.64 vs 0.50 ms.)
                    const bool xchg = !DBG_MODE(4) && !DBG_MODE(1024);
                    if (xchg) {
                        const uint32_t up_s = smem_u32(xup + q * 64 + c0), dn_s = smem_u32(xdn + q * 64 + c0);
#pragma unroll
                        for (int k = 0; k < W; k += 4) sts128_if(lane == 31, up_s + k * 4, r0[k], r0[k + 1], r0[k + 2], r0[k + 3]);
#pragma unroll
                        for (int k = 0; k < W; k += 4) sts128_if(lane == 0, dn_s + k * 4, r2[k], r2[k + 1], r2[k + 2], r2[k + 3]);
                    }
                    const int src_up = (lane + 31) & 31, src_dn = (lane + 1) & 31;
                    float acc[W];
#pragma unroll
                    for (int k = 0; k < W; k += 4) {
                        const float4 bv = *reinterpret_cast<const float4*>(sbias + c0 + k);
                        const float bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            // lane 0's `u` and lane 31's `d` come from the wrong row here; they are replaced below
                            const uint32_t u = DBG_MODE(4) ? r0[k + e] : __shfl_sync(0xffffffffu, r0[k + e], src_up);
                            const uint32_t d = DBG_MODE(4) ? r2[k + e] : __shfl_sync(0xffffffffu, r2[k + e], src_dn);
                            acc[k + e] = __uint_as_float(r1[k + e]) + bb[e];
                            r0[k + e] = u; r2[k + e] = d;
                        }
                    }
                    if (xchg) {
                        if (!DBG_MODE(512)) asm volatile("bar.sync %0, 128;" ::"r"(1 + eset) : "memory");
                        if (blockIdx.x == 0 && warp == kFirstEpiWarp<CB, K3>() && lane == 0) DBG_ADD(10, DBG_T() - tE1);
                        // lane 0: D_0 of the previous quarter's last row; lane 31: D_2 of the next quarter's first row
                        // (q = 0 / q = 3 keep a don't-care value: rows 0 and 127 of a sub-tile are never stored)
                        const uint32_t up_s = smem_u32(xup + ((q + 3) & 3) * 64 + c0), dn_s = smem_u32(xdn + ((q + 1) & 3) * 64 + c0);
#pragma unroll
                        for (int k = 0; k < W; k += 4) lds128_if(lane == 0, up_s + k * 4, r0[k], r0[k + 1], r0[k + 2], r0[k + 3]);
#pragma unroll
                        for (int k = 0; k < W; k += 4) lds128_if(lane == 31, dn_s + k * 4, r2[k], r2[k + 1], r2[k + 2], r2[k + 3]);
                    }
#pragma unroll
                    for (int k = 0; k < W; ++k) acc[k] += __uint_as_float(r0[k]) + __uint_as_float(r2[k]);
                    if (p.addsrc && valid) {
#pragma unroll
                        for (int k = 0; k < W; k += 8) {
                            float f[8];
                            unpack8(load8(p.addsrc + obase + (long long)row * 8 + ((c0 + k) >> 3) * cs8), f);
#pragma unroll
                            for (int e = 0; e < 8; ++e) acc[k + e] += f[e];
                        }
                    }
                    if (blockIdx.x == 0 && warp == kFirstEpiWarp<CB, K3>() && lane == 0) DBG_ADD(11, DBG_T() - tE1);
#pragma unroll
                    for (int k = 0; k < W; k += 8) {
                        Half8 h = pack8(acc + k);
                        if (!keep) h = Half8{};
                        __half* o = p.dst + obase + (long long)row * 8 + ((c0 + k) >> 3) * cs8;
                        if (do_store && !DBG_MODE(8)) store8(o, h);
                        if (p.write_halo) {      // uniform branch: the resblock convs (no halo copies) skip the address arithmetic too
                            if (halo_up) store8(o + (long long)halo_delta * 8, h);
                            if (halo_dn) store8(o - (long long)halo_delta * 8, h);
                        }
                    }
                    if (blockIdx.x == 0 && warp == kFirstEpiWarp<CB, K3>() && lane == 0) DBG_ADD(12, DBG_T() - tE1);
                    if (p.stats && valid && !DBG_MODE(16)) {
                        if (uniform) {
#pragma unroll
                            for (int k = 0; k < W; ++k) { gs[(cr + k) / CPG] += acc[k]; gq[(cr + k) / CPG] = fmaf(acc[k], acc[k], gq[(cr + k) / CPG]); }
                        } else {
                            // the warp's 32 rows straddle two samples (once per sample boundary): direct atomics
#pragma unroll
                            for (int k = 0; k < W; ++k) {
                                double* d = p.stats + ((long long)sample * p.stats_C + cb * CB + ((c0 + k) / CPG) * CPG) * 2;
                                atomicAdd(d, (double)acc[k]);
                                atomicAdd(d + 1, (double)(acc[k] * acc[k]));
                            }
                        }
                    }
                }
            } else {
                // 1x1x1: no recombination, no statistics: batches of up to 48 columns with all TMEM loads in flight at once
                constexpr int BW = (CH % 48 == 0) ? 48 : ((CH % 32 == 0) ? 32 : 16);
                // folded PreNorm: y = rstd*acc + (wb - mean*rstd*wsum), statistics of this row's sample
                float pn_a = 1.f, pn_m = 0.f;
                if (p.pn_stats && inrange) {
                    const double* st = p.pn_stats + (long long)sample * p.pn_C * 2;
                    const double mean = st[0] * (double)p.pn_inv_count;
                    double var = st[1] * (double)p.pn_inv_count - mean * mean;
                    if (var < 0.0) var = 0.0;
                    pn_a = rsqrtf((float)var + 1e-5f);
                    pn_m = -(float)mean * pn_a;
                }
#pragma unroll
                for (int cr = 0; cr < CH; cr += BW) {
                    const int c0 = half * CH + cr;
                    uint32_t r1[BW];
#pragma unroll
                    for (int k = 0; k < BW; k += 16) tmem_ld16_issue(taddr + c0 + k, r1 + k);
                    tmem_ld_wait();
                    if (cr + BW >= CH) {
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) mbar_arrive(smem_u32(&ctl->t_empty[slot]));
                    }
#pragma unroll
                    for (int k = 0; k < BW; k += 8) {
                        const float4 b0 = *reinterpret_cast<const float4*>(sbias + c0 + k);
                        const float4 b1 = *reinterpret_cast<const float4*>(sbias + c0 + k + 4);
                        float acc[8];
                        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                        if (p.pn_stats) {
                            const float4 w0 = *reinterpret_cast<const float4*>(swsum + c0 + k);
                            const float4 w1 = *reinterpret_cast<const float4*>(swsum + c0 + k + 4);
                            const float ws[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                            for (int e = 0; e < 8; ++e) acc[e] = fmaf(pn_a, __uint_as_float(r1[k + e]), fmaf(pn_m, ws[e], bb[e]));
                        } else {
#pragma unroll
                            for (int e = 0; e < 8; ++e) acc[e] = __uint_as_float(r1[k + e]) + bb[e];
                        }
                        Half8 h = pack8(acc);
                        if (!keep) h = Half8{};
                        __half* o = p.dst + obase + (long long)row * 8 + ((c0 + k) >> 3) * cs8;
                        if (do_store) store8(o, h);
                        if (p.write_halo) {
                            if (halo_up) store8(o + (long long)halo_delta * 8, h);
                            if (halo_dn) store8(o - (long long)halo_delta * 8, h);
                        }
                    }
                }
            }
            if (blockIdx.x == 0 && warp == kFirstEpiWarp<CB, K3>() && lane == 0) { DBG_ADD(4, tE1 - tE0); DBG_ADD(5, DBG_T() - tE1); DBG_ADD(6, 1); }
        }
        if (K3 && p.stats) flush();
    }
    DBG_FLUSH(blockIdx.x == 0 && lane == 0 && (warp == 1 || warp == kFirstEpiWarp<CB, K3>()));
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
// host side: weight packing, tile planning, launch
// ---------------------------------------------------------------------------------------------------------
static int pick_cb(int Cout, int taps) {
    if (taps == 1) return (Cout % 16 == 0 && Cout <= 128) ? Cout : 0;   // 1x1x1: one pass over the input, N = C_out
    const int cands[4] = {64, 48, 32, 16};
    for (int c : cands) if (Cout % c == 0) return c;
    return 0;
}

int mma_pick_cb(int Cout, int taps) { return pick_cb(Cout, taps); }

bool mma_conv_supported(int Cin, int Cout, int taps) {
    return (taps == 27 || taps == 1) && Cin % 16 == 0 && Cin >= 16 && Cout % 16 == 0 && pick_cb(Cout, taps) > 0;
}

size_t mma_packed_elems(int Cin, int Cout, int taps) { return (size_t)taps * Cin * Cout; }

// w: fp32 [tap][ci][co] (tap = (a*3+b)*3+c for 3x3x3).  out: fp16 [cb][shift][kc][n][8], n = g*CB + co_local
void mma_pack_weights(const float* w, int Cin, int Cout, int taps, __half* out) {
    const int CB = pick_cb(Cout, taps), ncb = Cout / CB, nkc = Cin / 8;
    const int nshift = taps == 27 ? 9 : 1, ng = taps == 27 ? 3 : 1, N = ng * CB;
    for (int cb = 0; cb < ncb; ++cb)
        for (int s = 0; s < nshift; ++s)
            for (int kc = 0; kc < nkc; ++kc)
                for (int n = 0; n < N; ++n)
                    for (int j = 0; j < 8; ++j) {
                        const int gi = n / CB, co = cb * CB + n % CB;
                        const int tap = s * ng + gi;
                        out[((((size_t)cb * nshift + s) * nkc + kc) * N + n) * 8 + j] =
                            __float2half_rn(w[((size_t)tap * Cin + kc * 8 + j) * Cout + co]);
                    }
}

template <int CB, bool K3, int CPG>
static cudaError_t launch_inst(const MmaParams& p, int grid, size_t smem, int max_smem, cudaStream_t st) {
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(conv_mma_kernel<CB, K3, CPG>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    conv_mma_kernel<CB, K3, CPG><<<grid, kThreads<CB, K3>(), smem, st>>>(p);
    return cudaSuccess;
}

int launch_conv_mma(const ConvParams& cp, const __half* packed_w, cudaStream_t st) {
    const int taps = cp.kz * cp.kp * cp.kr;
    if (cp.transposed || cp.sz != 1 || cp.sp != 1 || cp.sr != 1) return 1;
    const bool k3 = taps == 27;
    if (k3 && !(cp.kz == 3 && cp.kp == 3 && cp.kr == 3 && cp.pz == 1 && cp.cp == 1 && cp.pr == 1)) return 1;
    if (!mma_conv_supported(cp.Cin_total, cp.Cout, taps)) return 1;
    if (cp.gin.Z != cp.gout.Z || cp.gin.P != cp.gout.P || cp.gin.R != cp.gout.R) return 1;
    for (int s = 0; s < cp.nsrc; ++s) if (cp.src[s].C % 8 != 0) return 1;
    if (cp.stats && !k3) return 1;          // the 1x1x1 form carries no statistics epilogue
    if (cp.write_halo && cp.gout.P < 2) return 1;

    static int num_sms = 0, max_smem = 0;
    if (!num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    }
    const Geom& g = cp.gout;
    MmaParams p;
    p.g = g;
    p.src[0] = cp.src[0].base; p.src_chunks[0] = cp.src[0].C / 8;
    p.src[1] = cp.nsrc > 1 ? cp.src[1].base : nullptr; p.src_chunks[1] = cp.nsrc > 1 ? cp.src[1].C / 8 : 0;
    p.nchunks = cp.Cin_total / 8;
    if (p.src_chunks[0] + p.src_chunks[1] != p.nchunks) return 1;
    p.w = packed_w; p.bias = cp.bias; p.dst = cp.dst; p.stats = cp.stats;
    p.write_halo = cp.write_halo;
    if (cp.pn_stats && k3) return 1;        // the folded PreNorm exists in the 1x1x1 epilogue only
    if (cp.addsrc && !k3) return 1;         // accumulation onto a tensor exists in the 3x3x3 epilogue only
    p.addsrc = cp.addsrc;
    p.pn_stats = cp.pn_stats; p.pn_wsum = cp.pn_wsum; p.pn_wb = cp.pn_wb; p.pn_C = cp.pn_C;
    p.pn_inv_count = cp.pn_stats ? (float)(1.0 / cp.pn_count) : 0.f;
    p.Cout = cp.Cout; p.CB = pick_cb(cp.Cout, taps); p.ncb = cp.Cout / p.CB;
    p.stats_C = cp.stats_C > 0 ? cp.stats_C : cp.Cout;
    // statistics are accumulated per GroupNorm group: the group size must match an instantiated (CB, CPG) pair
    int cpg = 0;
    if (k3) {
        const int def_cpg[4] = {2, 4, 6, 8};                  // CB = 16, 32, 48, 64 with 8 groups over CB channels
        cpg = def_cpg[p.CB / 16 - 1];
        if (cp.stats) {
            cpg = cp.stats_cpg;
            const bool ok = (p.CB == 16 && cpg == 2) || (p.CB == 32 && (cpg == 4 || cpg == 8)) || (p.CB == 48 && (cpg == 6 || cpg == 12)) ||
                            (p.CB == 64 && (cpg == 8 || cpg == 16));
            if (!ok) return 1;
        }
    }
    p.N = k3 ? 3 * p.CB : p.CB;
    p.nshift = k3 ? 9 : 1;
    p.sub_stride = k3 ? 126 : 128;
    p.d_lead = k3 ? -1 : 0;
    p.nacc = 512 / p.N; if (p.nacc > kMaxAcc) p.nacc = kMaxAcc;
    // classes (issuer + epilogue group pairs) of this shape's kernel and the slots every class is sure to own
    const int groups = k3 ? (p.CB == 16 ? kGroups<16, true>() : p.CB == 32 ? kGroups<32, true>() : p.CB == 48 ? kGroups<48, true>() : kGroups<64, true>())
                          : kGroups<32, false>();
    const int class_slots = p.nacc / groups;
    if (class_slots < 1) return 1;
    const int mt_open = groups * class_slots < 4 ? groups * class_slots : 4;   // largest tile whose accumulators can all be open at once
    p.w_bytes_cb = p.nshift * cp.Cin_total * p.N * 2;
    p.kgroup_chunks = p.nchunks < 8 ? p.nchunks : 8;
    p.wstage_bytes = p.kgroup_chunks * p.N * 16;
    p.nwstages = 3;
    if (const char* e = getenv("CD_MMA_WSTAGES")) { const int v = atoi(e); if (v >= 2 && v <= kMaxWStages) p.nwstages = v; }   // diagnostics
    const int nwstages_plan = p.nwstages;
    p.nrows = (long long)cp.B * g.rows_per_sample;
    if (p.nrows + 4096 >= (1LL << 31)) return 1;

    const size_t fixed = kCtlBytes + kXchBytes + 256;
    bool found = false;
    size_t smem_total = 0;
    // window geometry of a tile of mt sub-tiles: one union window per chunk, or one window per dz when planes are large
    auto window = [&](int mt, int& nwin, int& wrows, int& row0, int& win_dz, int& off_dz, int& off_dp) {
        const int span = p.sub_stride * (mt - 1) + 128;
        nwin = 1; wrows = span; row0 = 0; win_dz = 0; off_dz = 0; off_dp = 0;
        if (k3) {
            const int uni = span + 2 * g.plane + 2 * g.Rp, three = 3 * (span + 2 * g.Rp);
            if (uni <= three) { nwin = 1; wrows = uni; row0 = -(g.plane + g.Rp); win_dz = 0; off_dz = g.plane; off_dp = g.Rp; }
            else { nwin = 3; wrows = span + 2 * g.Rp; row0 = -g.Rp; win_dz = g.plane; off_dz = 0; off_dp = g.Rp; }
        }
    };
    auto accept = [&](int resident, int ksplit, int mt, int astages) -> bool {
        int nwin, wrows, row0, win_dz, off_dz, off_dp;
        window(mt, nwin, wrows, row0, win_dz, off_dz, off_dp);
        const int gchunks = p.nchunks / ksplit;
        const size_t wbytes = resident ? (size_t)p.w_bytes_cb : (size_t)p.nwstages * p.wstage_bytes;
        const size_t a_stage = (size_t)gchunks * nwin * wrows * 16;
        const size_t total = fixed + (wbytes + 127) / 128 * 128 + astages * a_stage;
        if (total > (size_t)max_smem) return false;
        p.resident = resident; p.ksplit = ksplit; p.gchunks = gchunks; p.MT = mt; p.nwin = nwin; p.wrows = wrows; p.row0 = row0;
        p.win_dz = win_dz; p.off_dz = off_dz; p.off_dp = off_dp; p.a_stage_bytes = (int)a_stage; p.astages = astages;
        smem_total = total;
        return true;
    };
    // 1) weights resident, >= 2 activation stages; K staged in as few groups as fit.  Unsplit K: deepest activation ring first, then
    //    the largest tile (DS2 32 -> 32 at L0: 4 stages of 2 sub-tiles 0.423 ms, 3 x 3 0.436, 2 x 4 0.439 - a deeper ring hides the
    //    window loads better than a larger tile saves halo re-reads).  With K staged in groups all of a tile's accumulators stay
    //    open together (keep one slot free for the epilogue) and every stage is a fraction of a window: largest tile first.
    for (int ksplit = 1; ksplit <= 8 && !found; ksplit *= 2) {
        if (p.nchunks % ksplit != 0 || (p.nchunks / ksplit) % 2 != 0) continue;
        int mt_max = ksplit > 1 ? (p.nacc - 1 < mt_open ? p.nacc - 1 : mt_open) : 4;
        if (const char* e = getenv("CD_MMA_MT")) { const int v = atoi(e); if (v >= 1 && v < mt_max) mt_max = v; }   // diagnostics
        if (ksplit == 1 && !getenv("CD_MMA_TILE_FIRST")) {
            for (int astages = kAStages; astages >= 2 && !found; --astages)
                for (int mt = mt_max; mt >= (k3 ? 2 : 1) && !found; --mt) found = accept(1, ksplit, mt, astages);
        } else {
            for (int mt = mt_max; mt >= (k3 ? 2 : 1) && !found; --mt)
                for (int astages = kAStages; astages >= 2 && !found; --astages) found = accept(1, ksplit, mt, astages);
        }
    }
    // 2) otherwise: weights streamed through the 3-stage ring (K unsplit), two activation stages, largest tile that fits
    //    (a 2-stage ring would admit larger tiles for 128-wide inputs but faulted on multi-tile CTAs: not used)
    for (int mt = mt_open; mt >= 1 && !found; --mt) found = accept(0, 1, mt, 2);
    // 3) single-stage forms
    p.nwstages = nwstages_plan;
    for (int resident = 1; resident >= 0 && !found; --resident)
        for (int mt = 4; mt >= 1 && !found; --mt) {
            if (!resident && mt > mt_open) continue;
            if (resident && mt < 2 && k3) continue;   // a 1-sub-tile resident tile re-reads too much halo: stream instead
            found = accept(resident, 1, mt, 1);
        }
    if (!found) return 1;
    p.tile_rows = p.sub_stride * p.MT;
    p.ntiles = (int)((p.nrows + p.tile_rows - 1) / p.tile_rows);
    int ctas = num_sms;            // one persistent CTA per SM (the kernel owns all 512 TMEM columns)
    if (ctas > p.ntiles) ctas = p.ntiles;
    p.tiles_per_cta = (p.ntiles + ctas - 1) / ctas;
    ctas = (p.ntiles + p.tiles_per_cta - 1) / p.tiles_per_cta;
    p.idesc = (1u << 4) | ((uint32_t)(p.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);  // f16 x f16 -> f32, K-major A and B
    p.dbg_mode = 0;
#ifdef CD_MMA_DEBUG
    if (const char* e = getenv("CD_DBG_MODE")) p.dbg_mode = atoi(e);
#endif
    const int grid = ctas * p.ncb;
    cudaError_t e = cudaSuccess;
    if (k3) {
        switch (p.CB * 100 + cpg) {
            case 1602: e = launch_inst<16, true, 2>(p, grid, smem_total, max_smem, st); break;
            case 3204: e = launch_inst<32, true, 4>(p, grid, smem_total, max_smem, st); break;
            case 3208: e = launch_inst<32, true, 8>(p, grid, smem_total, max_smem, st); break;   // a 32-channel half of a 64-channel conv (8 per group)
            case 4806: e = launch_inst<48, true, 6>(p, grid, smem_total, max_smem, st); break;
            case 4812: e = launch_inst<48, true, 12>(p, grid, smem_total, max_smem, st); break;
            case 6408: e = launch_inst<64, true, 8>(p, grid, smem_total, max_smem, st); break;
            case 6416: e = launch_inst<64, true, 16>(p, grid, smem_total, max_smem, st); break;
            default: return 1;
        }
    } else {
        switch (p.CB) {
            case 16: e = launch_inst<16, false, 1>(p, grid, smem_total, max_smem, st); break;
            case 32: e = launch_inst<32, false, 1>(p, grid, smem_total, max_smem, st); break;
            case 48: e = launch_inst<48, false, 1>(p, grid, smem_total, max_smem, st); break;
            case 64: e = launch_inst<64, false, 1>(p, grid, smem_total, max_smem, st); break;
            case 80: e = launch_inst<80, false, 1>(p, grid, smem_total, max_smem, st); break;
            case 96: e = launch_inst<96, false, 1>(p, grid, smem_total, max_smem, st); break;
            case 112: e = launch_inst<112, false, 1>(p, grid, smem_total, max_smem, st); break;
            case 128: e = launch_inst<128, false, 1>(p, grid, smem_total, max_smem, st); break;
            default: return 1;
        }
    }
    return e == cudaSuccess ? 0 : 1;
}

#ifdef CD_MMA_DEBUG
extern "C" void cd_debug_counters(long long* out, int reset) {
    cudaMemcpyFromSymbol(out, g_mma_dbg, sizeof(long long) * 16);
    if (reset) { long long z[16] = {0}; cudaMemcpyToSymbol(g_mma_dbg, z, sizeof(z)); }
}
#endif

}  // namespace cd
