// Shared device/host definitions for the CaloDiffusion B200 denoiser library.
//
// Activation layout ("padded-linear, chunk-planar", PLCP), one tensor = C/8 chunk planes:
//   plane k holds, for every row g of the padded-linear voxel space, channels 8k..8k+7 as 8 halves
//   (16 bytes).  Rows enumerate (sample b, z in [0,Z], pp in [0,P+2), rr in [0,R+1)):
//       g = ((b*(Z+1) + z)*(P+2) + pp)*(R+1) + rr
//   pp = p+1 (pp=0 and pp=P+1 are circular phi copies of p=P-1 and p=0), rr = r (rr=R is a zero
//   column shared by r=R of this phi row and r=-1 of the next), z=Z is a zero plane shared by z=Z of
//   this sample and z=-1 of the next.  Every 3x3x3 neighbour of an interior voxel is then at the
//   constant row offset dz*(P+2)*(R+1) + dp*(R+1) + dr, which is what lets the tcgen05 implicit-GEMM
//   kernel fetch its A operand with plain 1-D bulk copies and address taps as shifted smem windows
//   (reference semantics: circular phi + zero Z/R padding, models/models.py:90-96).
//   `guard` zero rows precede row 0 and follow the last row of every chunk plane.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace cd {

struct Geom {       // one U-Net resolution level
    int Z, P, R;    // interior extents
    int Rp, Pp;     // R+1, P+2
    int plane;      // Pp*Rp rows per z-plane
    int rows_per_sample;  // (Z+1)*plane
    int guard;      // zero rows before/after the payload in each chunk plane
    long long chunk_stride;  // rows per chunk plane incl. guards (fixed by max batch)
    // exact division of n < 2^31 by plane / Rp / rows_per_sample:  q = (umulhi(n, mul) + n) >> shift
    uint32_t fd_mul[3], fd_shift[3];
};

__host__ inline void geom_init_fastdiv(Geom& g) {
    const int divs[3] = {g.plane, g.Rp, g.rows_per_sample};
    for (int i = 0; i < 3; ++i) {
        int l = 0;
        while ((1LL << l) < divs[i]) ++l;
        g.fd_mul[i] = (uint32_t)((((1ULL << l) - (unsigned long long)divs[i]) << 32) / (unsigned long long)divs[i] + 1ULL);
        g.fd_shift[i] = (uint32_t)l;
    }
}

struct TensorRef {  // a PLCP tensor
    __half* base;   // start of chunk plane 0 (guard included)
    int C;          // channels (multiple of 8)
};

__host__ __device__ inline long long plcp_offset(const Geom& g, int chunk, long long row) {
    // element offset (in halves) of channel 8*chunk of padded-linear row `row`
    return ((long long)chunk * g.chunk_stride + g.guard + row) * 8;
}

// Decode a padded-linear row. Returns kind: 0 interior, 1 phi-halo (src row returned), 2 zero row.
struct RowInfo {
    int b, z, pp, rr;
    int kind;
    long long src;  // for kind 1: the interior row this halo row mirrors; else the row itself
};

__device__ __forceinline__ RowInfo decode_row(const Geom& g, long long row) {
    RowInfo ri;
    int b = (int)(row / g.rows_per_sample);
    int rem = (int)(row - (long long)b * g.rows_per_sample);
    int z = rem / g.plane;
    int rem2 = rem - z * g.plane;
    int pp = rem2 / g.Rp;
    int rr = rem2 - pp * g.Rp;
    ri.b = b; ri.z = z; ri.pp = pp; ri.rr = rr;
    ri.src = row;
    if (z == g.Z || rr == g.R) {
        ri.kind = 2;
    } else if (pp == 0) {
        ri.kind = 1; ri.src = row + (long long)g.P * g.Rp;
    } else if (pp == g.P + 1) {
        ri.kind = 1; ri.src = row - (long long)g.P * g.Rp;
    } else {
        ri.kind = 0;
    }
    return ri;
}

// Row index within one sample (rem < rows_per_sample) -> z, pp, rr and kind, with 32-bit fast divisions.
struct SubRow { int z, pp, rr, kind; int src_delta; };
__device__ __forceinline__ uint32_t fdiv(const Geom& g, int which, uint32_t n) {
    return (__umulhi(n, g.fd_mul[which]) + n) >> g.fd_shift[which];
}
__device__ __forceinline__ SubRow decode_subrow(const Geom& g, uint32_t rem) {
    SubRow s;
    const uint32_t z = fdiv(g, 0, rem);
    const uint32_t rem2 = rem - z * (uint32_t)g.plane;
    const uint32_t pp = fdiv(g, 1, rem2);
    const uint32_t rr = rem2 - pp * (uint32_t)g.Rp;
    s.z = (int)z; s.pp = (int)pp; s.rr = (int)rr;
    s.src_delta = 0;
    if ((int)z == g.Z || (int)rr == g.R) s.kind = 2;
    else if (pp == 0) { s.kind = 1; s.src_delta = g.P * g.Rp; }
    else if ((int)pp == g.P + 1) { s.kind = 1; s.src_delta = -g.P * g.Rp; }
    else s.kind = 0;
    return s;
}

__host__ __device__ inline long long interior_row(const Geom& g, int b, int z, int p, int r) {
    return (((long long)b * (g.Z + 1) + z) * g.Pp + (p + 1)) * g.Rp + r;
}

// ---- conv descriptors -------------------------------------------------------------------------
struct ConvSrc {
    const __half* base;
    int C;
};

struct ConvParams {
    Geom gin, gout;
    ConvSrc src[2];
    int nsrc;
    int Cin_total;
    __half* dst;        // PLCP output (C_out channels)
    int Cout;
    const float* w;     // packed [tap][ci][co] fp32
    const float* bias;  // [co] or nullptr
    int kz, kp, kr;
    int sz, sp, sr;     // strides
    int pz, cp, pr;     // zero pad Z, circular pad phi, zero pad R   (transposed: padding 1, 3(kp-1), 1)
    int transposed;
    int write_halo;     // 1: produce all rows incl. halos (gather form); 0: interior rows only
    double* stats;      // [B][Cout][2] (sum, sumsq) or nullptr
    int stats_cpg;      // channels per GroupNorm group: producers may accumulate a whole group into its first channel slot
    // Folded PreNorm (1x1x1 tcgen05 form only): the input is x, the weights are W*diag(gamma) of a GroupNorm(1,C) and the
    // epilogue applies  y = rstd_b * acc - mean_b*rstd_b*pn_wsum[o] + pn_wb[o]  with the per-sample statistics pn_stats
    const double* pn_stats;   // [B][pn_C][2]: whole-sample (sum, sumsq) in the channel-0 slot; nullptr = plain conv
    const float* pn_wsum;     // [Cout] sum_c W'[o][c] (of the fp16-rounded folded weights)
    const float* pn_wb;       // [Cout] sum_c W[o][c]*beta[c]
    int pn_C;                 // channels of the normalised tensor
    double pn_count;          // elements per sample (C * voxels)
    // Optional C_out-channel PLCP tensor added to the result before rounding / statistics (3x3x3 tcgen05 form only): lets a
    // convolution over a wide concatenated input run as two resident-weight convolutions, the second accumulating onto the first.
    const __half* addsrc = nullptr;
    // channels per sample of the statistics array when this launch produces only a slice of the output channels (a 64 -> 64 conv
    // run as two 64 -> 32 halves: `dst`, `bias` and `stats` point at the slice, the per-sample stride stays the full width); 0: Cout
    int stats_C = 0;
    // Optional GroupNorm + SiLU + additive conditioning applied to the INPUT on its way into the GEMM (Z-stacked tcgen05 form only,
    // single source): src[0] is then the raw output of the previous convolution (phi halo rows written, zero rows zero) and
    // gn_stats its per-sample (sum, sumsq); the arithmetic is gn_apply_kernel's (act = SiLU), rounded to fp16 like its store.
    const double* gn_stats = nullptr;    // [B][Cin][2]
    const float* gn_gamma = nullptr; const float* gn_beta = nullptr;
    int gn_groups = 0;
    double gn_count = 0.0;               // voxels per sample (elements per channel)
    const float* gn_addvec = nullptr; int gn_addvec_stride = 0;    // [B][stride] conditioning vector at this block's offset, or nullptr
    int B;
};

// ---- small helpers ----------------------------------------------------------------------------
__device__ __forceinline__ float silu_f(float v) { return v / (1.0f + __expf(-v)); }
__device__ __forceinline__ float gelu_erf(float v) { return 0.5f * v * (1.0f + erff(v * 0.70710678118654752f)); }

struct alignas(16) Half8 { __half2 h[4]; };

// explicit 128-bit accesses: a struct copy through a Half8* may be split into four 32-bit accesses by the compiler
__device__ __forceinline__ void store8(void* p, const Half8& v) {
    uint4 u;
    u.x = *reinterpret_cast<const uint32_t*>(&v.h[0]); u.y = *reinterpret_cast<const uint32_t*>(&v.h[1]);
    u.z = *reinterpret_cast<const uint32_t*>(&v.h[2]); u.w = *reinterpret_cast<const uint32_t*>(&v.h[3]);
    *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ Half8 load8(const void* p) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    Half8 v;
    v.h[0] = *reinterpret_cast<const __half2*>(&u.x); v.h[1] = *reinterpret_cast<const __half2*>(&u.y);
    v.h[2] = *reinterpret_cast<const __half2*>(&u.z); v.h[3] = *reinterpret_cast<const __half2*>(&u.w);
    return v;
}

__device__ __forceinline__ void unpack8(const Half8& v, float* f) {
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 t = __half22float2(v.h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ Half8 pack8(const float* f) {
    Half8 v;
#pragma unroll
    for (int i = 0; i < 4; ++i) v.h[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
    return v;
}

}  // namespace cd
