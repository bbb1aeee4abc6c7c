// Launcher declarations for the hand-written kernels.
#pragma once
#include "cd_common.cuh"

namespace cd {

struct EmbedParams {
    const float* tw[3]; const float* tb[3];   // time MLP  Linear(1,h/2) GELU Linear(h/2,h) GELU Linear(h,h)
    const float* cw[3]; const float* cb[3];   // cond MLP  Linear(cs,hid) GELU Linear(hid,h) GELU Linear(h,h)
    int cond_size, cond_hidden, cond_dim, time_mode;  // time_mode 0: sigma/sqrt(1+sigma^2), 1: 0.5*log(sigma)
    int n_proj;
    const float* pw[16]; const float* pb[16]; int pC[16]; int poff[16];
    int addvec_stride;
};

struct FinalParams {
    Geom g;
    const __half* h; int C;
    const float* w; float bias;         // final conv weights (C) + bias
    const float* x;                     // denoiser input (B, voxels) fp32
    const float* sigma; int sigma_stride; float sigma_data;
    float* pred_out;                    // optional raw network output
    float* x0_out;                      // optional denoised output
    float* x_out;                       // optional sampler-updated state
    const float* coef;                  // k0..k4, [7] = d_scale (device)
    float* d_out;                       // optional: d_scale*(x - x0), the derivative multistep samplers keep
    const float* a1; const float* a2; const float* a3;
    int rng_enabled; unsigned long long seed; int shower_offset; int step;
    // optional device-resident Philox key {seed lo, seed hi, shower_offset} (overrides seed / shower_offset): keeps a captured
    // sampling graph independent of the per-call seed
    const uint32_t* rng_key;
    // optional fault word: atomicMin(eval_index) when the raw network output of a voxel is non-finite (fp16 range exceeded)
    int* fault; int eval_index;
};

// CSR linear map between the caller's flat voxel vector and the regular grid (device pointers)
struct GeomMap { const int* rowptr; const int* col; const float* val; int n_rows, n_cols; };

// dense ResNet of the layer-energy model; all matrices transposed to [in][out] fp32 (device pointers)
struct LayerMlpParams {
    int dim_in, hidden, n_hidden, cond_size, cond_dim, time_mode;
    const float* tw[3]; const float* tb[3];
    const float* cw[3]; const float* cb[3];
    const float* in_w; const float* in_b;
    const float* ew[8]; const float* eb[8];
    const float* w1[8]; const float* b1[8];
    const float* w2[8]; const float* b2[8];
    const float* out_w; const float* out_b;
};

struct RevNormParams {
    const float* vox; const float* e; int e_stride; const float* layerE;   // (B,total), (B,e_stride), (B,nseg+1) or nullptr
    float* out; float* energy_out;                                         // (B,total), (B) or nullptr
    int nseg, total;                      // calorimeter layers ("segments") per shower, values per shower
    const int* seg;                       // device [nseg+1] segment offsets, or nullptr: nseg equal segments
    int map_kind;                         // 0 reverse logit, 1 exp (log maps), 2 identity
    float a_scale, a_shift, alpha, log_cap, post_scale, post_shift;
    float totalE_mean, totalE_std, layers_mean, layers_std, guard_eps;
    int energy_mode;                      // 0: emin*(emax/emin)^e, 1: emin + (emax-emin)*e
    float emin, emax, max_deposit, ecut;
    int final_scale;                      // 1: multiply by max_deposit * E_inc (0 for an intermediate stage)
};
// sampled ("sparse") GLaM decoding: CSC of the positive decoder entries, see cd_kernels.cu
struct SparseDecParams {
    const int* colptr; const int* row; const float* val;   // per grid bin c: cells row[k] (global index l*N+n) with weights val[k]
    int n_cols, n_rows, cells_per_layer, bins_per_layer;
    const float* x; float* out;            // (B, n_cols) grid energies -> (B, n_rows) cell energies (zeroed by the launcher)
    int per_batch;                         // 1: one sampled matrix shared by the whole batch
    unsigned long long seed;
    const float* u_bank;                   // optional uniforms laid out like the reference's rand_like: [b][l][n][e]
};
void launch_sparse_decode(const SparseDecParams& p, int B, cudaStream_t st);
void launch_reverse_norm(const RevNormParams& p, int B, cudaStream_t st);

void launch_prep_input(const Geom& g, const float* x, const float* sigma, int sigma_stride, float sigma_data,
                       const float* r_vals, int rz_input, int phi_input, __half* dst, int out_chunks, int B, cudaStream_t st);
void launch_prep_input_geom(const Geom& g, const GeomMap& enc, const float* x, const float* sigma, int sigma_stride,
                            float sigma_data, const float* r_vals, int rz_input, int phi_input, __half* dst, int out_chunks,
                            int B, cudaStream_t st);
void launch_layer_mlp(const LayerMlpParams& m, const FinalParams& p, const float* cond, int B, int time_is_raw, cudaStream_t st);
void launch_geom_apply(const GeomMap& m, const float* x, float* y, int B, cudaStream_t st);
void launch_geom_decode(const FinalParams& p, const GeomMap& dec, const float* grid_pred, int B, cudaStream_t st);
void launch_conv_generic(const ConvParams& p, cudaStream_t st);
void launch_gn_apply(const Geom& g, const __half* src, const double* stats, const float* gamma, const float* beta, int groups,
                     double count_per_channel, const float* addvec, int addvec_stride, const __half* residual,
                     const __half* residual2, int res_split, __half* dst, int C, int act, double* out_stats, int B,
                     cudaStream_t st);
// linear attention + to_out 1x1x1 conv + GroupNorm(1,C) statistics (cd_attention.cu); 3 launches; returns 0 if C is supported
int launch_attention(const Geom& g, const __half* qkv, unsigned* kmax, float* ctx, const float* wout, const float* bout,
                     __half* dst, int C, double* stats, int B, cudaStream_t st);
bool attention_out_supported(int C);
// PreNorm + to_qkv + attention + to_out without the qkv tensor (cd_attention_fused.cu); 4 launches; returns 0 if launched
bool attention_fused_supported(int C);
int launch_attention_fused(const Geom& g, const __half* x, const __half* wf, const float* wsum, const float* wb,
                           const double* pn_stats, double pn_count, unsigned* kmax, float* ctx, float* part, __half* fold,
                           const float* wout, const float* bout, __half* dst, int C, double* stats, int B, cudaStream_t st);
constexpr int kAttnMaxBlocks = 32;   // block partials per sample merged by the online-softmax form
int attn_part_floats();
void launch_embed(const EmbedParams& p, const float* sigma, int sigma_stride, const float* cond, float* addvec, int B,
                  cudaStream_t st);
void launch_final(const FinalParams& p, int B, cudaStream_t st);
void launch_axpby(const float* x, const float* noise, float* out, const float* coef, long long n, int vps,
                  unsigned long long seed, int shower_offset, int step, int rng_enabled, cudaStream_t st,
                  const uint32_t* rng_key = nullptr);
void launch_copy_bytes(const void* src /* nullptr: clear */, void* dst, size_t bytes, cudaStream_t st);
void launch_fill_normal(float* out, long long n, int vps, unsigned long long seed, int shower_offset, int step,
                        cudaStream_t st);

// tcgen05 implicit-GEMM convolution (cd_conv_mma.cu); returns 0 if the shape is supported and was launched
int launch_conv_mma(const ConvParams& p, const __half* packed_w, cudaStream_t st);
// tcgen05 gather implicit GEMM for strided / transposed / thin-input convolutions (cd_conv_gather.cu)
int launch_conv_gather(const ConvParams& p, const __half* packed_w, cudaStream_t st);
bool gather_conv_supported(int Cin, int Cout);
size_t gather_packed_elems(int Cin, int Cout, int taps);
void gather_pack_weights(const float* w, int Cin, int Cout, int taps, __half* out);
// transposed stride-2 convs (k_r = 4): merged (r, r+1) output-pair taps, see cd_conv_gather.cu
size_t gather_packed_elems_pairT(int Cin, int Cout, int kz, int kp);
void gather_pack_weights_pairT(const float* w, int Cin, int Cout, int kz, int kp, int kr, __half* out);
bool mma_conv_supported(int Cin, int Cout, int taps);
size_t mma_packed_elems(int Cin, int Cout, int taps);
// w: fp32 [tap][ci][co]; out: fp16 blocks in the layout the tcgen05 kernel streams (see cd_conv_mma.cu)
void mma_pack_weights(const float* w, int Cin, int Cout, int taps, __half* out);

}  // namespace cd
