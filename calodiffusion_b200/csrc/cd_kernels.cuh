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
    const float* bias_ptr;              // optional device copy of the bias (training: read in place from the master parameters)
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
// Z-stacked tcgen05 conv (cd_conv_zstack.cu): 3x3x3 stride 1, C_out = 32; force = 1 takes it whatever its row efficiency
bool zstack_conv_supported(int Cin, int Cout, int taps);
size_t zstack_packed_elems(int Cin, int Cout, int taps);
void zstack_pack_weights(const float* w, int Cin, int Cout, __half* out);
int launch_conv_zstack(const ConvParams& p, const __half* packed_w, int force, cudaStream_t st);
bool zstack_geom_efficient(const Geom& g);     // the level's plane fills its 128-row strips well enough for the Z-stacked form


// ---- training step (cd_backward.cu) ---------------------------------------------------------------
constexpr int kBwdMaxC = 128;

struct LossBwdParams {
    Geom g;
    const __half* h; int C;                 // input of the final 1x1x1 conv
    const float* w; const float* bias;      // its weights (C) and bias (device)
    const float* x; const float* data;      // noised showers fed to the denoiser, clean showers (B, voxels) fp32
    const float* sigma; float sigma_data;
    const float* coef; const float* lossw;  // per shower: loss_scale*2*w_b/(mean(w)*numel), w_b/(mean(w)*numel)
    __half* dh;                             // gradient of h (interior rows)
    float* dw; float* db; double* loss; float* x0_out;
};

struct GnBwdParams {
    Geom g;
    const __half* x; const __half* dy;      // pre-norm tensor, gradient of the op's output
    const double* stats; const float* gamma; const float* beta;
    int C, groups, act; double count_per_channel;
    float* acc;                             // scratch [B][C][4]
    float* m12;                             // scratch [B][groups][2]
    float* mr;                              // scratch [B][C][2]: mean, rstd of the channel's group
    float* dgamma; float* dbeta; float* dbias;      // parameter gradients (dbias: bias of the convolution that produced x, or null)
    float* daddvec; int addvec_stride;      // gradient of the additive conditioning (already offset to this block), or null
    __half* dx; int dx_accumulate;
    __half* dres; __half* dres2; int res_split, dres_accumulate, dres2_accumulate;   // residual gradient(s), or null
    int rows_per_block;
};

struct WgradParams {
    Geom g;
    const __half* x0; const __half* x1; int src_chunks0;   // forward input (one or two concatenated sources)
    const __half* dy;                                      // gradient of the conv output (Cout channels)
    float* dw;                                             // master layout [co][ci][tap], Cin_real input channels
    int Cin_pad, Cin_real, Cout, taps, B, ncob;
};

struct AttnBwdParams {
    Geom g;
    const __half* qkv; const __half* dy;    // raw 96-channel qkv tensor, gradient of the to_out conv output (C channels)
    const unsigned* kmax; const float* ctx; // forward statistics: ordered-uint max of k, [B][33][32] context + S
    const float* wout; int C;               // to_out weights [C][32]
    float* dcx;                             // scratch [B][32][32]
    float* dwout; float* dbout;             // parameter gradients
    __half* dqkv;                           // gradient of the qkv tensor (96 channels)
    int rows_per_block;
};

struct EmbedBwdLayout { int stride, cvec, dt3, t2, dc3, c2, dt2, t1, dt1, t0, dc2, c1, dc1, c0; };

struct RepackParams {
    const float* master; float* out_f; __half* out_h;
    long long n; int mode, taps, Cin_real, Cin_pad, Cout, transposed, CB, ci_lo, n_out;
    // modes 4 / 5 (gather-tcgen05 layouts): pk = the fp32 [tap][Cin_pad][Cout] array of mode 0; swap: pack the convolution with
    // the channel roles exchanged (the data gradient of a strided / transposed conv is the opposite kind with the same taps)
    const float* pk; int swap, kz, kp, kr;
};

struct LinearJob { const float* dY; const float* X; float* dW; float* dbias; long long sy, sx; int n_out, n_in; };
void launch_linear_wgrad_table(const LinearJob* table_dev, int njobs, int max_n, int B, cudaStream_t st);
void launch_repack_table(const RepackParams* table_dev, int njobs, long long max_n, cudaStream_t st);
void launch_loss_backward(const LossBwdParams& p, int B, cudaStream_t st);
void launch_gn_backward(GnBwdParams p, int B, cudaStream_t st);
int launch_wgrad_mma(const WgradParams& p, cudaStream_t st);
void launch_conv_strided_backward(const ConvParams& p, const __half* x, const __half* dy, __half* dx, int dx_accumulate, float* dw,
                                  int Cin_real, int master_transposed, cudaStream_t st);
void launch_channel_sum(const Geom& g, const __half* t, int C, float* out, int B, cudaStream_t st);
void launch_plcp_accumulate(const Geom& g, const __half* src, __half* dst, int C, int accumulate, int B, cudaStream_t st);
int launch_attn_backward(AttnBwdParams p, int B, cudaStream_t st);
void launch_embed_backward(const EmbedParams& p, const EmbedBwdLayout& L, const float* sigma, const float* cond, const float* daddvec,
                           float* rec, int B, cudaStream_t st);
void launch_linear_wgrad(const float* dY, long long sy, const float* X, long long sx, int B, int n_out, int n_in, float* dW, float* dbias,
                         cudaStream_t st);
void launch_adam(float* w, float* m, float* v, float* g, long long n, float lr, float beta1, float beta2, float eps, int step,
                 float inv_scale, cudaStream_t st);
void launch_repack_conv(const RepackParams& p, cudaStream_t st);
void launch_fill_zero(float* p, long long n, cudaStream_t st);
int mma_pick_cb(int Cout, int taps);     // channel block the tcgen05 weight layout uses (cd_conv_mma.cu)

}  // namespace cd
