/* calodiffusion_b200 — C ABI of the B200-native CaloDiffusion denoiser / sampling-loop library.
 *
 * Drop-in boundary for the one hot path of OzAmram/CaloDiffusion: the U-Net denoiser evaluated at
 * every step of the reverse-diffusion sampling loop.  Plain pointers and sizes only; device pointers
 * are CUDA device addresses on the current device, `stream` is a cudaStream_t passed as void*.
 * All int-returning functions return 0 on success, non-zero on failure; cd_last_error() describes
 * the failure.  There is no CPU fallback: every compute entry point fails if no sm_100 device exists.
 *
 * Reference interfaces replaced (paths relative to the reference repo, calodiffusion/):
 *   cd_unet_create / cd_unet_set_weight / cd_unet_finalize
 *        -> CaloDiffusion.init_model (models/calodiffusion.py:39-81) + CondUnet.__init__
 *           (models/models.py:523-699) + load_state_dict (models/calodiffusion.py:31-37)
 *   cd_unet_forward  -> CaloDiffusion.forward (models/calodiffusion.py:86-98) = add_RZPhi + CondUnet.forward
 *           (models/models.py:701-748)
 *   cd_denoise       -> CaloDiffusion.denoise / __call__ (models/calodiffusion.py:154-173) with
 *           Loss.get_scaling (models/loss.py:29-41) and do_time_embed (models/calodiffusion.py:144-152)
 *   cd_sample_run    -> the per-step loops of DDim/DDPM.__call__ (models/sample.py:41-109) and
 *           EDMAbstract.for_loop + Euler/Heun.in_loop_sampler (models/sample.py:631-662, 784-821),
 *           expressed as a table of evaluations (see cd_eval) built by the host-side sampler classes
 *   cd_fill_normal   -> Diffusion.noise_generation (models/diffusion.py:58-61)
 *   cd_reverse_norm  -> ReverseNormCaloChall (utils/utils.py:446-573; called at models/diffusion.py:171-191) for the
 *           CaloChallenge '*logit-norm' shower maps (dataset 2 / 3)
 */
#ifndef CALODIFFUSION_B200_H
#define CALODIFFUSION_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CD_MAX_LEVELS 4

typedef struct cd_unet cd_unet; /* opaque handle: one model on one device */

typedef struct cd_unet_desc {
    int Z, P, R;                 /* SHAPE_FINAL[2:5] */
    int n_sizes;                 /* len(LAYER_SIZE_UNET) (= levels + 1), <= CD_MAX_LEVELS + 1 */
    int layer_sizes[CD_MAX_LEVELS + 1];
    int in_channels;             /* 1 + 2*R_Z_INPUT + PHI_INPUT */
    int rz_input, phi_input;
    int cond_size;               /* width of the conditioning vector E||layers */
    int cond_dim;                /* COND_SIZE_UNET */
    int groups;                  /* BLOCK_GROUPS (default 8) */
    int dataset_num;             /* DATASET_NUM: selects the R bin centres */
    int time_embed;              /* 0 = 'sigma', 1 = 'log' (TIME_EMBED) */
    float sigma_data;            /* 1.0 if 'log' in NOISE_SCHED else 0.5 (models/loss.py:16-25) */
    int max_batch;               /* largest batch a single call will carry */
    float r_centres[64];         /* scaled R bin centres (R values), utils/utils.py:134-150 */
} cd_unet_desc;

/* One denoiser evaluation inside a sampling loop plus the fused state update that follows it:
 *   [optional]  buf[x_in] = pre_a*buf[pre_src] + pre_b*noise(pre_noise_index)      (pre_src < 0: in place on buf[x_in])
 *   x0 = denoise(buf[x_in], sigma)
 *   [optional]  buf[x0_buf] = x0
 *   [optional]  buf[d_out] = d_scale*(buf[x_in] - x0)        (the ODE derivative (x - D)/sigma kept by multistep samplers)
 *   [optional]  buf[x_out] = k0*x0 + k1*buf[x_in] + k2*A1 + k3*buf[a2] + k4*buf[a3]
 * A1 is buf[a1] if a1 >= 0, noise slot `noise_index` if a1 == -2, absent if a1 == -1.
 * Noise comes from the caller's bank (noise_bank != NULL: slot i = bank + i*B*voxels) or from the
 * in-kernel Philox generator keyed by (seed, shower index, slot, voxel). */
typedef struct cd_eval {
    float sigma;
    int x_in, x_out, x0_buf;
    float k[5];
    int a1, a2, a3;
    int noise_index;
    float pre_a, pre_b;
    int pre_noise_index;
    int rec_x0;    /* >= 0: also copy x0 to x0s[rec_x0] */
    int rec_x;     /* >= 0: copy a state buffer to xs[rec_x] ... */
    int rec_x_buf; /* ... this buffer, after the update (a buffer that receives pre-noise is recorded before the noise) */
    int pre_src;   /* >= 0: the pre-noise step reads this buffer and writes buf[x_in]; -1: in place */
    int d_out;     /* >= 0: buf[d_out] = d_scale*(buf[x_in] - x0); -1: none */
    float d_scale;
} cd_eval;

#define CD_NUM_STATE_BUFS 8

const char* cd_last_error(void);
int cd_version(void);
/* 0 if a usable sm_100 device is current, else non-zero (and cd_last_error says why) */
int cd_device_check(void);

cd_unet* cd_unet_create(const cd_unet_desc* desc);

/* Layer-energy denoiser of LayerDiffusion (models/layerdiffusion.py:34-37 `layer_model = ResNet(dim_in = Z+1, num_layers = 5,
 * cond_size)`, 111-114 `layer_forward`; ResNet / ResDense models/models.py:372-457): a dense residual MLP over (B, dim_in)
 * vectors.  Returns the same handle type: cd_unet_set_weight (names "model.time_mlp.1.weight", "model.in_lay.weight",
 * "model.hidden_layers.0.dense1.0.weight", ...), cd_unet_finalize, cd_unet_forward, cd_denoise, cd_sample_run and
 * cd_unet_destroy apply unchanged, with voxels = dim_in and cond = E only (layerdiffusion.py:111-114). */
typedef struct cd_layer_desc {
    int dim_in;                  /* SHAPE_FINAL[2] + 1 (total energy + one value per layer), <= 64 */
    int hidden;                  /* ResNet hidden_dim (256), <= 256 */
    int num_layers;              /* ResNet num_layers (5): num_layers - 1 ResDense blocks, <= 9 */
    int cond_size;               /* 1 (3 for HGCal) */
    int cond_dim;                /* cond_emb_dim (128) */
    int time_embed;              /* 0 = 'sigma', 1 = 'log' */
    float sigma_data;
    int max_batch;
} cd_layer_desc;
cd_unet* cd_layer_model_create(const cd_layer_desc* desc);
void cd_unet_destroy(cd_unet* h);
/* fp32 host tensor under its reference state-dict name (e.g. "model.downs.0.0.block1.proj.conv.weight") */
int cd_unet_set_weight(cd_unet* h, const char* name, const float* host_data, const int64_t* shape, int ndim);
/* In-loop geometry embedding: the two linear maps between the caller's flat voxel vector (B, n_flat) and the regular
 * (Z,P,R) grid the U-Net works on, applied inside EVERY denoiser evaluation.  Replaces NNConverter.enc/dec
 * (utils/utils.py:610-653; Dataset 1, called from CaloDiffusion.forward models/calodiffusion.py:86-98) and
 * HGCalConverter.enc/dec = Embeder/Decoder.forward (utils/HGCal_utils.py:315-320, 340-349, 636-663; non-sparse decoding).
 * Each map is a host CSR matrix: which = 0 "enc" (n_rows = Z*P*R grid cells, n_cols = n_flat),
 * which = 1 "dec" (n_rows = n_flat, n_cols = Z*P*R).  Both must be set, before cd_unet_finalize; afterwards every x / pred /
 * x0 / state / noise tensor of cd_unet_forward, cd_denoise and cd_sample_run is (B, n_flat) instead of (B, Z*P*R). */
int cd_unet_set_geom_map(cd_unet* h, int which, int n_rows, int n_cols, const int32_t* rowptr, const int32_t* col,
                         const float* val);
/* One geometry map applied on its own, y (B, n_rows) = M x (B, n_cols), M a CSR matrix in DEVICE memory: the converters'
 * enc / dec outside the sampling loop (utils/utils.py:610-653, HGCal_utils.py:636-663 enc/dec, enc_batches/dec_batches). */
int cd_geom_apply(const int32_t* rowptr_dev, const int32_t* col_dev, const float* val_dev, int n_rows, int n_cols,
                  const float* x_dev, float* y_dev, int B, void* stream);
/* voxels per shower in the caller-facing tensors: n_flat when geometry maps are set, else Z*P*R */
int cd_unet_voxels(const cd_unet* h);
/* validates that every tensor of the topology was supplied, packs weights, allocates workspaces */
int cd_unet_finalize(cd_unet* h);
/* options: "conv_backend" 0 = tcgen05 implicit GEMM where supported (default), 1 = generic SIMT everywhere;
 *          "profile" 1 = time every op of each forward with CUDA events (eager, diagnostic), 2 = same + reset, 0 = off */
int cd_unet_set_option(cd_unet* h, const char* key, int value);
int64_t cd_unet_workspace_bytes(const cd_unet* h);
/* number of kernel launches issued by this handle since creation (graph replays count their nodes) */
int64_t cd_unet_launch_count(const cd_unet* h);

/* per-op timing gathered while "profile" is on: lines "name count total_ms flops bytes" (flops/bytes = algorithmic totals) */
int cd_unet_profile_report(const cd_unet* h, char* buf, int cap);

/* pred = CondUnet(add_RZPhi(x), cond, time); x and time (B values, already embedded) are used as given */
int cd_unet_forward(cd_unet* h, const float* x_dev, const float* time_dev, const float* cond_dev,
                    float* pred_dev, int B, void* stream);
/* x0 = c_skip*x + c_out*CondUnet(c_in*x, ...); sigma_dev holds B values */
int cd_denoise(cd_unet* h, const float* x_dev, const float* sigma_dev, const float* cond_dev,
               float* x0_dev, int B, void* stream);
/* buf[0] = start*start_scale, then the evaluations in order; x_final = buf[final_buf].
 * xs / x0s (optional, device) receive the per-step records.  use_graph: 0 eager launches,
 * 1 the whole loop captured and replayed as one CUDA graph (cached while the arguments repeat). */
int cd_sample_run(cd_unet* h, const float* start_dev, float start_scale, const float* cond_dev,
                  const cd_eval* evals, int n_evals, const float* noise_bank_dev, uint64_t seed,
                  int shower_offset, int final_buf, float* x_final_dev, float* xs_dev, float* x0s_dev,
                  int B, void* stream, int use_graph);
/* out[b][v] = N(0,1) from the library's Philox stream (slot = `slot`) */
int cd_fill_normal(float* out_dev, int64_t n_showers, int64_t voxels, uint64_t seed, int shower_offset,
                   int slot, void* stream);

/* Post-processing of sampled showers.  Constants come from utils/consts.py:dataset_params[dataset_num]. */
typedef struct cd_reverse_norm_desc {
    int Z, PR;                        /* calorimeter layers, voxels per layer (P*R) */
    float logit_mean, logit_std;
    float totalE_mean, totalE_std, layers_mean, layers_std;   /* used only with layerE ('layer' shower maps) */
    float emin, emax;                 /* incident energy = emin*(emax/emin)^e (logE) */
    float max_deposit, ecut;          /* ecut <= 0: no threshold */
} cd_reverse_norm_desc;
/* ReverseNormCaloChall for datasets 2 / 3 in network shape (utils/utils.py:446-573): vox_dev (B,Z,PR) network-space showers,
 * e_dev (B) scaled incident energies, layerE_dev (B,Z+1) = [totalE, layers] or NULL; writes out_dev (B,Z*PR) deposited
 * energies and, if not NULL, energy_out_dev (B) incident energies. */
int cd_reverse_norm(const cd_reverse_norm_desc* desc, const float* vox_dev, const float* e_dev, const float* layerE_dev,
                    float* out_dev, float* energy_out_dev, int B, void* stream);

/* General form of the same pipeline: ReverseNormCaloChall on the flat Dataset-0/1 voxel vector (ragged layers from the binning
 * file, utils/utils.py:549-558) and ReverseNormHGCal (utils/HGCal_utils.py:167-292; its three stages - inverse map on the grid,
 * GLaM decoding, per-layer renormalisation on the cells - are three calls with final_scale only on the last).
 *   data = inverse(v*a_scale + a_shift) * post_scale + post_shift ; ['layer'] per-segment renormalisation ; * max_deposit*E_inc */
typedef struct cd_reverse_norm_ex_desc {
    int nseg, total;                  /* layers per shower, values per shower (all layers) */
    int map_kind;                     /* 0 reverse logit, 1 exp of min(x, log_cap) ('log' maps), 2 identity */
    float a_scale, a_shift;           /* '*-norm': std, mean;  '*-scaled': 0.5*(max-min), 0.5*(max-min)+min */
    float alpha;                      /* reverse_logit alpha: 1e-6 (utils.py:233), 1e-8 (HGCal_utils.py:13) */
    float log_cap;                    /* log(max_deposit) */
    float post_scale, post_shift;     /* HGCalConverter.dec normalisation (embed_std, embed_mean); 1, 0 otherwise */
    float totalE_mean, totalE_std, layers_mean, layers_std;
    float guard_eps;                  /* layers / previous sums below this are not rescaled: 1e-6 (utils.py:528), 1e-8 (HGCal) */
    int energy_mode;                  /* 0: emin*(emax/emin)^e (logE), 1: emin + (emax-emin)*e */
    float emin, emax, max_deposit, ecut;
    int e_stride;                     /* floats per shower in e_dev (HGCal gen info has 3 columns; column 0 is used) */
    int final_scale;                  /* 1: multiply by max_deposit*E_inc */
} cd_reverse_norm_ex_desc;
/* seg_dev: device int32 [nseg+1] segment offsets into the `total` values of one shower, or NULL for nseg equal segments. */
int cd_reverse_norm_ex(const cd_reverse_norm_ex_desc* desc, const int32_t* seg_dev, const float* vox_dev, const float* e_dev,
                       const float* layerE_dev, float* out_dev, float* energy_out_dev, int B, void* stream);

/* Sampled ("sparse") GLaM decoding (HGCal_utils.py:340-349 with sparse_decoding, generate_sparse_mat :355-407): every grid
 * bin's energy is split equally over a random subset of its cells - cell n of bin e is kept when u + w[n,e] > 1 (u uniform,
 * w the decoder weight, only w > 1e-6 take part) and the entry with the largest u + w is always kept.  The decoder comes as
 * a CSC matrix over (grid bin -> cells) holding its positive entries.  per_batch != 0: one draw shared by the batch.
 * u_bank_dev (optional, tests): the uniforms, laid out like the reference's rand_like(in_mat): [b or 0][layer][cell][bin]. */
int cd_geom_decode_sparse(const int32_t* colptr_dev, const int32_t* row_dev, const float* val_dev, int n_cols, int n_rows,
                          int cells_per_layer, int bins_per_layer, const float* x_dev, float* out_dev, int B, int per_batch,
                          uint64_t seed, const float* u_bank_dev, void* stream);

/* ---- training step -------------------------------------------------------------------------------------
 * Replaces, for TRAINING_OBJ hybrid_weight / LOSS_TYPE l2, the body of train/train_diffusion.py:42-66:
 *   noise -> x = data + sigma*noise -> x0 = denoise(x) -> L = sum w (x0-data)^2 / (mean(w) numel), w = 1 + 1/sigma^2
 *   (models/loss.py:118-142, 168-179, 103-104) -> backward -> optimizer step (torch.optim.Adam, train/train.py:144).
 * The handle must be created as usual and given cd_unet_set_option(h, "train", 1) before cd_unet_finalize.  Parameters,
 * gradients and the Adam moments live in four flat fp32 device buffers (PyTorch layouts; cd_train_param_info gives every
 * parameter's name / offset / size; cd_train_buffer(h, 0..3) the buffers P, G, M, V) so that a data-parallel caller can
 * all-reduce G in place with one ncclAllReduce.  Gradients in G carry `loss_scale` (activation gradients are fp16). */
int cd_train_step(cd_unet* h, const float* data_dev /* (B, voxels) clean showers */, const float* noise_dev, const float* sigma_dev /* (B) */,
                  const float* cond_dev /* (B, cond_size) */, float loss_scale, int B, void* stream);   /* forward + backward; G, loss */
int cd_train_loss(cd_unet* h, double* loss_host, void* stream);                     /* loss of the last step (synchronises) */
int cd_train_adam(cd_unet* h, float lr, float beta1, float beta2, float eps, float loss_scale, void* stream);   /* P -= Adam(G / loss_scale) */
int cd_train_refresh(cd_unet* h, void* stream);                                     /* re-derive the kernels' weight layouts after P changed */
int64_t cd_train_flat_size(const cd_unet* h);
int cd_train_num_params(const cd_unet* h);
int cd_train_param_info(const cd_unet* h, int index, char* name, int cap, int64_t* offset, int64_t* numel);
void* cd_train_buffer(cd_unet* h, int which /* 0 P, 1 G, 2 M, 3 V, 4 x0 of the last step (B, voxels) */);

/* test hooks: one layer in isolation on PLCP tensors held by the handle (used by tests/ only) */
int cd_test_conv(cd_unet* h, int level_in, int level_out, int Cin, int Cout, const int* ksize,
                 const int* stride, int transposed, const float* w_host, const float* bias_host,
                 const float* x_dev /* (B,Cin,Z,P,R) fp32 */, float* y_dev /* (B,Cout,Zo,Po,Ro) fp32 */,
                 int B, int backend /* 0 R-stacked tcgen05, 1 SIMT, 2 gather tcgen05, 3 Z-stacked tcgen05 */, void* stream);
int cd_level_shape(const cd_unet* h, int level, int* zpr /* [3] */);

/* cd_conv_noalloc: no counterpart in the reference - added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same single
 * CylindricalConv / CylindricalConvTrans as cd_test_conv's SIMT backend (backend 1), but every byte of scratch - the repacked
 * fp32 weights and the two fp16 staging tensors in the engine's activation layout - lives in ONE static device array of
 * cd_conv_noalloc_max_bytes() = 16 MiB that every GPU loading the library holds until exit.  The call allocates nothing
 * (no cudaMalloc / cudaFree / allocating library routine), does not synchronise, and only enqueues work on `stream`; the handle
 * is read for the level geometry only.  Weights come in the PyTorch layout ON THE DEVICE: (Cout,Cin,kz,kp,kr), transposed
 * (Cin,Cout,kz,kp,kr); bias_dev (Cout) or NULL.  Needs taps*Cin*Cout*4 + (Cin + Cout)/8 * 16 * rows(level, max_batch) bytes
 * (each part rounded up to 256); a request beyond the 16 MiB is rejected with an error naming the three parts.  All calls on
 * a device share the array, which keeps the previous call's contents (the staging tensors are cleared by every call): calls
 * that overlap on the GPU - two streams, or two threads - race; serialise them on one stream. */
int64_t cd_conv_noalloc_max_bytes(void);
int cd_conv_noalloc(const cd_unet* h, int level_in, int level_out, int Cin, int Cout, const int* ksize, const int* stride,
                    int transposed, const float* w_dev, const float* bias_dev, const float* x_dev /* (B,Cin,Z,P,R) fp32 */,
                    float* y_dev /* (B,Cout,Zo,Po,Ro) fp32 */, int B, void* stream);

#ifdef __cplusplus
}
#endif
#endif
