// Engine: builds the CondUnet evaluation plan from a cd_unet_desc + reference-named weights, owns the
// PLCP activation workspaces, and exposes the C ABI of include/calodiffusion_b200.h.
#include <cmath>
#include <csignal>
#include <execinfo.h>
#include <unistd.h>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/calodiffusion_b200.h"
#include "cd_kernels.cuh"

namespace cd {

static thread_local char g_err[1024] = "";
static int fail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}
#define CD_CUDA(call)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

struct HostTensor {
    std::vector<float> data;
    std::vector<int64_t> shape;
};

struct ConvW {
    float* w = nullptr;      // [tap][ci][co] fp32 (generic kernel)
    float* bias = nullptr;   // [co]
    __half* wmma = nullptr;  // fp16 blocks for the R-stacked tcgen05 kernel (stride-1 3x3x3 / 1x1x1 convs)
    __half* wgat = nullptr;  // fp16 [tap][kc][co][8] for the gather tcgen05 kernel (strided / transposed / thin-input convs)
    // 64 -> 64 stride-1 3x3x3 conv as two 64 -> 32 halves (output channels 0-31 / 32-63): each half's 110 KB of weights stays
    // resident in shared memory (K-staged form, 0.62 of peak) where the full 221 KB has to stream through the weight ring (0.39)
    __half* wmma_lo = nullptr; __half* wmma_hi = nullptr; float* bias_hi = nullptr;
    __half* wz = nullptr;        // Z-stacked tcgen05 layout (cd_conv_zstack.cu): 3x3x3, C_out = 32, sampling handles only
    int Cin = 0, Cout = 0, kz = 1, kp = 1, kr = 1;
    int transposed = 0;
};

struct Run {  // per-call state
    int B;
    cudaStream_t st;
    const float* x;        // (B, voxels) fp32 denoiser input
    const float* sigma;    // per-sample (stride 1) or broadcast (stride 0)
    int sigma_stride;
    const float* cond;     // (B, cond_size)
    int scale_input;       // 1: multiply x by c_in (denoise); 0: raw forward
    int time_is_raw;       // 1: `sigma` already holds the embedded time value (CaloDiffusion.forward's `time`)
    FinalParams fin;       // output routing of the final kernel
};

struct Buf { int level, C; __half* ptr; bool busy; __half* grad = nullptr; };

// ---- training tape (option "train" before cd_unet_finalize): what the backward pass needs to know about every op ----
struct TapeOp {
    int kind = 0;                 // 0 conv, 1 GroupNorm(+act+cond+residual), 2 attention core, 3 final conv + loss, 4 init marker
    // conv
    std::string wname, bname;
    int Cin_real = 0, Cin_pad = 0, Cout = 0, kz = 1, kp = 1, kr = 1, stride = 1, transposed = 0, level_in = 0, level_out = 0;
    int pz = 0, cp = 0, pr = 0;
    std::vector<int> srcs; int dst = -1;
    float* w = nullptr; __half* wmma = nullptr;              // forward layouts (refreshed from the master after every optimizer step)
    float* wd[2] = {nullptr, nullptr}; __half* wdm[2] = {nullptr, nullptr};   // data-gradient layouts, one per source
    __half* wgat = nullptr; __half* wgat_d = nullptr;        // gather-tcgen05 layouts: forward, and the data gradient (opposite conv kind)
    bool need_dgrad = true, bias_via_gn = false;
    // GroupNorm
    std::string gname, betaname;
    int src = -1, C = 0, groups = 1, act = 0, addvec_off = -1, residual = -1, residual2 = -1, conv_of_src = -1;
    size_t stats_off = 0;
    // attention core
    int qkv = -1, y = -1; size_t skm = 0, sctx = 0; std::string woname, boname;
};
struct ParamSlot { std::string name; long long offset, numel; };

constexpr int kRow = 12;   // floats per sampler coefficient-table row: sigma, k0..k4, pre_a, pre_b, d_scale, 3 spare

struct OpInfo { std::string name; double flops, bytes; };          // per shower
struct ProfAcc { long long count = 0; double ms = 0, flops = 0, bytes = 0; };

}  // namespace cd

using namespace cd;

struct cd_unet {
    cd_unet_desc d;
    int n_levels = 0;
    Geom geom[CD_MAX_LEVELS];
    std::map<std::string, HostTensor> weights;
    std::vector<void*> dev_allocs;
    std::vector<Buf> bufs;
    std::vector<std::function<void(Run&)>> ops;
    std::vector<cd::OpInfo> op_info;      // parallel to ops: name + algorithmic flops/bytes per shower
    std::map<std::string, cd::ProfAcc> prof;
    int profile = 0;
    int debug_sync = 0;
    bool dense = false;           // layer-energy ResNet (cd_layer_model_create) instead of the U-Net
    cd_layer_desc ld = {};
    LayerMlpParams lm = {};
    int online_attention = 1;     // option "online_attention" 0: separate max pass + atomically merged context
    float* attn_part = nullptr;   // [max_batch][kAttnMaxBlocks][attn_part_floats()] block partials of the online form (not zeroed)
    int no_fused_attention = 0;   // option "fused_attention" 0: materialise qkv (1x1x1 conv) and run the three-pass attention on it
    std::vector<cudaEvent_t> prof_evs;
    bool finalized = false;
    int conv_backend = 0;
    // training
    bool train = false;
    std::vector<cd::TapeOp> tape;
    std::vector<cd::ParamSlot> params;
    std::map<std::string, int> param_index;
    float* P = nullptr; float* G = nullptr; float* M = nullptr; float* V = nullptr;   // flat master params, grads, Adam moments
    long long n_params = 0;
    int adam_step = 0;
    float* gn_acc = nullptr; float* gn_m12 = nullptr; float* attn_dcx = nullptr; float* daddvec = nullptr; float* embed_rec = nullptr;
    float* x_noisy = nullptr; float* x0_buf = nullptr; float* coef = nullptr; double* loss_dev = nullptr;
    cd::EmbedBwdLayout embed_layout = {};
    cd::LinearJob* linear_tab = nullptr; int linear_jobs = 0, linear_max_n = 0;                 // device table of the Linear weight-gradient jobs
    cd::RepackParams* repack_tab = nullptr; int repack_jobs = 0; long long repack_max_n = 0;   // device table of weight re-layout jobs
    std::vector<__half*> bwd_tmp;  // per level: scratch gradient tensor for data gradients accumulated outside the conv epilogue
    int tin_buf_index = -1;
    std::vector<std::string> proj_names;        // state-dict prefix of every conditioned block's Linear (e.g. model.downs.0.0.mlp.1)
    float* pptr(const std::string& name) { auto it = param_index.find(name); return it == param_index.end() ? nullptr : P + params[it->second].offset; }
    float* gptr(const std::string& name) { auto it = param_index.find(name); return it == param_index.end() ? nullptr : G + params[it->second].offset; }
    int64_t workspace_bytes = 0;
    int64_t launches = 0;
    int launches_per_forward = 0;
    // scratch (zeroed per forward): per-sample byte size and base
    unsigned char* scratch = nullptr;
    size_t scratch_per_sample = 0;
    // persistent small buffers
    float* r_vals = nullptr;
    float* addvec = nullptr;  // [maxB][addvec_stride]
    int addvec_stride = 0;
    float* final_w = nullptr;
    float final_b = 0.f;
    int final_C = 0;
    const __half* final_h = nullptr;
    EmbedParams embed;
    // in-loop geometry embedding (cd_unet_set_geom_map): host CSR until finalize, device CSR afterwards
    struct HostCsr { std::vector<int> rowptr, col; std::vector<float> val; int n_rows = 0, n_cols = 0; } geom_host[2];
    GeomMap geom_map[2] = {};
    bool has_geom = false;
    float* grid_pred = nullptr;   // (max_batch, Z*P*R) raw network output before decoding
    int voxels() const { return has_geom ? geom_map[1].n_rows : d.Z * d.P * d.R; }
    // sampling state
    float* state[CD_NUM_STATE_BUFS] = {};
    float* coef_tab = nullptr;  // [max_evals][kRow]; row n_evals: {start_scale, 0, seed lo, seed hi, shower_offset, ...} (raw bits)
    int coef_cap = 0;
    int* fault = nullptr;       // device word: first evaluation whose raw network output was non-finite (INT_MAX: none)
    int* fault_host = nullptr;  // pinned mirror
    int check_finite = 1;       // option "check_finite": cd_sample_run reads the fault word back and fails on a fault
    float* unit_sigma = nullptr;
    // graph cache
    cudaGraphExec_t graph_exec = nullptr;
    cudaStream_t capture_stream = nullptr;
    std::vector<unsigned char> graph_key;
    int64_t graph_nodes = 0;

    template <typename T>
    T* dalloc(size_t n, bool zero = true) {
        void* p = nullptr;
        if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
        if (zero) cudaMemset(p, 0, n * sizeof(T));
        dev_allocs.push_back(p);
        workspace_bytes += (int64_t)(n * sizeof(T));
        return (T*)p;
    }
    float* upload(const std::vector<float>& v) {
        float* p = dalloc<float>(v.size(), false);
        if (p) cudaMemcpy(p, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice);
        return p;
    }
};

namespace cd {

static Geom make_geom(int Z, int P, int R, int max_batch) {
    Geom g;
    g.Z = Z; g.P = P; g.R = R;
    g.Rp = R + 1; g.Pp = P + 2;
    g.plane = g.Pp * g.Rp;
    g.rows_per_sample = (Z + 1) * g.plane;
    g.guard = ((g.plane + g.Rp + 1 + 7) / 8) * 8 + 1024;
    long long rows = (long long)max_batch * g.rows_per_sample + 2LL * g.guard;
    g.chunk_stride = (rows + 7) / 8 * 8;
    geom_init_fastdiv(g);
    return g;
}

static int acquire(cd_unet* h, int level, int C) {
    // training keeps every activation for the backward pass: no reuse
    for (size_t i = 0; i < h->bufs.size() && !h->train; ++i)
        if (!h->bufs[i].busy && h->bufs[i].level == level && h->bufs[i].C == C) { h->bufs[i].busy = true; return (int)i; }
    Buf b;
    b.level = level; b.C = C; b.busy = true;
    size_t n = (size_t)(C / 8) * (size_t)h->geom[level].chunk_stride * 8;
    b.ptr = h->dalloc<__half>(n, true);
    if (h->train) b.grad = h->dalloc<__half>(n, true);
    h->bufs.push_back(b);
    return (int)h->bufs.size() - 1;
}
static void release(cd_unet* h, int idx) { h->bufs[idx].busy = false; }

static const HostTensor* find_w(cd_unet* h, const std::string& name) {
    auto it = h->weights.find(name);
    return it == h->weights.end() ? nullptr : &it->second;
}

// pack a conv weight; returns false (with error set) if missing / wrong shape
static bool pack_conv(cd_unet* h, const std::string& wname, const std::string& bname, int Cin_real, int Cin_pad, int Cout,
                      int kz, int kp, int kr, int transposed, bool want_mma, ConvW* out, bool want_gather = false) {
    const HostTensor* w = find_w(h, wname);
    if (!w) { fail("missing weight %s", wname.c_str()); return false; }
    int64_t e0 = transposed ? Cin_real : Cout, e1 = transposed ? Cout : Cin_real;
    if (w->shape.size() != 5 || w->shape[0] != e0 || w->shape[1] != e1 || w->shape[2] != kz || w->shape[3] != kp ||
        w->shape[4] != kr) {
        fail("weight %s has unexpected shape", wname.c_str());
        return false;
    }
    int taps = kz * kp * kr;
    std::vector<float> pk((size_t)taps * Cin_pad * Cout, 0.f);
    for (int co = 0; co < Cout; ++co)
        for (int ci = 0; ci < Cin_real; ++ci)
            for (int t = 0; t < taps; ++t) {
                size_t src = transposed ? (((size_t)ci * Cout + co) * taps + t) : (((size_t)co * Cin_real + ci) * taps + t);
                pk[((size_t)t * Cin_pad + ci) * Cout + co] = w->data[src];
            }
    out->w = h->upload(pk);
    out->Cin = Cin_pad; out->Cout = Cout; out->kz = kz; out->kp = kp; out->kr = kr; out->transposed = transposed;
    if (!bname.empty()) {
        const HostTensor* b = find_w(h, bname);
        if (!b) { fail("missing weight %s", bname.c_str()); return false; }
        if ((int)b->data.size() != Cout) { fail("bias %s has unexpected size", bname.c_str()); return false; }
        out->bias = (h->train && h->pptr(bname)) ? h->pptr(bname) : h->upload(b->data);
    }
    if (want_mma && !transposed && mma_conv_supported(Cin_pad, Cout, taps)) {
        std::vector<__half> hp(mma_packed_elems(Cin_pad, Cout, taps));
        mma_pack_weights(pk.data(), Cin_pad, Cout, taps, hp.data());
        out->wmma = h->dalloc<__half>(hp.size(), false);
        if (out->wmma) cudaMemcpy(out->wmma, hp.data(), hp.size() * sizeof(__half), cudaMemcpyHostToDevice);
    }
    // Z-stacked layout for the large-plane levels (taken per launch where the strip decomposition is row-efficient): not in
    // training (the weight layouts are re-derived on the device there)
    if (want_mma && !transposed && !h->train && zstack_conv_supported(Cin_pad, Cout, taps) && !getenv("CD_NO_ZSTACK")) {
        std::vector<__half> hp(zstack_packed_elems(Cin_pad, Cout, taps));
        zstack_pack_weights(pk.data(), Cin_pad, Cout, hp.data());
        out->wz = h->dalloc<__half>(hp.size(), false);
        if (out->wz) cudaMemcpy(out->wz, hp.data(), hp.size() * sizeof(__half), cudaMemcpyHostToDevice);
    }
    // output-channel halves of a 64 -> 64 conv (see ConvW): not in training (the weight layouts are re-derived on the device there)
    if (want_mma && !transposed && !h->train && taps == 27 && Cin_pad == 64 && Cout == 64 && out->wmma && out->bias &&
        mma_conv_supported(Cin_pad, Cout / 2, taps) && !getenv("CD_NO_OSPLIT")) {
        const int half = Cout / 2;
        std::vector<float> ph((size_t)taps * Cin_pad * half);
        std::vector<__half> hp(mma_packed_elems(Cin_pad, half, taps));
        __half** dstp[2] = {&out->wmma_lo, &out->wmma_hi};
        for (int part = 0; part < 2; ++part) {
            for (size_t r = 0; r < (size_t)taps * Cin_pad; ++r)
                for (int co = 0; co < half; ++co) ph[r * half + co] = pk[r * Cout + part * half + co];
            mma_pack_weights(ph.data(), Cin_pad, half, taps, hp.data());
            *dstp[part] = h->dalloc<__half>(hp.size(), false);
            if (*dstp[part]) cudaMemcpy(*dstp[part], hp.data(), hp.size() * sizeof(__half), cudaMemcpyHostToDevice);
        }
        out->bias_hi = out->bias + half;
        if (!out->wmma_lo || !out->wmma_hi) { out->wmma_lo = out->wmma_hi = nullptr; }
    }
    if (want_gather && gather_conv_supported(Cin_pad, Cout) && taps <= 64) {   // (training re-derives it on the device: repack modes 4 / 5)
        const bool pairT = transposed && kr == 4;     // transposed convs use the merged output-pair taps
        std::vector<__half> hp(pairT ? gather_packed_elems_pairT(Cin_pad, Cout, kz, kp) : gather_packed_elems(Cin_pad, Cout, taps));
        if (pairT) gather_pack_weights_pairT(pk.data(), Cin_pad, Cout, kz, kp, kr, hp.data());
        else gather_pack_weights(pk.data(), Cin_pad, Cout, taps, hp.data());
        out->wgat = h->dalloc<__half>(hp.size(), false);
        if (out->wgat) cudaMemcpy(out->wgat, hp.data(), hp.size() * sizeof(__half), cudaMemcpyHostToDevice);
    }
    return true;
}

static float* upload_vec(cd_unet* h, const std::string& name, int n) {
    const HostTensor* t = find_w(h, name);
    if (!t) { fail("missing weight %s", name.c_str()); return nullptr; }
    if ((int)t->data.size() != n) { fail("weight %s has unexpected size %zu (want %d)", name.c_str(), t->data.size(), n); return nullptr; }
    if (h->train && h->pptr(name)) return h->pptr(name);      // training: every fp32 consumer reads the master copy in place
    return h->upload(t->data);
}

}  // namespace cd

// -------------------------------------------------------------------------------------------------
// plan construction
// -------------------------------------------------------------------------------------------------
namespace cd {

// A GroupNorm + SiLU (+ conditioning) whose only consumer is the next convolution: the Z-stacked tcgen05 conv applies it to its
// staged input windows (cd_conv_zstack.cu), otherwise the op launches gn_apply first.  raw = the normalised tensor's source,
// the conv's own source buffer is then where gn_apply would put its result.
struct FusedGn {
    int raw = -1;
    float* gamma = nullptr; float* beta = nullptr;
    int groups = 0; long long stats_off = -1; int addvec_off = -1;
};


struct Builder {
    cd_unet* h;
    bool ok = true;
    size_t scratch_prefix = 0;  // bytes per sample
    // training: names / real input width of the conv about to be added (set by the call sites right before conv())
    std::string t_wname, t_bname; int t_cin_real = 0;
    void tag(const std::string& w, const std::string& b, int cin_real) { t_wname = w; t_bname = b; t_cin_real = cin_real; }

    size_t slot(size_t bytes_per_sample) {
        size_t off = scratch_prefix;
        scratch_prefix += (bytes_per_sample + 15) / 16 * 16;
        return off;
    }

    // conv: srcs are buffer indices at level_in; dst buffer at level_out
    void conv(const ConvW& cw, int level_in, int level_out, std::vector<int> srcs, int dst, int stride, int pz, int cp, int pr,
              int write_halo, long long stats_off /* -1: none */, FusedGn fg = FusedGn()) {
        cd_unet* hh = h;
        ConvW w = cw;
        static const int zs_force = getenv("CD_ZS_FORCE") ? 1 : 0;     // tests: Z-stacked form whatever the row efficiency / batch
        if (hh->train) {
            TapeOp t;
            t.kind = 0; t.wname = t_wname; t.bname = t_bname; t.Cin_real = t_cin_real ? t_cin_real : w.Cin; t.Cin_pad = w.Cin; t.Cout = w.Cout;
            t.kz = w.kz; t.kp = w.kp; t.kr = w.kr; t.stride = stride; t.transposed = w.transposed; t.level_in = level_in; t.level_out = level_out;
            t.pz = pz; t.cp = cp; t.pr = pr; t.srcs = srcs; t.dst = dst; t.w = w.w; t.wmma = w.wmma; t.wgat = w.wgat;
            hh->tape.push_back(t);
            t_wname.clear(); t_bname.clear(); t_cin_real = 0;
        }
        {
            const Geom& go = hh->geom[level_out];
            const Geom& gi = hh->geom[level_in];
            double mo = (double)go.Z * go.P * go.R, mi = (double)gi.Z * gi.P * gi.R;
            int taps = w.kz * w.kp * w.kr;
            char nm[96];
            // ("gn": this step also applies the GroupNorm + SiLU + conditioning of its input, fused or as a pass of its own)
            snprintf(nm, sizeof(nm), "conv_k%d%s_c%d_%d_L%d%s", taps, w.transposed ? "T" : (stride > 1 ? "s2" : (fg.raw >= 0 ? "gn" : "")), w.Cin, w.Cout,
                     level_out, (w.wmma && stride == 1 && !w.transposed) ? "_mma" : (w.wgat ? "_gather" : "_simt"));
            // FLOP = 2*MAC as executed: every output voxel x all taps for a (strided) conv; every INPUT voxel x all
            // taps for a transposed conv (an output voxel of a stride-2 transposed conv receives 1/8 of the taps); bytes = fp16 in + out
            hh->op_info.push_back({nm, 2.0 * (w.transposed ? mi : mo) * w.Cin * w.Cout * taps, 2.0 * (mi * w.Cin + mo * w.Cout)});
        }
        hh->ops.push_back([=](Run& r) {
            ConvParams p;
            p.gin = hh->geom[level_in];
            p.gout = hh->geom[level_out];
            p.nsrc = (int)srcs.size();
            p.Cin_total = w.Cin;
            for (int i = 0; i < p.nsrc; ++i) { p.src[i].base = hh->bufs[srcs[i]].ptr; p.src[i].C = hh->bufs[srcs[i]].C; }
            if (p.nsrc == 1) { p.src[1].base = nullptr; p.src[1].C = 0; }
            p.dst = hh->bufs[dst].ptr;
            p.Cout = w.Cout;
            p.w = w.w; p.bias = w.bias;
            p.kz = w.kz; p.kp = w.kp; p.kr = w.kr;
            p.sz = p.sp = p.sr = stride;
            p.pz = pz; p.cp = cp; p.pr = pr;
            p.transposed = w.transposed;
            p.write_halo = write_halo;
            p.stats = stats_off >= 0 ? (double*)(hh->scratch + (size_t)stats_off * r.B) : nullptr;
            p.stats_cpg = w.Cout / (hh->d.groups > 0 ? hh->d.groups : 1);
            p.pn_stats = nullptr; p.pn_wsum = nullptr; p.pn_wb = nullptr; p.pn_C = 0; p.pn_count = 1.0;
            p.B = r.B;
            hh->launches += 1;
            // (measured: DS2 L1 at B = 1000, 1.2 M rows, 0.287 -> 0.263 ms; neutral at 0.3 M rows; slower on the 30 k rows of
            //  the HGCal L2 level at B = 128, where the second launch's set-up is not amortised)
            if (hh->conv_backend == 0 && w.wmma_lo && w.wmma_hi && stride == 1 && !w.transposed && p.nsrc == 1 &&
                (long long)r.B * p.gout.rows_per_sample >= 500000) {
                ConvParams lo = p, hi = p;
                const int half = w.Cout / 2;
                lo.Cout = hi.Cout = half; lo.stats_C = hi.stats_C = w.Cout;
                hi.dst = p.dst + (long long)(half / 8) * p.gout.chunk_stride * 8;
                hi.bias = w.bias_hi;
                hi.stats = p.stats ? p.stats + (size_t)half * 2 : nullptr;
                if (launch_conv_mma(lo, w.wmma_lo, r.st) == 0) {
                    if (launch_conv_mma(hi, w.wmma_hi, r.st) == 0) { hh->launches += 1; return; }
                    // (the first half is overwritten by whichever full-width form runs below)
                }
            }
            if (fg.raw >= 0) {
                const Geom& g = p.gout;
                const double cnt = (double)g.Z * g.P * g.R;
                const double* gst = (const double*)(hh->scratch + (size_t)fg.stats_off * r.B);
                const float* av = fg.addvec_off >= 0 ? hh->addvec + fg.addvec_off : nullptr;
                if (hh->conv_backend == 0 && w.wz) {
                    ConvParams pf = p;
                    pf.src[0].base = hh->bufs[fg.raw].ptr; pf.src[0].C = hh->bufs[fg.raw].C;
                    pf.gn_stats = gst; pf.gn_gamma = fg.gamma; pf.gn_beta = fg.beta; pf.gn_groups = fg.groups; pf.gn_count = cnt;
                    pf.gn_addvec = av; pf.gn_addvec_stride = hh->addvec_stride;
                    if (launch_conv_zstack(pf, w.wz, zs_force, r.st) == 0) return;
                }
                // not taken (small batch, other backend): the stand-alone GroupNorm pass, then the convolution on its result
                launch_gn_apply(g, hh->bufs[fg.raw].ptr, gst, fg.gamma, fg.beta, fg.groups, cnt, av, hh->addvec_stride, nullptr, nullptr, 0,
                                hh->bufs[srcs[0]].ptr, w.Cin, 1, nullptr, r.B, r.st);
                hh->launches += 1;
            }
            if (hh->conv_backend == 0 && w.wz && stride == 1 && !w.transposed) {
                if (launch_conv_zstack(p, w.wz, zs_force, r.st) == 0) return;
            }
            if (hh->conv_backend == 0 && w.wmma && stride == 1 && !w.transposed) {
                if (launch_conv_mma(p, w.wmma, r.st) == 0) return;
            }
            if (hh->conv_backend == 0 && w.wgat) {
                if (launch_conv_gather(p, w.wgat, r.st) == 0) return;
            }
            launch_conv_generic(p, r.st);
        });
    }

    // stride-1 3x3x3 conv over two concatenated sources as two tcgen05 convs (see resblock); falls back to the single conv
    void conv_split(const ConvW& full, const ConvW& wa, const ConvW& wb, int level, std::vector<int> srcs, int dst, long long stats_off) {
        cd_unet* hh = h;
        {
            const Geom& g = hh->geom[level];
            double m = (double)g.Z * g.P * g.R;
            char nm[96];
            snprintf(nm, sizeof(nm), "conv_k27_c%d+%d_%d_L%d_mma", wa.Cin, wb.Cin, full.Cout, level);
            hh->op_info.push_back({nm, 2.0 * m * full.Cin * full.Cout * 27, 2.0 * m * (full.Cin + full.Cout)});
        }
        hh->ops.push_back([=](Run& r) {
            ConvParams p;
            p.gin = hh->geom[level]; p.gout = hh->geom[level];
            p.dst = hh->bufs[dst].ptr; p.Cout = full.Cout;
            p.kz = p.kp = p.kr = 3; p.sz = p.sp = p.sr = 1; p.pz = p.cp = p.pr = 1;
            p.transposed = 0; p.write_halo = 0;
            p.stats_cpg = full.Cout / (hh->d.groups > 0 ? hh->d.groups : 1);
            p.pn_stats = nullptr; p.pn_wsum = nullptr; p.pn_wb = nullptr; p.pn_C = 0; p.pn_count = 1.0;
            p.B = r.B;
            double* stats = stats_off >= 0 ? (double*)(hh->scratch + (size_t)stats_off * r.B) : nullptr;
            if (hh->conv_backend == 0) {
                ConvParams pa = p;
                pa.nsrc = 1; pa.Cin_total = wa.Cin; pa.src[0].base = hh->bufs[srcs[0]].ptr; pa.src[0].C = wa.Cin;
                pa.src[1].base = nullptr; pa.src[1].C = 0; pa.w = wa.w; pa.bias = wa.bias; pa.stats = nullptr; pa.addsrc = nullptr;
                ConvParams pb = pa;
                pb.Cin_total = wb.Cin; pb.src[0].base = hh->bufs[srcs[1]].ptr; pb.src[0].C = wb.Cin; pb.w = wb.w; pb.bias = nullptr;
                pb.stats = stats; pb.addsrc = p.dst;
                if (launch_conv_mma(pa, wa.wmma, r.st) == 0 && launch_conv_mma(pb, wb.wmma, r.st) == 0) { hh->launches += 2; return; }
            }
            p.nsrc = 2; p.Cin_total = full.Cin;
            for (int i = 0; i < 2; ++i) { p.src[i].base = hh->bufs[srcs[i]].ptr; p.src[i].C = hh->bufs[srcs[i]].C; }
            p.w = full.w; p.bias = full.bias; p.stats = stats; p.addsrc = nullptr;
            hh->launches += 1;
            if (hh->conv_backend == 0 && full.wmma && launch_conv_mma(p, full.wmma, r.st) == 0) return;
            launch_conv_generic(p, r.st);
        });
    }

    // GroupNorm finalize + apply:  dst = act(GN(src)) + addvec + residual   (all rows incl. halos)
    // out_stats_off >= 0: also accumulate the per-sample sum / sum of squares of dst (GroupNorm(1,C) of a following PreNorm)
    void gn(const std::string& gname, const std::string& bname, int level, int src, int dst, int C, int groups, long long stats_off,
            int act, int addvec_off /* -1 none */, int residual /* -1 none */, int residual2 = -1, long long out_stats_off = -1) {
        cd_unet* hh = h;
        float* gamma = upload_vec(hh, gname, C);
        float* beta = upload_vec(hh, bname, C);
        if (!gamma || !beta) { ok = false; return; }
        if (C > 256) { fail("GroupNorm over more than 256 channels is not supported"); ok = false; return; }
        if (hh->train) {
            TapeOp t;
            t.kind = 1; t.gname = gname; t.betaname = bname; t.level_in = t.level_out = level; t.src = src; t.dst = dst; t.C = C;
            t.groups = groups; t.stats_off = (size_t)stats_off; t.act = act; t.addvec_off = addvec_off; t.residual = residual;
            t.residual2 = residual2;
            for (int i = (int)hh->tape.size() - 1; i >= 0; --i)      // the convolution that produced `src` (its bias gradient comes from here)
                if (hh->tape[i].kind == 0 && hh->tape[i].dst == src) { t.conv_of_src = i; hh->tape[i].bias_via_gn = true; break; }
            hh->tape.push_back(t);
        }
        {
            const Geom& g = hh->geom[level];
            double m = (double)g.Z * g.P * g.R;
            char nm[96];
            snprintf(nm, sizeof(nm), "gn_apply_c%d_L%d%s", C, level, residual >= 0 ? "_res" : "");
            hh->op_info.push_back({nm, 0.0, 2.0 * m * C * (residual >= 0 ? 3 : 2)});
        }
        hh->ops.push_back([=](Run& r) {
            const Geom& g = hh->geom[level];
            double cnt = (double)g.Z * g.P * g.R;
            launch_gn_apply(g, hh->bufs[src].ptr, (const double*)(hh->scratch + (size_t)stats_off * r.B), gamma, beta, groups, cnt,
                            addvec_off >= 0 ? hh->addvec + addvec_off : nullptr, hh->addvec_stride,
                            residual >= 0 ? hh->bufs[residual].ptr : nullptr, residual2 >= 0 ? hh->bufs[residual2].ptr : nullptr,
                            residual >= 0 ? hh->bufs[residual].C / 8 : 0, hh->bufs[dst].ptr, C, act,
                            out_stats_off >= 0 ? (double*)(hh->scratch + (size_t)out_stats_off * r.B) : nullptr, r.B, r.st);
            hh->launches += 1;
        });
    }

    // ResnetBlock (models.py:172-200).  srcs at `level`; returns output buffer (caller releases inputs)
    int resblock(const std::string& pre, int level, std::vector<int> srcs, int Cin, int Cout, bool has_cond, long long out_stats_off = -1) {
        int groups = h->d.groups;
        ConvW c1, c2, cres;
        if (!pack_conv(h, pre + "block1.proj.conv.weight", pre + "block1.proj.conv.bias", Cin, Cin, Cout, 3, 3, 3, 0, true, &c1) ||
            !pack_conv(h, pre + "block2.proj.conv.weight", pre + "block2.proj.conv.bias", Cout, Cout, Cout, 3, 3, 3, 0, true, &c2)) {
            ok = false; return -1;
        }
        bool has_res = find_w(h, pre + "res_conv.conv.weight") != nullptr;
        if (!has_res && Cin != Cout) { fail("missing weight %sres_conv.conv.weight", pre.c_str()); ok = false; return -1; }
        if (has_res && !pack_conv(h, pre + "res_conv.conv.weight", pre + "res_conv.conv.bias", Cin, Cin, Cout, 1, 1, 1, 0, true, &cres)) {
            ok = false; return -1;
        }
        int addvec_off = -1;
        if (has_cond) {
            int k = h->embed.n_proj;
            if (k >= 16) { fail("too many conditioned blocks"); ok = false; return -1; }
            h->embed.pw[k] = upload_vec(h, pre + "mlp.1.weight", Cout * h->d.cond_dim);
            h->embed.pb[k] = upload_vec(h, pre + "mlp.1.bias", Cout);
            if (!h->embed.pw[k] || !h->embed.pb[k]) { ok = false; return -1; }
            h->embed.pC[k] = Cout;
            h->proj_names.push_back(pre + "mlp.1");
            h->embed.poff[k] = h->addvec_stride;
            addvec_off = h->addvec_stride;
            h->addvec_stride += Cout;
            h->embed.n_proj = k + 1;
        }
        size_t s1 = slot((size_t)Cout * 16), s2 = slot((size_t)Cout * 16);
        int t1 = acquire(h, level, Cout);
        // A 3x3x3 conv over a concatenated input whose weights cannot stay resident in shared memory (27*Cin*Cout fp16)
        // runs as two resident-weight convs, one per source, the second accumulating onto the first (conv is linear in Cin).
        ConvW ca, cb2;
        bool split = false;
        // (criterion: resident weights + two activation windows of the full input width exceed what one CTA can hold, so the
        //  single conv would stream its weights or stage K in many small groups; measured on DS2 L1 128->32 and DS3 L0 64->32)
        // Round 2 re-measured the 64 -> 32 case after the ring bookkeeping rework: the single K-staged conv (110 KB of resident
        // weights, K in four groups) is now the faster form (DS3 L0 1.39 vs 1.46-1.51 ms, HGCal L0 0.153 vs 0.176 ms), so the
        // split is kept for inputs whose weights cannot stay resident at all (27*Cin*Cout fp16 > 160 KB: 128 -> 32 and wider).
        if (srcs.size() == 2 && (size_t)27 * Cin * Cout * 2 > (size_t)160 * 1024 && c1.wmma && !getenv("CD_NO_SPLIT") && !h->train) {
            const int Ca = h->bufs[srcs[0]].C, Cb = h->bufs[srcs[1]].C;
            const HostTensor* wfull = find_w(h, pre + "block1.proj.conv.weight");
            if (wfull && Ca + Cb == Cin && mma_conv_supported(Ca, Cout, 27) && mma_conv_supported(Cb, Cout, 27)) {
                HostTensor wa, wb;
                wa.shape = {Cout, Ca, 3, 3, 3}; wb.shape = {Cout, Cb, 3, 3, 3};
                wa.data.resize((size_t)Cout * Ca * 27); wb.data.resize((size_t)Cout * Cb * 27);
                for (int co = 0; co < Cout; ++co) {
                    std::copy_n(&wfull->data[((size_t)co * Cin) * 27], (size_t)Ca * 27, &wa.data[(size_t)co * Ca * 27]);
                    std::copy_n(&wfull->data[((size_t)co * Cin + Ca) * 27], (size_t)Cb * 27, &wb.data[(size_t)co * Cb * 27]);
                }
                h->weights["__split_a"] = wa; h->weights["__split_b"] = wb;
                split = pack_conv(h, "__split_a", pre + "block1.proj.conv.bias", Ca, Ca, Cout, 3, 3, 3, 0, true, &ca) &&
                        pack_conv(h, "__split_b", "", Cb, Cb, Cout, 3, 3, 3, 0, true, &cb2) && ca.wmma && cb2.wmma;
                h->weights.erase("__split_a"); h->weights.erase("__split_b");
            }
        }
        // block1's GroupNorm + SiLU + conditioning has one consumer, block2's conv: where that conv can run Z-stacked (sampling
        // handles, C = 32, a plane that fills its strips) the normalisation is applied on the conv's staged input instead of by a
        // pass of its own; block1's conv then writes its phi halo rows itself
        static const bool zs_fuse_off = getenv("CD_NO_GN_FUSE") != nullptr, zs_forced = getenv("CD_ZS_FORCE") != nullptr;
        const bool fuse = !h->train && !split && !zs_fuse_off && c2.wz && Cout == 32 && (zs_forced || zstack_geom_efficient(h->geom[level]));
        tag(pre + "block1.proj.conv.weight", pre + "block1.proj.conv.bias", Cin);
        if (split) conv_split(c1, ca, cb2, level, srcs, t1, (long long)s1);
        else
        conv(c1, level, level, srcs, t1, 1, 1, 1, 1, fuse ? 1 : 0, (long long)s1);
        int hbuf = acquire(h, level, Cout);
        FusedGn fg;
        if (fuse) {
            fg.raw = t1; fg.groups = groups; fg.stats_off = (long long)s1; fg.addvec_off = addvec_off;
            fg.gamma = upload_vec(h, pre + "block1.norm.weight", Cout);
            fg.beta = upload_vec(h, pre + "block1.norm.bias", Cout);
            if (!fg.gamma || !fg.beta) { ok = false; return -1; }
        } else {
            gn(pre + "block1.norm.weight", pre + "block1.norm.bias", level, t1, hbuf, Cout, groups, (long long)s1, 1, addvec_off, -1);
        }
        // (fused: block2's conv reads t1, so its output needs a buffer of its own - hbuf, which then also holds the block's result)
        int t2 = h->train ? acquire(h, level, Cout) : (fuse ? acquire(h, level, Cout) : t1);     // training keeps block1's pre-norm tensor
        tag(pre + "block2.proj.conv.weight", pre + "block2.proj.conv.bias", Cout);
        conv(c2, level, level, {hbuf}, t2, 1, 1, 1, 1, 0, (long long)s2, fg);
        int res = srcs[0];
        int rbuf = -1;
        if (has_res) {
            rbuf = acquire(h, level, Cout);
            tag(pre + "res_conv.conv.weight", pre + "res_conv.conv.bias", Cin);
            conv(cres, level, level, srcs, rbuf, 1, 0, 0, 0, 0, -1);
            res = rbuf;
        }
        // hbuf is free again as the destination: gn_apply reads t1 + residual only
        int res2 = (!has_res && srcs.size() == 2) ? srcs[1] : -1;  // identity residual over a concatenated input
        int obuf = h->train ? acquire(h, level, Cout) : hbuf;     // training keeps block2's input (hbuf)
        gn(pre + "block2.norm.weight", pre + "block2.norm.bias", level, t2, obuf, Cout, groups, (long long)s2, 1, -1, res, res2, out_stats_off);
        release(h, t1);
        if (rbuf >= 0) release(h, rbuf);
        return obuf;
    }

    // Residual(PreNorm(LinearAttention)) (models.py:111-117, 281-329); returns output buffer.
    // s0: scratch slot holding the GroupNorm(1,C) statistics of x (accumulated by the gn() that produced x)
    int attention(const std::string& pre, int level, int x, int C, size_t s0) {
        cd_unet* hh = h;
        ConvW cq;
        if (!pack_conv(h, pre + "fn.fn.to_qkv.conv.weight", "", C, C, 96, 1, 1, 1, 0, true, &cq)) { ok = false; return -1; }
        if (!attention_out_supported(C)) { fail("attention over %d channels is not supported", C); ok = false; return -1; }
        const HostTensor* wo = find_w(h, pre + "fn.fn.to_out.0.conv.weight");
        if (!wo || wo->shape.size() != 5 || wo->shape[0] != C || wo->shape[1] != 32) {
            fail("weight %sfn.fn.to_out.0.conv.weight missing or of unexpected shape", pre.c_str()); ok = false; return -1;
        }
        float* wout = upload_vec(h, pre + "fn.fn.to_out.0.conv.weight", C * 32);
        float* bout = upload_vec(h, pre + "fn.fn.to_out.0.conv.bias", C);
        if (!wout || !bout) { ok = false; return -1; }
        size_t s1 = slot((size_t)C * 16);
        size_t skm = slot(32 * 4), sctx = slot(33 * 32 * 4);
        size_t sfold = slot((size_t)C * 32 * 2);     // fused attention: per-sample folded (Ctx/S, to_out) matrix, fp16
        int xn = acquire(h, level, C);
        int qkv = acquire(h, level, 96);
        if (h->train) {
            // training: the unfused chain, one tape entry per step - PreNorm GroupNorm(1,C) -> to_qkv 1x1x1 conv -> attention core
            // (+ to_out conv) -> GroupNorm(1,C) + residual
            gn(pre + "fn.norm.weight", pre + "fn.norm.bias", level, x, xn, C, 1, (long long)s0, 0, -1, -1);
            tag(pre + "fn.fn.to_qkv.conv.weight", "", C);
            conv(cq, level, level, {xn}, qkv, 1, 0, 0, 0, 0, -1);
            int y = acquire(h, level, C);
            {
                const Geom& g = hh->geom[level];
                double m = (double)g.Z * g.P * g.R;
                char nm[96];
                snprintf(nm, sizeof(nm), "attention_L%d", level);
                hh->op_info.push_back({nm, 2.0 * 2.0 * 32 * 32 * m + 2.0 * 32 * C * m, 2.0 * m * (32 * 4 + C)});
                TapeOp t;
                t.kind = 2; t.level_in = t.level_out = level; t.qkv = qkv; t.y = y; t.C = C; t.skm = skm; t.sctx = sctx;
                t.woname = pre + "fn.fn.to_out.0.conv.weight"; t.boname = pre + "fn.fn.to_out.0.conv.bias";
                hh->tape.push_back(t);
            }
            hh->ops.push_back([=](Run& r) {
                launch_attention(hh->geom[level], hh->bufs[qkv].ptr, (unsigned*)(hh->scratch + skm * r.B),
                                 (float*)(hh->scratch + sctx * r.B), wout, bout, hh->bufs[y].ptr, C,
                                 (double*)(hh->scratch + s1 * r.B), r.B, r.st);
                hh->launches += 3;
            });
            int out = acquire(h, level, C);
            gn(pre + "fn.fn.to_out.1.weight", pre + "fn.fn.to_out.1.bias", level, y, out, C, 1, (long long)s1, 0, -1, x);
            return out;
        }
        // PreNorm (GroupNorm(1,C), models.py:321-329) folded into the to_qkv 1x1x1 conv on the tcgen05 path:
        //   W (gamma*(x-mean)*rstd + beta) = rstd * (W diag(gamma)) x - mean*rstd * rowsum(W diag(gamma)) + W beta
        // so the conv reads x directly (no normalised copy) and its epilogue applies the per-sample scalars.
        ConvW cqf;
        float *pn_wsum = nullptr, *pn_wb = nullptr;
        __half* wf_h = nullptr;
        {
            const HostTensor* wq = find_w(h, pre + "fn.fn.to_qkv.conv.weight");
            const HostTensor* gm = find_w(h, pre + "fn.norm.weight");
            const HostTensor* bt = find_w(h, pre + "fn.norm.bias");
            if (!wq || !gm || !bt || (int)gm->data.size() != C || (int)bt->data.size() != C) { fail("missing PreNorm weights %s", pre.c_str()); ok = false; return -1; }
            HostTensor folded = *wq;
            std::vector<float> wsum(96, 0.f), wb(96, 0.f);
            for (int o = 0; o < 96; ++o)
                for (int ci = 0; ci < C; ++ci) {
                    const float wf = wq->data[(size_t)o * C + ci] * gm->data[ci];
                    folded.data[(size_t)o * C + ci] = wf;
                    wsum[o] += __half2float(__float2half_rn(wf));     // of the weights as the tensor cores see them
                    wb[o] += wq->data[(size_t)o * C + ci] * bt->data[ci];
                }
            h->weights["__folded_qkv"] = folded;
            bool okp = pack_conv(h, "__folded_qkv", "", C, C, 96, 1, 1, 1, 0, true, &cqf);
            h->weights.erase("__folded_qkv");
            if (!okp) { ok = false; return -1; }
            pn_wsum = h->upload(wsum);
            pn_wb = h->upload(wb);
            if (attention_fused_supported(C)) {     // row-major fp16 [96][C] for the fused attention kernels
                std::vector<__half> wh((size_t)96 * C);
                for (size_t i = 0; i < wh.size(); ++i) wh[i] = __float2half_rn(folded.data[i]);
                wf_h = h->dalloc<__half>(wh.size(), false);
                if (!wf_h) { ok = false; return -1; }
                cudaMemcpy(wf_h, wh.data(), wh.size() * sizeof(__half), cudaMemcpyHostToDevice);
            }
        }
        float* pn_gamma = upload_vec(h, pre + "fn.norm.weight", C);
        float* pn_beta = upload_vec(h, pre + "fn.norm.bias", C);
        if (!pn_gamma || !pn_beta) { ok = false; return -1; }
        {
            const Geom& g = hh->geom[level];
            double m = (double)g.Z * g.P * g.R;
            char nm[96];
            snprintf(nm, sizeof(nm), "conv_k1_prenorm_c%d_96_L%d_mma", C, level);
            hh->op_info.push_back({nm, 2.0 * m * C * 96, 2.0 * m * (C + 96)});
        }
        // fused form (cd_attention_fused.cu): q/k/v are recomputed from x inside the attention passes, no qkv tensor
        const bool can_fuse = wf_h != nullptr;
        hh->ops.push_back([=](Run& r) {
            if (can_fuse && hh->conv_backend == 0 && !hh->no_fused_attention) return;
            const Geom& g = hh->geom[level];
            ConvParams p;
            p.gin = g; p.gout = g; p.nsrc = 1; p.Cin_total = C;
            p.src[0].C = C; p.src[1].base = nullptr; p.src[1].C = 0;
            p.dst = hh->bufs[qkv].ptr; p.Cout = 96; p.bias = nullptr;
            p.kz = p.kp = p.kr = 1; p.sz = p.sp = p.sr = 1; p.pz = p.cp = p.pr = 0;
            p.transposed = 0; p.write_halo = 0; p.stats = nullptr; p.stats_cpg = 1; p.B = r.B;
            p.pn_stats = nullptr; p.pn_wsum = nullptr; p.pn_wb = nullptr; p.pn_C = 0; p.pn_count = 1.0;
            const double* st = (const double*)(hh->scratch + s0 * r.B);
            if (hh->conv_backend == 0 && cqf.wmma) {
                p.src[0].base = hh->bufs[x].ptr; p.w = cqf.w;
                p.pn_stats = st; p.pn_wsum = pn_wsum; p.pn_wb = pn_wb; p.pn_C = C; p.pn_count = (double)C * g.Z * g.P * g.R;
                hh->launches += 1;
                if (launch_conv_mma(p, cqf.wmma, r.st) == 0) return;
                p.pn_stats = nullptr;
            }
            // unfused form: normalised copy, then the plain conv (SIMT backend / unsupported widths)
            launch_gn_apply(g, hh->bufs[x].ptr, st, pn_gamma, pn_beta, 1, (double)g.Z * g.P * g.R, nullptr, hh->addvec_stride, nullptr,
                            nullptr, 0, hh->bufs[xn].ptr, C, 0, nullptr, r.B, r.st);
            p.src[0].base = hh->bufs[xn].ptr; p.w = cq.w;
            hh->launches += 2;
            if (hh->conv_backend == 0 && cq.wmma && launch_conv_mma(p, cq.wmma, r.st) == 0) return;
            launch_conv_generic(p, r.st);
        });
        {
            const Geom& g = hh->geom[level];
            double m = (double)g.Z * g.P * g.R;
            char nm[96];
            snprintf(nm, sizeof(nm), "attention_L%d", level);
            // two 32x32xn einsums + the to_out 1x1x1 conv (+ the three projections the fused form recomputes).  Bytes: the fused
            // form moves the C-channel tensor three times (x read by the context pass and by the output pass, y written) =
            // 192 B/voxel at C = 32; the unfused form reads k twice, v and q once from the 96-channel tensor and writes C channels
            const bool fused_form = can_fuse && !hh->no_fused_attention;
            hh->op_info.push_back({nm, 2.0 * 2.0 * 32 * 32 * m + 2.0 * 32 * C * m + (fused_form ? 2.0 * (64 + 32) * C * m : 0.0),
                                   fused_form ? 2.0 * m * 3 * C : 2.0 * m * (32 * 4 + C)});
        }
        // attention + to_out conv -> xn reused as the raw output (+ its GroupNorm(1,C) statistics), then GN(1) + residual x
        hh->ops.push_back([=](Run& r) {
            if (can_fuse && hh->conv_backend == 0 && !hh->no_fused_attention) {
                const Geom& g = hh->geom[level];
                launch_attention_fused(g, hh->bufs[x].ptr, wf_h, pn_wsum, pn_wb, (const double*)(hh->scratch + s0 * r.B),
                                       (double)C * g.Z * g.P * g.R, (unsigned*)(hh->scratch + skm * r.B),
                                       (float*)(hh->scratch + sctx * r.B), hh->online_attention ? hh->attn_part : nullptr,
                                       (__half*)(hh->scratch + sfold * r.B), wout, bout,
                                       hh->bufs[xn].ptr, C, (double*)(hh->scratch + s1 * r.B), r.B, r.st);
                hh->launches += 4;
                return;
            }
            launch_attention(hh->geom[level], hh->bufs[qkv].ptr, (unsigned*)(hh->scratch + skm * r.B),
                             (float*)(hh->scratch + sctx * r.B), wout, bout, hh->bufs[xn].ptr, C,
                             (double*)(hh->scratch + s1 * r.B), r.B, r.st);
            hh->launches += 3;
        });
        int out = acquire(h, level, C);
        gn(pre + "fn.fn.to_out.1.weight", pre + "fn.fn.to_out.1.bias", level, xn, out, C, 1, (long long)s1, 0, -1, x);
        release(h, xn); release(h, qkv);
        return out;
    }
};

// [out][in] row-major fp32 -> device [in][out]
static const float* upload_transposed(cd_unet* h, const std::string& name, int n_out, int n_in) {
    const HostTensor* t = find_w(h, name);
    if (!t) { fail("missing weight %s", name.c_str()); return nullptr; }
    if ((int)t->data.size() != n_out * n_in) { fail("weight %s has unexpected size %zu (want %dx%d)", name.c_str(), t->data.size(), n_out, n_in); return nullptr; }
    std::vector<float> w((size_t)n_in * n_out);
    for (int o = 0; o < n_out; ++o)
        for (int i = 0; i < n_in; ++i) w[(size_t)i * n_out + o] = t->data[(size_t)o * n_in + i];
    return h->upload(w);
}

// dense ResNet of the layer-energy model (models/models.py:390-457): one kernel per evaluation
static int build_dense_plan(cd_unet* h) {
    const cd_layer_desc& d = h->ld;
    const std::string M = "model.";
    LayerMlpParams& m = h->lm;
    const int half = d.cond_dim / 2;
    m.dim_in = d.dim_in; m.hidden = d.hidden; m.n_hidden = d.num_layers - 1; m.cond_size = d.cond_size; m.cond_dim = d.cond_dim;
    m.time_mode = d.time_embed;
    const int tidx[3] = {1, 3, 5}, cidx[3] = {0, 2, 4};
    const int tin[3] = {1, half / 2, half}, cin[3] = {d.cond_size, half / 2, half}, outs[3] = {half / 2, half, half};
    for (int i = 0; i < 3; ++i) {
        m.tw[i] = upload_transposed(h, M + "time_mlp." + std::to_string(tidx[i]) + ".weight", outs[i], tin[i]);
        m.tb[i] = upload_vec(h, M + "time_mlp." + std::to_string(tidx[i]) + ".bias", outs[i]);
        m.cw[i] = upload_transposed(h, M + "cond_mlp." + std::to_string(cidx[i]) + ".weight", outs[i], cin[i]);
        m.cb[i] = upload_vec(h, M + "cond_mlp." + std::to_string(cidx[i]) + ".bias", outs[i]);
        if (!m.tw[i] || !m.tb[i] || !m.cw[i] || !m.cb[i]) return 1;
    }
    m.in_w = upload_transposed(h, M + "in_lay.weight", d.hidden, d.dim_in);
    m.in_b = upload_vec(h, M + "in_lay.bias", d.hidden);
    m.out_w = upload_transposed(h, M + "out_lay.weight", d.dim_in, d.hidden);
    m.out_b = upload_vec(h, M + "out_lay.bias", d.dim_in);
    if (!m.in_w || !m.in_b || !m.out_w || !m.out_b) return 1;
    for (int l = 0; l < m.n_hidden; ++l) {
        const std::string L = M + "hidden_layers." + std::to_string(l) + ".";
        m.ew[l] = upload_transposed(h, L + "embeder.1.weight", d.hidden, d.cond_dim);
        m.eb[l] = upload_vec(h, L + "embeder.1.bias", d.hidden);
        m.w1[l] = upload_transposed(h, L + "dense1.0.weight", d.hidden, d.hidden);
        m.b1[l] = upload_vec(h, L + "dense1.0.bias", d.hidden);
        m.w2[l] = upload_transposed(h, L + "dense2.0.weight", d.hidden, d.hidden);
        m.b2[l] = upload_vec(h, L + "dense2.0.bias", d.hidden);
        if (!m.ew[l] || !m.eb[l] || !m.w1[l] || !m.b1[l] || !m.w2[l] || !m.b2[l]) return 1;
    }
    const double mac = (double)d.dim_in * d.hidden * 2 + (double)m.n_hidden * ((double)d.cond_dim * d.hidden + 2.0 * d.hidden * d.hidden);
    h->op_info.push_back({"layer_mlp", 2.0 * mac, 8.0 * d.dim_in});
    h->ops.push_back([h](Run& r) {
        FinalParams f = r.fin;
        f.x = r.x; f.sigma = r.sigma; f.sigma_stride = r.sigma_stride;
        f.sigma_data = r.scale_input ? h->ld.sigma_data : -1.f;
        launch_layer_mlp(h->lm, f, r.cond, r.B, r.time_is_raw, r.st);
        h->launches += 1;
    });
    h->scratch_per_sample = 0;
    return 0;
}

static int build_plan(cd_unet* h) {
    if (h->dense) return build_dense_plan(h);
    const cd_unet_desc& d = h->d;
    Builder b;
    b.h = h;
    const std::string M = "model.";
    int nl = h->n_levels;
    const int* ls = d.layer_sizes;

    if (h->train) {
        // flat fp32 master copy of every parameter (PyTorch layouts, name order), its gradient and the Adam moments
        long long off = 0;
        for (auto& kv : h->weights) {
            h->param_index[kv.first] = (int)h->params.size();
            h->params.push_back({kv.first, off, (long long)kv.second.data.size()});
            off += ((long long)kv.second.data.size() + 3) / 4 * 4;      // 16-byte aligned slots
        }
        h->n_params = off;
        h->P = h->dalloc<float>((size_t)off, true); h->G = h->dalloc<float>((size_t)off, true);
        h->M = h->dalloc<float>((size_t)off, true); h->V = h->dalloc<float>((size_t)off, true);
        if (!h->P || !h->G || !h->M || !h->V) return fail("out of device memory for the training parameters");
        for (auto& ps : h->params)
            cudaMemcpy(h->P + ps.offset, h->weights[ps.name].data.data(), (size_t)ps.numel * sizeof(float), cudaMemcpyHostToDevice);
    }
    // embedding MLPs
    EmbedParams& e = h->embed;
    memset(&e, 0, sizeof(e));
    int half = d.cond_dim / 2;
    int hid = d.cond_size > half / 2 ? d.cond_size : half / 2;
    e.cond_size = d.cond_size; e.cond_hidden = hid; e.cond_dim = d.cond_dim; e.time_mode = d.time_embed;
    if (d.cond_dim > 256 || hid > 256 || d.cond_size > 256) return fail("conditioning sizes above 256 are not supported");
    const int tidx[3] = {1, 3, 5}, cidx[3] = {0, 2, 4};
    const int tin[3] = {1, half / 2, half}, tout[3] = {half / 2, half, half};
    const int cin[3] = {d.cond_size, hid, half}, cout[3] = {hid, half, half};
    for (int i = 0; i < 3; ++i) {
        e.tw[i] = upload_vec(h, M + "time_mlp." + std::to_string(tidx[i]) + ".weight", tin[i] * tout[i]);
        e.tb[i] = upload_vec(h, M + "time_mlp." + std::to_string(tidx[i]) + ".bias", tout[i]);
        e.cw[i] = upload_vec(h, M + "cond_mlp." + std::to_string(cidx[i]) + ".weight", cin[i] * cout[i]);
        e.cb[i] = upload_vec(h, M + "cond_mlp." + std::to_string(cidx[i]) + ".bias", cout[i]);
        if (!e.tw[i] || !e.tb[i] || !e.cw[i] || !e.cb[i]) return 1;
    }
    h->addvec_stride = 0;
    h->op_info.push_back({"embed_mlp", 0.0, 0.0});
    h->ops.push_back([h](Run& r) {
        EmbedParams ep = h->embed;
        if (r.time_is_raw) ep.time_mode = 2;
        launch_embed(ep, r.sigma, r.sigma_stride, r.cond, h->addvec, r.B, r.st);
        h->launches += 1;
    });

    // input + init conv (geometry channels are generated in-kernel; Cin zero-padded to the tensor-core K granularity 16)
    int tin_buf = acquire(h, 0, 16);
    {
        double m = (double)d.Z * d.P * d.R;
        h->op_info.push_back({"prep_input", 0.0, m * 4 + m * 16});
    }
    h->ops.push_back([h, tin_buf](Run& r) {
        if (h->has_geom)
            launch_prep_input_geom(h->geom[0], h->geom_map[0], r.x, r.sigma, r.sigma_stride, r.scale_input ? h->d.sigma_data : -1.f,
                                   h->r_vals, h->d.rz_input, h->d.phi_input, h->bufs[tin_buf].ptr, 2, r.B, r.st);
        else
            launch_prep_input(h->geom[0], r.x, r.sigma, r.sigma_stride, r.scale_input ? h->d.sigma_data : -1.f, h->r_vals,
                              h->d.rz_input, h->d.phi_input, h->bufs[tin_buf].ptr, 2, r.B, r.st);
        h->launches += 1;
    });
    ConvW cinit;
    if (!pack_conv(h, M + "init_conv.conv.weight", M + "init_conv.conv.bias", d.in_channels, 16, ls[0], 3, 3, 3, 0, true, &cinit, true)) return 1;
    int x = acquire(h, 0, ls[0]);
    b.tag(M + "init_conv.conv.weight", M + "init_conv.conv.bias", d.in_channels);
    b.conv(cinit, 0, 0, {tin_buf}, x, 1, 1, 1, 1, 1, -1);
    if (h->train) { h->tape.back().need_dgrad = false; h->tin_buf_index = tin_buf; }
    release(h, tin_buf);

    std::vector<int> skips;
    for (int i = 0; i < nl; ++i) {
        std::string p = M + "downs." + std::to_string(i) + ".";
        int cin_ = ls[i], cout_ = ls[i + 1];
        int y = b.resblock(p + "0.", i, {x}, cin_, cout_, true);
        if (!b.ok) return 1;
        release(h, x);
        size_t sa = b.slot((size_t)cout_ * 16);
        int y2 = b.resblock(p + "1.", i, {y}, cout_, cout_, true, (long long)sa);
        if (!b.ok) return 1;
        release(h, y);
        int a = b.attention(M + "downs_attn." + std::to_string(i) + ".", i, y2, cout_, sa);
        if (!b.ok) return 1;
        release(h, y2);
        skips.push_back(a);
        if (i < nl - 1) {
            ConvW cd_;
            if (!pack_conv(h, p + "2.conv.weight", p + "2.conv.bias", cout_, cout_, cout_, 3, 4, 4, 0, false, &cd_, true)) return 1;
            x = acquire(h, i + 1, cout_);
            b.tag(p + "2.conv.weight", p + "2.conv.bias", cout_);
            b.conv(cd_, i, i + 1, {a}, x, 2, 1, 1, 1, 1, -1);
        } else {
            x = a;
        }
    }
    int mid = ls[nl];
    int lv = nl - 1;
    {
        size_t sa = b.slot((size_t)mid * 16);
        int y = b.resblock(M + "mid_block1.", lv, {x}, mid, mid, true, (long long)sa);
        if (!b.ok) return 1;
        // x is the last skip: keep it
        int a = b.attention(M + "mid_attn.", lv, y, mid, sa);
        if (!b.ok) return 1;
        release(h, y);
        int y2 = b.resblock(M + "mid_block2.", lv, {a}, mid, mid, true);
        if (!b.ok) return 1;
        release(h, a);
        x = y2;
    }
    for (int i = 0; i < nl; ++i) {
        int lev = nl - 1 - i;
        int dim_in = ls[lev], dim_out = ls[lev + 1];
        std::string p = M + "ups." + std::to_string(i) + ".";
        int skip = skips.back();
        skips.pop_back();
        int y = b.resblock(p + "0.", lev, {x, skip}, dim_out * 2, dim_in, true);
        if (!b.ok) return 1;
        release(h, x); release(h, skip);
        size_t sa = b.slot((size_t)dim_in * 16);
        int y2 = b.resblock(p + "1.", lev, {y}, dim_in, dim_in, true, (long long)sa);
        if (!b.ok) return 1;
        release(h, y);
        int a = b.attention(M + "ups_attn." + std::to_string(i) + ".", lev, y2, dim_in, sa);
        if (!b.ok) return 1;
        release(h, y2);
        if (i < nl - 1) {
            const HostTensor* wt = find_w(h, p + "2.convTrans.weight");
            if (!wt || wt->shape.size() != 5) return fail("missing weight %s2.convTrans.weight", p.c_str());
            ConvW cu;
            if (!pack_conv(h, p + "2.convTrans.weight", p + "2.convTrans.bias", dim_in, dim_in, dim_in, (int)wt->shape[2], 4, 4, 1, false, &cu, true)) return 1;
            x = acquire(h, lev - 1, dim_in);
            b.tag(p + "2.convTrans.weight", p + "2.convTrans.bias", dim_in);
            b.conv(cu, lev, lev - 1, {a}, x, 2, 1, 1, 1, 1, -1);
            release(h, a);
        } else {
            x = a;
        }
    }
    {
        int y = b.resblock(M + "final_conv.0.", 0, {x}, ls[1], ls[0], false);
        if (!b.ok) return 1;
        release(h, x);
        x = y;
    }
    h->final_w = upload_vec(h, M + "final_conv.1.conv.weight", ls[0]);
    const HostTensor* fb = find_w(h, M + "final_conv.1.conv.bias");
    if (!h->final_w || !fb) return fail("missing final conv weights");
    h->final_b = fb->data[0];
    h->final_C = ls[0];
    h->final_h = h->bufs[x].ptr;
    if (h->train) { TapeOp t; t.kind = 3; t.src = x; t.C = ls[0]; h->tape.push_back(t); }
    {
        double m = (double)d.Z * d.P * d.R;
        h->op_info.push_back({"final_conv_update", 2.0 * m * ls[0], m * (2.0 * ls[0] + 16)});
    }
    h->ops.push_back([h](Run& r) {
        FinalParams f = r.fin;
        f.g = h->geom[0];
        f.h = h->final_h; f.C = h->final_C; f.w = h->final_w; f.bias = h->final_b;
        f.bias_ptr = h->train ? h->pptr("model.final_conv.1.conv.bias") : nullptr;
        f.x = r.x; f.sigma = r.sigma; f.sigma_stride = r.sigma_stride;
        f.sigma_data = r.scale_input ? h->d.sigma_data : -1.f;
        if (h->has_geom) {
            // grid-space raw output first, then decode + pre-conditioning + sampler update on the flat voxels
            FinalParams fg;
            memset(&fg, 0, sizeof(fg));
            fg.g = f.g; fg.h = f.h; fg.C = f.C; fg.w = f.w; fg.bias = f.bias; fg.bias_ptr = f.bias_ptr;
            fg.sigma = r.sigma; fg.sigma_stride = r.sigma_stride; fg.sigma_data = -1.f;
            fg.pred_out = h->grid_pred;
            launch_final(fg, r.B, r.st);
            launch_geom_decode(f, h->geom_map[1], h->grid_pred, r.B, r.st);
            h->launches += 2;
            return;
        }
        launch_final(f, r.B, r.st);
        h->launches += 1;
    });

    h->scratch_per_sample = b.scratch_prefix;
    h->scratch = h->dalloc<unsigned char>(h->scratch_per_sample * (size_t)d.max_batch, true);
    h->addvec = h->dalloc<float>((size_t)d.max_batch * (h->addvec_stride > 0 ? h->addvec_stride : 1), true);
    h->embed.addvec_stride = h->addvec_stride;
    if (!h->scratch || !h->addvec) return fail("out of device memory for scratch");
    return 0;
}

static int run_forward(cd_unet* h, Run& r) {
    if (h->scratch_per_sample)
        { launch_copy_bytes(nullptr, h->scratch, h->scratch_per_sample * (size_t)r.B, r.st); h->launches += 1; }
    if (h->profile) {
        // eager per-op timing with CUDA events on the launch stream (diagnostic mode; not capturable):
        // events are recorded between the ops without host synchronisation, read back after the last op
        const size_t n = h->ops.size();
        while (h->prof_evs.size() < n + 1) { cudaEvent_t ev; cudaEventCreate(&ev); h->prof_evs.push_back(ev); }
        cudaEventRecord(h->prof_evs[0], r.st);
        for (size_t i = 0; i < n; ++i) { h->ops[i](r); cudaEventRecord(h->prof_evs[i + 1], r.st); }
        cudaEventSynchronize(h->prof_evs[n]);
        for (size_t i = 0; i < n; ++i) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, h->prof_evs[i], h->prof_evs[i + 1]);
            ProfAcc& a = h->prof[h->op_info[i].name];
            a.count += 1; a.ms += ms;
            a.flops += h->op_info[i].flops * r.B; a.bytes += h->op_info[i].bytes * r.B;
        }
    } else if (h->debug_sync) {
        // diagnostic: synchronise after every op so a faulting kernel is reported by name
        for (size_t i = 0; i < h->ops.size(); ++i) {
            h->ops[i](r);
            cudaError_t es = cudaStreamSynchronize(r.st);
            if (es != cudaSuccess) return fail("op %zu (%s) failed: %s", i, h->op_info[i].name.c_str(), cudaGetErrorString(es));
        }
    } else {
        for (auto& op : h->ops) op(r);
    }
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) return fail("kernel launch failed: %s", cudaGetErrorString(e));
    return 0;
}

}  // namespace cd

// -------------------------------------------------------------------------------------------------
// dense <-> PLCP converters (test hooks only)
// -------------------------------------------------------------------------------------------------
namespace cd {
__global__ void dense_to_plcp_kernel(Geom g, const float* __restrict__ x, __half* __restrict__ dst, int C, long long nrows) {
    long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    RowInfo ri = decode_row(g, row);
    int p = ri.pp - 1;
    if (p < 0) p += g.P;
    if (p >= g.P) p -= g.P;
    long long sp = (long long)g.Z * g.P * g.R;
    for (int k = 0; k < C / 8; ++k) {
        float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (ri.kind != 2)
            for (int j = 0; j < 8; ++j)
                f[j] = x[((long long)ri.b * C + k * 8 + j) * sp + ((long long)ri.z * g.P + p) * g.R + ri.rr];
        store8(dst + plcp_offset(g, k, row), pack8(f));
    }
}
__global__ void plcp_to_dense_kernel(Geom g, const __half* __restrict__ src, float* __restrict__ y, int C, long long nvox) {
    long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nvox) return;
    long long sp = (long long)g.Z * g.P * g.R;
    int b = (int)(v / sp);
    int rem = (int)(v - b * sp);
    int z = rem / (g.P * g.R);
    int rem2 = rem - z * g.P * g.R;
    int p = rem2 / g.R, r = rem2 - p * g.R;
    long long row = interior_row(g, b, z, p, r);
    for (int k = 0; k < C / 8; ++k) {
        Half8 hv = load8(src + plcp_offset(g, k, row));
        float f[8];
        unpack8(hv, f);
        for (int j = 0; j < 8; ++j) y[((long long)b * C + k * 8 + j) * sp + rem] = f[j];
    }
}
}  // namespace cd

namespace cd { static int train_setup(cd_unet* h); }

// -------------------------------------------------------------------------------------------------
// C ABI
// -------------------------------------------------------------------------------------------------
extern "C" {

const char* cd_last_error(void) { return cd::g_err; }
int cd_version(void) { return 1; }

int cd_device_check(void) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail("no CUDA device: %s", cudaGetErrorString(e));
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, dev);
    if (e != cudaSuccess) return fail("cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    if (prop.major != 10) return fail("device %s is sm_%d%d; this library is built for sm_100a only", prop.name, prop.major, prop.minor);
    return 0;
}

cd_unet* cd_unet_create(const cd_unet_desc* desc) {
    if (!desc) { fail("null desc"); return nullptr; }
    if (cd_device_check()) return nullptr;
    if (desc->n_sizes < 2 || desc->n_sizes > CD_MAX_LEVELS + 1) { fail("unsupported LAYER_SIZE_UNET length %d", desc->n_sizes); return nullptr; }
    for (int i = 0; i < desc->n_sizes; ++i)
        if (desc->layer_sizes[i] % 8 != 0 || desc->layer_sizes[i] % desc->groups != 0) {
            fail("LAYER_SIZE_UNET[%d]=%d must be a multiple of 8 and of BLOCK_GROUPS", i, desc->layer_sizes[i]);
            return nullptr;
        }
    if (desc->layer_sizes[0] != desc->layer_sizes[1]) { fail("final ResnetBlock needs LAYER_SIZE_UNET[0]==[1] (models.py:697-699)"); return nullptr; }
    if (desc->in_channels > 8 || desc->R > 64 || desc->max_batch < 1) { fail("unsupported desc"); return nullptr; }
    cd_unet* h = new cd_unet();
    h->d = *desc;
    h->n_levels = desc->n_sizes - 1;
    int Z = desc->Z, P = desc->P, R = desc->R;
    for (int l = 0; l < h->n_levels; ++l) {
        if (Z < 1 || P < 1 || R < 1) { fail("level %d has an empty extent", l); delete h; return nullptr; }
        h->geom[l] = make_geom(Z, P, R, desc->max_batch);
        Z = (Z + 1) / 2; P = P / 2; R = R / 2;  // models.py:629-634
    }
    return h;
}

cd_unet* cd_layer_model_create(const cd_layer_desc* desc) {
    if (!desc) { fail("null desc"); return nullptr; }
    if (cd_device_check()) return nullptr;
    if (desc->dim_in < 1 || desc->dim_in > 64 || desc->hidden < 8 || desc->hidden > 256 || desc->num_layers < 2 || desc->num_layers > 9 ||
        desc->cond_size < 1 || desc->cond_size > 64 || desc->cond_dim < 8 || desc->cond_dim > 128 || desc->cond_dim % 4 || desc->max_batch < 1) {
        fail("unsupported layer-model shape (dim_in %d, hidden %d, layers %d, cond %d/%d)", desc->dim_in, desc->hidden, desc->num_layers,
             desc->cond_size, desc->cond_dim);
        return nullptr;
    }
    cd_unet* h = new cd_unet();
    memset(&h->d, 0, sizeof(h->d));
    h->dense = true;
    h->ld = *desc;
    // the shared machinery (state buffers, sampling loop, Philox keys) sees a (dim_in, 1, 1) "grid"
    h->d.Z = desc->dim_in; h->d.P = 1; h->d.R = 1; h->d.n_sizes = 2; h->d.max_batch = desc->max_batch;
    h->d.cond_size = desc->cond_size; h->d.cond_dim = desc->cond_dim; h->d.sigma_data = desc->sigma_data; h->d.time_embed = desc->time_embed;
    h->n_levels = 1;
    h->geom[0] = make_geom(desc->dim_in, 1, 1, desc->max_batch);
    return h;
}

void cd_unet_destroy(cd_unet* h) {
    if (!h) return;
    if (h->graph_exec) cudaGraphExecDestroy(h->graph_exec);
    if (h->capture_stream) cudaStreamDestroy(h->capture_stream);
    for (cudaEvent_t ev : h->prof_evs) cudaEventDestroy(ev);
    if (h->fault_host) cudaFreeHost(h->fault_host);
    for (void* p : h->dev_allocs) cudaFree(p);
    delete h;
}

int cd_unet_set_weight(cd_unet* h, const char* name, const float* host_data, const int64_t* shape, int ndim) {
    if (!h || !name || !host_data) return fail("null argument");
    if (h->finalized) return fail("weights are frozen after cd_unet_finalize; create a new handle");
    HostTensor t;
    int64_t n = 1;
    for (int i = 0; i < ndim; ++i) { t.shape.push_back(shape[i]); n *= shape[i]; }
    t.data.assign(host_data, host_data + n);
    h->weights[name] = std::move(t);
    return 0;
}

int cd_unet_set_geom_map(cd_unet* h, int which, int n_rows, int n_cols, const int32_t* rowptr, const int32_t* col,
                         const float* val) {
    if (!h || !rowptr) return fail("null argument");
    if (h->finalized) return fail("geometry maps are frozen after cd_unet_finalize; create a new handle");
    if (which < 0 || which > 1 || n_rows < 1 || n_cols < 1) return fail("bad geometry map request (which=%d, %dx%d)", which, n_rows, n_cols);
    if (rowptr[0] != 0) return fail("geometry map: rowptr[0] must be 0");
    for (int i = 0; i < n_rows; ++i)
        if (rowptr[i + 1] < rowptr[i]) return fail("geometry map: rowptr is not non-decreasing at row %d", i);
    const int nnz = rowptr[n_rows];
    if (nnz > 0 && (!col || !val)) return fail("null argument");
    for (int k = 0; k < nnz; ++k)
        if (col[k] < 0 || col[k] >= n_cols) return fail("geometry map: column %d outside [0,%d) at entry %d", col[k], n_cols, k);
    auto& m = h->geom_host[which];
    m.n_rows = n_rows; m.n_cols = n_cols;
    m.rowptr.assign(rowptr, rowptr + n_rows + 1);
    m.col.assign(col, col + nnz);
    m.val.assign(val, val + nnz);
    return 0;
}

int cd_unet_voxels(const cd_unet* h) {
    if (!h) return 0;
    if (!h->finalized && h->geom_host[1].n_rows > 0) return h->geom_host[1].n_rows;
    return h->voxels();
}

int cd_unet_set_option(cd_unet* h, const char* key, int value) {
    if (!h || !key) return fail("null argument");
    if (!strcmp(key, "conv_backend")) {
        h->conv_backend = value;
        if (h->graph_exec) { cudaGraphExecDestroy(h->graph_exec); h->graph_exec = nullptr; h->graph_key.clear(); }
        return 0;
    }
    if (!strcmp(key, "debug_sync")) { h->debug_sync = value; return 0; }
    if (!strcmp(key, "train")) {
        if (h->finalized) return fail("option train must be set before cd_unet_finalize");
        h->train = value != 0;
        return 0;
    }
    if (!strcmp(key, "check_finite")) { h->check_finite = value ? 1 : 0; return 0; }
    if (!strcmp(key, "online_attention")) {
        h->online_attention = value ? 1 : 0;
        if (h->graph_exec) { cudaGraphExecDestroy(h->graph_exec); h->graph_exec = nullptr; h->graph_key.clear(); }
        return 0;
    }
    if (!strcmp(key, "fused_attention")) {
        h->no_fused_attention = value ? 0 : 1;
        if (h->graph_exec) { cudaGraphExecDestroy(h->graph_exec); h->graph_exec = nullptr; h->graph_key.clear(); }
        return 0;
    }
    if (!strcmp(key, "profile")) {
        h->profile = value;
        if (value == 2) { h->prof.clear(); h->profile = 1; }   // 2 = enable and reset
        return 0;
    }
    return fail("unknown option %s", key);
}

int cd_unet_finalize(cd_unet* h) {
    if (!h) return fail("null handle");
    if (h->finalized) return 0;
    std::vector<float> rv(h->d.r_centres, h->d.r_centres + h->d.R);
    h->r_vals = h->upload(rv);
    const int ngrid = h->d.Z * h->d.P * h->d.R;
    if ((h->geom_host[0].n_rows > 0) != (h->geom_host[1].n_rows > 0)) return fail("geometry maps: both enc (0) and dec (1) must be set");
    if (h->geom_host[0].n_rows > 0) {
        const auto& en = h->geom_host[0];
        const auto& de = h->geom_host[1];
        if (en.n_rows != ngrid || de.n_cols != ngrid || en.n_cols != de.n_rows)
            return fail("geometry maps: enc is %dx%d, dec is %dx%d; expected %dxN and Nx%d", en.n_rows, en.n_cols, de.n_rows, de.n_cols,
                        ngrid, ngrid);
        for (int w = 0; w < 2; ++w) {
            const auto& m = h->geom_host[w];
            int* rp = h->dalloc<int>(m.rowptr.size(), false);
            int* cl = h->dalloc<int>(m.col.size() ? m.col.size() : 1, false);
            float* vl = h->dalloc<float>(m.val.size() ? m.val.size() : 1, false);
            if (!rp || !cl || !vl) return fail("out of device memory for geometry maps");
            cudaMemcpy(rp, m.rowptr.data(), m.rowptr.size() * sizeof(int), cudaMemcpyHostToDevice);
            if (!m.col.empty()) {
                cudaMemcpy(cl, m.col.data(), m.col.size() * sizeof(int), cudaMemcpyHostToDevice);
                cudaMemcpy(vl, m.val.data(), m.val.size() * sizeof(float), cudaMemcpyHostToDevice);
            }
            h->geom_map[w] = GeomMap{rp, cl, vl, m.n_rows, m.n_cols};
        }
        h->grid_pred = h->dalloc<float>((size_t)h->d.max_batch * ngrid, true);
        if (!h->grid_pred) return fail("out of device memory for the grid output");
        h->has_geom = true;
    }
    if (build_plan(h)) return 1;
    if (!h->dense) {
        h->attn_part = h->dalloc<float>((size_t)h->d.max_batch * kAttnMaxBlocks * attn_part_floats(), false);
        if (!h->attn_part) return fail("out of device memory for the attention partials");
    }
    size_t nv = (size_t)h->d.max_batch * h->voxels();
    for (int i = 0; i < CD_NUM_STATE_BUFS; ++i) {
        h->state[i] = h->dalloc<float>(nv, true);
        if (!h->state[i]) return fail("out of device memory for sampler state");
    }
    h->coef_cap = 4096;
    h->coef_tab = h->dalloc<float>((size_t)h->coef_cap * kRow, true);
    h->fault = h->dalloc<int>(1, true);
    if (!h->coef_tab || !h->fault || cudaMallocHost((void**)&h->fault_host, sizeof(int)) != cudaSuccess)
        return fail("out of memory for the sampler tables");
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return fail("finalize: %s", cudaGetErrorString(e));
    for (auto& b : h->bufs) if (!b.ptr) return fail("out of device memory for activations");
    if (h->train && cd::train_setup(h)) return 1;
    h->weights.clear();
    h->finalized = true;
    return 0;
}

int64_t cd_unet_workspace_bytes(const cd_unet* h) { return h ? h->workspace_bytes : 0; }
int64_t cd_unet_launch_count(const cd_unet* h) { return h ? h->launches : 0; }

int cd_unet_profile_report(const cd_unet* h, char* buf, int cap) {
    if (!h || !buf || cap < 1) return fail("bad argument");
    int n = 0;
    buf[0] = 0;
    for (auto& kv : h->prof) {
        int w = snprintf(buf + n, (size_t)(cap - n), "%s %lld %.6f %.6e %.6e\n", kv.first.c_str(), kv.second.count, kv.second.ms,
                         kv.second.flops, kv.second.bytes);
        if (w < 0 || w >= cap - n) return fail("profile report buffer too small");
        n += w;
    }
    return 0;
}

int cd_level_shape(const cd_unet* h, int level, int* zpr) {
    if (!h || level < 0 || level >= h->n_levels) return fail("bad level");
    zpr[0] = h->geom[level].Z; zpr[1] = h->geom[level].P; zpr[2] = h->geom[level].R;
    return 0;
}

static int check_call(cd_unet* h, int B) {
    if (!h) return fail("null handle");
    if (!h->finalized) return fail("cd_unet_finalize has not been called");
    if (B < 1 || B > h->d.max_batch) return fail("batch %d outside [1, max_batch=%d]", B, h->d.max_batch);
    return 0;
}

int cd_unet_forward(cd_unet* h, const float* x_dev, const float* time_dev, const float* cond_dev, float* pred_dev, int B,
                    void* stream) {
    if (check_call(h, B)) return 1;
    Run r;
    memset(&r.fin, 0, sizeof(r.fin));
    r.B = B; r.st = (cudaStream_t)stream; r.x = x_dev; r.sigma = time_dev; r.sigma_stride = 1; r.cond = cond_dev;
    r.scale_input = 0; r.time_is_raw = 1;
    r.fin.pred_out = pred_dev;
    return run_forward(h, r);
}

int cd_denoise(cd_unet* h, const float* x_dev, const float* sigma_dev, const float* cond_dev, float* x0_dev, int B, void* stream) {
    if (check_call(h, B)) return 1;
    Run r;
    memset(&r.fin, 0, sizeof(r.fin));
    r.B = B; r.st = (cudaStream_t)stream; r.x = x_dev; r.sigma = sigma_dev; r.sigma_stride = 1; r.cond = cond_dev;
    r.scale_input = 1; r.time_is_raw = 0;
    r.fin.x0_out = x0_dev;
    return run_forward(h, r);
}

int cd_fill_normal(float* out_dev, int64_t n_showers, int64_t voxels, uint64_t seed, int shower_offset, int slot, void* stream) {
    if (cd_device_check()) return 1;
    launch_fill_normal(out_dev, n_showers * voxels, (int)voxels, seed, shower_offset, slot, (cudaStream_t)stream);
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) return fail("fill_normal: %s", cudaGetErrorString(e));
    return 0;
}

int cd_geom_apply(const int32_t* rowptr_dev, const int32_t* col_dev, const float* val_dev, int n_rows, int n_cols,
                  const float* x_dev, float* y_dev, int B, void* stream) {
    if (!rowptr_dev || !col_dev || !val_dev || !x_dev || !y_dev) return fail("null argument");
    if (cd_device_check()) return 1;
    if (n_rows < 1 || n_cols < 1 || B < 1) return fail("bad geom_apply shape");
    GeomMap m{rowptr_dev, col_dev, val_dev, n_rows, n_cols};
    launch_geom_apply(m, x_dev, y_dev, B, (cudaStream_t)stream);
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) return fail("geom_apply: %s", cudaGetErrorString(e));
    return 0;
}

int cd_reverse_norm_ex(const cd_reverse_norm_ex_desc* d, const int32_t* seg_dev, const float* vox_dev, const float* e_dev,
                       const float* layerE_dev, float* out_dev, float* energy_out_dev, int B, void* stream) {
    if (!d || !vox_dev || !e_dev || !out_dev) return fail("null argument");
    if (cd_device_check()) return 1;
    if (B < 1 || B > 65535 || d->nseg < 1 || d->nseg > 65535 || d->total < d->nseg) return fail("bad reverse_norm shape");
    if (!seg_dev && d->total % d->nseg != 0) return fail("reverse_norm: %d values do not split into %d equal layers", d->total, d->nseg);
    if (d->map_kind < 0 || d->map_kind > 2) return fail("reverse_norm: unknown map kind %d", d->map_kind);
    if (d->energy_mode == 0 && (!(d->emin > 0.f) || !(d->emax > 0.f))) return fail("bad reverse_norm constants");
    if (d->e_stride < 1) return fail("reverse_norm: e_stride must be >= 1");
    RevNormParams p;
    p.vox = vox_dev; p.e = e_dev; p.e_stride = d->e_stride; p.layerE = layerE_dev; p.out = out_dev; p.energy_out = energy_out_dev;
    p.nseg = d->nseg; p.total = d->total; p.seg = seg_dev; p.map_kind = d->map_kind;
    p.a_scale = d->a_scale; p.a_shift = d->a_shift; p.alpha = d->alpha; p.log_cap = d->log_cap;
    p.post_scale = d->post_scale; p.post_shift = d->post_shift;
    p.totalE_mean = d->totalE_mean; p.totalE_std = d->totalE_std; p.layers_mean = d->layers_mean; p.layers_std = d->layers_std;
    p.guard_eps = d->guard_eps; p.energy_mode = d->energy_mode; p.emin = d->emin; p.emax = d->emax;
    p.max_deposit = d->max_deposit; p.ecut = d->ecut; p.final_scale = d->final_scale;
    launch_reverse_norm(p, B, (cudaStream_t)stream);
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) return fail("reverse_norm: %s", cudaGetErrorString(e));
    return 0;
}

int cd_reverse_norm(const cd_reverse_norm_desc* d, const float* vox_dev, const float* e_dev, const float* layerE_dev,
                    float* out_dev, float* energy_out_dev, int B, void* stream) {
    if (!d) return fail("null argument");
    if (d->Z < 1 || d->Z > 65535 || d->PR < 1) return fail("bad reverse_norm shape");
    if (!(d->emin > 0.f) || !(d->emax > 0.f) || !(d->logit_std > 0.f)) return fail("bad reverse_norm constants");
    cd_reverse_norm_ex_desc x;
    memset(&x, 0, sizeof(x));
    x.nseg = d->Z; x.total = d->Z * d->PR; x.map_kind = 0; x.a_scale = d->logit_std; x.a_shift = d->logit_mean; x.alpha = 1e-6f;
    x.post_scale = 1.f; x.post_shift = 0.f;
    x.totalE_mean = d->totalE_mean; x.totalE_std = d->totalE_std; x.layers_mean = d->layers_mean; x.layers_std = d->layers_std;
    x.guard_eps = 1e-6f; x.energy_mode = 0; x.emin = d->emin; x.emax = d->emax; x.max_deposit = d->max_deposit; x.ecut = d->ecut;
    x.e_stride = 1; x.final_scale = 1;
    return cd_reverse_norm_ex(&x, nullptr, vox_dev, e_dev, layerE_dev, out_dev, energy_out_dev, B, stream);
}

int cd_geom_decode_sparse(const int32_t* colptr_dev, const int32_t* row_dev, const float* val_dev, int n_cols, int n_rows,
                          int cells_per_layer, int bins_per_layer, const float* x_dev, float* out_dev, int B, int per_batch,
                          uint64_t seed, const float* u_bank_dev, void* stream) {
    if (!colptr_dev || !row_dev || !val_dev || !x_dev || !out_dev) return fail("null argument");
    if (cd_device_check()) return 1;
    if (n_cols < 1 || n_rows < 1 || B < 1 || cells_per_layer < 1 || bins_per_layer < 1 || n_cols % bins_per_layer ||
        n_rows % cells_per_layer || n_cols / bins_per_layer != n_rows / cells_per_layer)
        return fail("bad sparse decode shape (%d bins, %d cells, %d / %d per layer)", n_cols, n_rows, bins_per_layer, cells_per_layer);
    SparseDecParams p;
    p.colptr = colptr_dev; p.row = row_dev; p.val = val_dev; p.n_cols = n_cols; p.n_rows = n_rows;
    p.cells_per_layer = cells_per_layer; p.bins_per_layer = bins_per_layer; p.x = x_dev; p.out = out_dev;
    p.per_batch = per_batch; p.seed = seed; p.u_bank = u_bank_dev;
    launch_sparse_decode(p, B, (cudaStream_t)stream);
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) return fail("sparse decode: %s", cudaGetErrorString(e));
    return 0;
}

}  // extern "C"

// -------------------------------------------------------------------------------------------------
// sampling loop
// -------------------------------------------------------------------------------------------------
namespace cd {

struct SampleArgs {
    const float* start; float start_scale; const float* cond; const cd_eval* evals; int n_evals;
    const float* bank; uint64_t seed; int shower_offset; int final_buf; float* x_final; float* xs; float* x0s; int B;
};

static int enqueue_sample(cd_unet* h, const SampleArgs& a, cudaStream_t st) {
    const int V = h->voxels();
    const long long n = (long long)a.B * V;
    const size_t bytes = (size_t)n * sizeof(float);
    float* tab = h->coef_tab;
    // Philox key {seed lo, seed hi, shower_offset}: device-resident (row n_evals, slots 2..4) so the graph does not bake it
    const uint32_t* rng_key = reinterpret_cast<const uint32_t*>(tab + (size_t)a.n_evals * kRow + 2);
    // buf[0] = start * start_scale  (row n_evals of the table holds [start_scale, 0])
    launch_axpby(a.start, nullptr, h->state[0], tab + (size_t)a.n_evals * kRow, n, V, a.seed, a.shower_offset, 0, 0, st);
    h->launches += 1;
    for (int i = 0; i < a.n_evals; ++i) {
        const cd_eval& e = a.evals[i];
        float* row = tab + (size_t)i * kRow;
        bool recorded_x = false;
        if (e.pre_b != 0.f || e.pre_src >= 0) {
            const int src = e.pre_src >= 0 ? e.pre_src : e.x_in;
            if (e.rec_x >= 0 && a.xs && e.rec_x_buf == e.x_in && src == e.x_in) {
                // the reference records the state before the churn noise is added (models/sample.py:660)
                launch_copy_bytes(h->state[e.rec_x_buf], a.xs + (size_t)e.rec_x * n, bytes, st); h->launches += 1;
                recorded_x = true;
            }
            const float* nz = a.bank ? a.bank + (size_t)e.pre_noise_index * n : nullptr;
            launch_axpby(h->state[src], nz, h->state[e.x_in], row + 6, n, V, a.seed, a.shower_offset, e.pre_noise_index, 1, st,
                         rng_key);
            h->launches += 1;
        }
        Run r;
        memset(&r.fin, 0, sizeof(r.fin));
        r.B = a.B; r.st = st; r.x = h->state[e.x_in]; r.sigma = row; r.sigma_stride = 0; r.cond = a.cond; r.scale_input = 1; r.time_is_raw = 0;
        float* x0_dst = nullptr;
        if (e.x0_buf >= 0) x0_dst = h->state[e.x0_buf];
        else if (e.rec_x0 >= 0 && a.x0s) x0_dst = a.x0s + (size_t)e.rec_x0 * n;
        r.fin.x0_out = x0_dst;
        r.fin.fault = h->fault; r.fin.eval_index = i;
        r.fin.coef = row + 1;
        if (e.d_out >= 0) r.fin.d_out = h->state[e.d_out];
        r.fin.rng_key = rng_key;
        if (e.x_out >= 0) {
            r.fin.x_out = h->state[e.x_out];
            r.fin.coef = row + 1;
            if (e.a1 >= 0) r.fin.a1 = h->state[e.a1];
            else if (e.a1 == -2) {
                if (a.bank) r.fin.a1 = a.bank + (size_t)e.noise_index * n;
                else { r.fin.rng_enabled = 1; r.fin.seed = a.seed; r.fin.shower_offset = a.shower_offset; r.fin.step = e.noise_index; }
            }
            if (e.a2 >= 0) r.fin.a2 = h->state[e.a2];
            if (e.a3 >= 0) r.fin.a3 = h->state[e.a3];
        }
        if (run_forward(h, r)) return 1;
        if (e.x0_buf >= 0 && e.rec_x0 >= 0 && a.x0s)
            { launch_copy_bytes(h->state[e.x0_buf], a.x0s + (size_t)e.rec_x0 * n, bytes, st); h->launches += 1; }
        if (e.rec_x >= 0 && a.xs && !recorded_x)
            { launch_copy_bytes(h->state[e.rec_x_buf], a.xs + (size_t)e.rec_x * n, bytes, st); h->launches += 1; }
    }
    launch_copy_bytes(h->state[a.final_buf], a.x_final, bytes, st);
    h->launches += 1;
    cudaError_t err = cudaPeekAtLastError();
    if (err != cudaSuccess) return fail("sample enqueue: %s", cudaGetErrorString(err));
    return 0;
}

}  // namespace cd

static void cd_segv_handler(int sig) {
    void* frames[64];
    int n = backtrace(frames, 64);
    const char msg[] = "calodiffusion_b200: fatal signal, native backtrace:\n";
    (void)!write(2, msg, sizeof(msg) - 1);
    backtrace_symbols_fd(frames, n, 2);
    signal(sig, SIG_DFL);
    raise(sig);
}

extern "C" int cd_sample_run(cd_unet* h, const float* start_dev, float start_scale, const float* cond_dev, const cd_eval* evals,
                             int n_evals, const float* noise_bank_dev, uint64_t seed, int shower_offset, int final_buf,
                             float* x_final_dev, float* xs_dev, float* x0s_dev, int B, void* stream, int use_graph) {
    if (getenv("CD_SEGV_TRACE")) signal(SIGSEGV, cd_segv_handler);   // diagnostic: native backtrace on a crash
    if (check_call(h, B)) return 1;
    if (n_evals < 1 || n_evals + 1 > h->coef_cap) return fail("n_evals %d outside [1, %d]", n_evals, h->coef_cap - 1);
    if (final_buf < 0 || final_buf >= CD_NUM_STATE_BUFS) return fail("bad final_buf");
    for (int i = 0; i < n_evals; ++i) {
        const cd_eval& e = evals[i];
        auto okb = [](int b, int lo) { return b >= lo && b < CD_NUM_STATE_BUFS; };
        if (!okb(e.x_in, 0) || !okb(e.x_out, -1) || !okb(e.x0_buf, -1) || !okb(e.a1, -2) || !okb(e.a2, -1) || !okb(e.a3, -1) ||
            !okb(e.pre_src, -1) || !okb(e.d_out, -1) ||
            (e.rec_x >= 0 && !okb(e.rec_x_buf, 0)))
            return fail("eval %d references a state buffer outside [0,%d)", i, CD_NUM_STATE_BUFS);
        if (!(e.sigma > 0.f)) return fail("eval %d: sigma must be > 0 (got %g)", i, (double)e.sigma);  // sample.py:166-167
    }
    cudaStream_t st = (cudaStream_t)stream;
    std::vector<float> tab((size_t)(n_evals + 1) * kRow, 0.f);
    for (int i = 0; i < n_evals; ++i) {
        float* row = &tab[(size_t)i * kRow];
        row[0] = evals[i].sigma;
        for (int j = 0; j < 5; ++j) row[1 + j] = evals[i].k[j];
        row[6] = evals[i].pre_a; row[7] = evals[i].pre_b; row[8] = evals[i].d_scale;
    }
    tab[(size_t)n_evals * kRow] = start_scale;
    {
        const uint32_t key[3] = {(uint32_t)seed, (uint32_t)(seed >> 32), (uint32_t)shower_offset};
        memcpy(&tab[(size_t)n_evals * kRow + 2], key, sizeof(key));
    }
    CD_CUDA(cudaMemcpyAsync(h->coef_tab, tab.data(), tab.size() * sizeof(float), cudaMemcpyHostToDevice, st));
    *h->fault_host = 0x7fffffff;
    CD_CUDA(cudaMemcpyAsync(h->fault, h->fault_host, sizeof(int), cudaMemcpyHostToDevice, st));
    // after the loop: one 4-byte readback (the only synchronisation; nothing inside the graph waits on the host)
    auto check_fault = [&]() -> int {
        if (!h->check_finite) return 0;
        CD_CUDA(cudaMemcpyAsync(h->fault_host, h->fault, sizeof(int), cudaMemcpyDeviceToHost, st));
        CD_CUDA(cudaStreamSynchronize(st));
        if (*h->fault_host != 0x7fffffff)
            return fail("non-finite network output at evaluation %d of %d (sigma %g): activations left the fp16 range "
                        "(supported domain |c_in*x| < 65504); set option check_finite=0 to get the non-finite showers instead",
                        *h->fault_host, n_evals, (double)evals[*h->fault_host].sigma);
        return 0;
    };
    SampleArgs a{start_dev, start_scale, cond_dev, evals, n_evals, noise_bank_dev, seed, shower_offset, final_buf,
                 x_final_dev, xs_dev, x0s_dev, B};
    if (!use_graph || h->profile) { if (enqueue_sample(h, a, st)) return 1; return check_fault(); }

    // key: everything baked into the graph (structure + pointers); numeric schedule values live in coef_tab
    std::vector<unsigned char> key;
    auto put = [&key](const void* p, size_t n) { const unsigned char* c = (const unsigned char*)p; key.insert(key.end(), c, c + n); };
    put(&start_dev, sizeof(void*)); put(&cond_dev, sizeof(void*)); put(&noise_bank_dev, sizeof(void*));
    put(&x_final_dev, sizeof(void*)); put(&xs_dev, sizeof(void*)); put(&x0s_dev, sizeof(void*));
    put(&final_buf, sizeof(int)); put(&B, sizeof(int));     // seed / shower_offset are device-resident (coef_tab), not baked
    put(&h->conv_backend, sizeof(int)); put(&n_evals, sizeof(int));
    for (int i = 0; i < n_evals; ++i) {
        const cd_eval& e = evals[i];
        int s[14] = {e.x_in, e.x_out, e.x0_buf, e.a1, e.a2, e.a3, e.noise_index, e.pre_b != 0.f, e.pre_noise_index, e.rec_x0, e.rec_x,
                     e.rec_x_buf, e.pre_src, e.d_out};
        put(s, sizeof(s));
    }
    if (!h->graph_exec || key != h->graph_key) {
        if (h->graph_exec) { cudaGraphExecDestroy(h->graph_exec); h->graph_exec = nullptr; }
        h->graph_key.clear();
        int64_t before = h->launches;
        // capture on a private stream: the caller's stream may be the legacy default stream, which cannot capture
        if (!h->capture_stream) CD_CUDA(cudaStreamCreateWithFlags(&h->capture_stream, cudaStreamNonBlocking));
        CD_CUDA(cudaStreamBeginCapture(h->capture_stream, cudaStreamCaptureModeThreadLocal));
        int rc = enqueue_sample(h, a, h->capture_stream);
        cudaGraph_t graph = nullptr;
        cudaError_t ce = cudaStreamEndCapture(h->capture_stream, &graph);
        h->graph_nodes = h->launches - before;
        h->launches = before;
        if (rc) { if (graph) cudaGraphDestroy(graph); return 1; }
        if (ce != cudaSuccess) return fail("graph capture failed: %s", cudaGetErrorString(ce));
        ce = cudaGraphInstantiate(&h->graph_exec, graph, 0);
        cudaGraphDestroy(graph);
        if (ce != cudaSuccess) { h->graph_exec = nullptr; return fail("graph instantiate failed: %s", cudaGetErrorString(ce)); }
        h->graph_key = key;
    }
    CD_CUDA(cudaGraphLaunch(h->graph_exec, st));
    h->launches += h->graph_nodes;
    return check_fault();
}

// -------------------------------------------------------------------------------------------------
// test hook: one convolution in isolation, dense fp32 in/out
// -------------------------------------------------------------------------------------------------
extern "C" int cd_test_conv(cd_unet* h, int level_in, int level_out, int Cin, int Cout, const int* ksize, const int* stride,
                            int transposed, const float* w_host, const float* bias_host, const float* x_dev, float* y_dev,
                            int B, int backend, void* stream) {
    if (check_call(h, B)) return 1;
    if (level_in < 0 || level_in >= h->n_levels || level_out < 0 || level_out >= h->n_levels) return fail("bad level");
    if (Cin % 8 || Cout % 8) return fail("channels must be multiples of 8");
    cudaStream_t st = (cudaStream_t)stream;
    const Geom& gi = h->geom[level_in];
    const Geom& go = h->geom[level_out];
    HostTensor wt;
    int taps = ksize[0] * ksize[1] * ksize[2];
    wt.shape = {transposed ? Cin : Cout, transposed ? Cout : Cin, ksize[0], ksize[1], ksize[2]};
    wt.data.assign(w_host, w_host + (size_t)Cin * Cout * taps);
    h->weights["__test.w"] = wt;
    if (bias_host) {
        HostTensor bt;
        bt.shape = {Cout};
        bt.data.assign(bias_host, bias_host + Cout);
        h->weights["__test.b"] = bt;
    }
    ConvW cw;
    bool okp = pack_conv(h, "__test.w", bias_host ? "__test.b" : "", Cin, Cin, Cout, ksize[0], ksize[1], ksize[2], transposed, true, &cw, true);
    h->weights.clear();
    if (!okp) return 1;
    __half *xin = nullptr, *yout = nullptr;
    size_t nin = (size_t)(Cin / 8) * gi.chunk_stride * 8, nout = (size_t)(Cout / 8) * go.chunk_stride * 8;
    CD_CUDA(cudaMalloc(&xin, nin * sizeof(__half)));
    CD_CUDA(cudaMalloc(&yout, nout * sizeof(__half)));
    cudaMemsetAsync(xin, 0, nin * sizeof(__half), st);
    cudaMemsetAsync(yout, 0, nout * sizeof(__half), st);
    long long nrows = (long long)B * gi.rows_per_sample;
    dense_to_plcp_kernel<<<(unsigned)((nrows + 127) / 128), 128, 0, st>>>(gi, x_dev, xin, Cin, nrows);
    ConvParams p;
    p.gin = gi; p.gout = go; p.nsrc = 1; p.Cin_total = Cin;
    p.src[0].base = xin; p.src[0].C = Cin; p.src[1].base = nullptr; p.src[1].C = 0;
    p.dst = yout; p.Cout = Cout; p.w = cw.w; p.bias = cw.bias;
    p.kz = ksize[0]; p.kp = ksize[1]; p.kr = ksize[2];
    p.sz = stride[0]; p.sp = stride[1]; p.sr = stride[2];
    bool is1 = taps == 1;
    p.pz = is1 ? 0 : 1; p.cp = is1 ? 0 : 1; p.pr = is1 ? 0 : 1;
    p.transposed = transposed; p.write_halo = 0; p.stats = nullptr; p.stats_cpg = 1; p.B = B;
    p.pn_stats = nullptr; p.pn_wsum = nullptr; p.pn_wb = nullptr; p.pn_C = 0; p.pn_count = 1.0;
    int rc = 0;
    if (backend == 0) {
        if (!cw.wmma || launch_conv_mma(p, cw.wmma, st) != 0) rc = fail("tcgen05 conv does not support this shape");
    } else if (backend == 3) {
        if (!cw.wz || launch_conv_zstack(p, cw.wz, 1, st) != 0) rc = fail("Z-stacked tcgen05 conv does not support this shape");
    } else if (backend == 2) {
        p.write_halo = 1;
        if (!cw.wgat || launch_conv_gather(p, cw.wgat, st) != 0) rc = fail("gather tcgen05 conv does not support this shape");
    } else {
        launch_conv_generic(p, st);
    }
    long long nvox = (long long)B * go.Z * go.P * go.R;
    if (!rc) plcp_to_dense_kernel<<<(unsigned)((nvox + 127) / 128), 128, 0, st>>>(go, yout, y_dev, Cout, nvox);
    cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(xin); cudaFree(yout);
    if (e != cudaSuccess) return fail("test conv: %s", cudaGetErrorString(e));
    return rc;
}

// -------------------------------------------------------------------------------------------------
// The same convolution with NO allocation: all scratch (repacked weights + the two fp16 PLCP staging tensors) lives in one
// file-scope device array.  Every call on a device shares it (it keeps the previous call's contents; what must start at zero
// is cleared per call), so calls that overlap on the GPU race.  No counterpart in the reference.
// -------------------------------------------------------------------------------------------------
#define CD_NOALLOC_SCRATCH_BYTES (16u << 20)
static __device__ __align__(256) unsigned char g_noalloc_scratch[CD_NOALLOC_SCRATCH_BYTES];

namespace cd {
// torch layout (Cout,Cin,taps) / transposed (Cin,Cout,taps) fp32 -> the SIMT kernel's [tap][ci][co]
__global__ void noalloc_pack_w_kernel(const float* __restrict__ w, float* __restrict__ pk, int Cin, int Cout, int taps, int transposed) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= taps * Cin * Cout) return;
    int co = i % Cout, ci = (i / Cout) % Cin, t = i / (Cout * Cin);
    size_t src = transposed ? (((size_t)ci * Cout + co) * taps + t) : (((size_t)co * Cin + ci) * taps + t);
    pk[i] = w[src];
}
// device address of the scratch array, looked up once per device with that device current (the array's name in host code is
// a host-side stand-in, not a device address)
static unsigned char* noalloc_scratch_ptr() {
    static std::mutex mu;
    static unsigned char* ptrs[64] = {};
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { fail("no-alloc conv: cannot query the current device"); return nullptr; }
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) { fail("no-alloc conv: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return nullptr; }
        ptrs[dev] = static_cast<unsigned char*>(p);
    }
    return ptrs[dev];
}
}  // namespace cd

extern "C" int64_t cd_conv_noalloc_max_bytes(void) { return (int64_t)CD_NOALLOC_SCRATCH_BYTES; }

extern "C" int cd_conv_noalloc(const cd_unet* h, int level_in, int level_out, int Cin, int Cout, const int* ksize, const int* stride,
                               int transposed, const float* w_dev, const float* bias_dev, const float* x_dev, float* y_dev,
                               int B, void* stream) {
    if (!h) return fail("null handle");
    if (!h->finalized) return fail("cd_unet_finalize has not been called");
    if (B < 1 || B > h->d.max_batch) return fail("batch %d outside [1, max_batch=%d]", B, h->d.max_batch);
    if (!ksize || !stride || !w_dev || !x_dev || !y_dev) return fail("null argument");
    if (level_in < 0 || level_in >= h->n_levels || level_out < 0 || level_out >= h->n_levels) return fail("bad level");
    if (Cin < 8 || Cout < 8 || Cin % 8 || Cout % 8) return fail("channels must be multiples of 8");
    const int taps = ksize[0] * ksize[1] * ksize[2];
    if (taps < 1 || taps > 64) return fail("bad kernel size");
    cudaStream_t st = (cudaStream_t)stream;
    const Geom& gi = h->geom[level_in];
    const Geom& go = h->geom[level_out];
    auto up = [](size_t b) { return (b + 255) / 256 * 256; };
    const size_t wbytes = up((size_t)taps * Cin * Cout * sizeof(float));
    const size_t nin = (size_t)(Cin / 8) * gi.chunk_stride * 8, nout = (size_t)(Cout / 8) * go.chunk_stride * 8;
    const size_t inbytes = up(nin * sizeof(__half)), outbytes = up(nout * sizeof(__half));
    const size_t need = wbytes + inbytes + outbytes;
    if (need > (size_t)CD_NOALLOC_SCRATCH_BYTES)
        return fail("no-alloc conv needs %zu bytes of scratch, the static array holds %u (weights %zu + input %zu + output %zu; "
                    "the staging tensors scale with the handle's max_batch)", need, CD_NOALLOC_SCRATCH_BYTES, wbytes, inbytes, outbytes);
    unsigned char* base = noalloc_scratch_ptr();
    if (!base) return 1;
    float* pk = reinterpret_cast<float*>(base);
    __half* xin = reinterpret_cast<__half*>(base + wbytes);
    __half* yout = reinterpret_cast<__half*>(base + wbytes + inbytes);
    // halo rows / guards of both staging tensors must read as zero: the array keeps the previous call's contents
    CD_CUDA(cudaMemsetAsync(xin, 0, inbytes + outbytes, st));
    const int nw = taps * Cin * Cout;
    noalloc_pack_w_kernel<<<(nw + 255) / 256, 256, 0, st>>>(w_dev, pk, Cin, Cout, taps, transposed);
    long long nrows = (long long)B * gi.rows_per_sample;
    dense_to_plcp_kernel<<<(unsigned)((nrows + 127) / 128), 128, 0, st>>>(gi, x_dev, xin, Cin, nrows);
    ConvParams p;
    p.gin = gi; p.gout = go; p.nsrc = 1; p.Cin_total = Cin;
    p.src[0].base = xin; p.src[0].C = Cin; p.src[1].base = nullptr; p.src[1].C = 0;
    p.dst = yout; p.Cout = Cout; p.w = pk; p.bias = bias_dev;
    p.kz = ksize[0]; p.kp = ksize[1]; p.kr = ksize[2];
    p.sz = stride[0]; p.sp = stride[1]; p.sr = stride[2];
    const bool is1 = taps == 1;
    p.pz = is1 ? 0 : 1; p.cp = is1 ? 0 : 1; p.pr = is1 ? 0 : 1;
    p.transposed = transposed; p.write_halo = 0; p.stats = nullptr; p.stats_cpg = 1; p.B = B;
    p.pn_stats = nullptr; p.pn_wsum = nullptr; p.pn_wb = nullptr; p.pn_C = 0; p.pn_count = 1.0;
    launch_conv_generic(p, st);
    long long nvox = (long long)B * go.Z * go.P * go.R;
    plcp_to_dense_kernel<<<(unsigned)((nvox + 127) / 128), 128, 0, st>>>(go, yout, y_dev, Cout, nvox);
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) return fail("no-alloc conv: %s", cudaGetErrorString(e));
    return 0;
}

#include "cd_train.inl"
