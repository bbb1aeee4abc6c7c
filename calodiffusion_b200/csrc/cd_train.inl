// Training step on top of the inference engine (included at the end of cd_engine.cu).
//   forward  = the engine's own plan, built with option "train" (no buffer reuse, unfused attention chain, SIMT strided convs)
//   backward = the tape recorded by the plan builder, walked in reverse with the kernels of cd_backward.cu
//   update   = fused Adam on the flat master parameters, then the conv weight layouts are re-derived on the device
// Reference: Loss.__call__ / hybrid_weight.loss_function (models/loss.py:118-142, 168-179), training_loop
// (train/train_diffusion.py:42-66), torch.optim.Adam (train/train.py:144).
namespace cd {

__global__ void noisy_kernel(const float* __restrict__ data, const float* __restrict__ noise, const float* __restrict__ sigma,
                             float* __restrict__ out, long long n, int vps) {
    const long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) out[v] = data[v] + sigma[v / vps] * noise[v];          // x_noisy = data + sigma * noise (loss.py:169)
}
// coef[b] = loss_scale * 2 * w_b / (mean(w) * numel), lossw[b] = w_b / (mean(w) * numel), w_b = 1 + 1/sigma_b^2 (loss.py:175, 103-104)
__global__ void loss_coef_kernel(const float* __restrict__ sigma, int B, double numel, float loss_scale, float* __restrict__ coef) {
    __shared__ double red[256];
    double s = 0.0;
    for (int b = threadIdx.x; b < B; b += blockDim.x) s += 1.0 + 1.0 / ((double)sigma[b] * sigma[b]);
    red[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
    const double mean = red[0] / B;
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        const double w = 1.0 + 1.0 / ((double)sigma[b] * sigma[b]);
        coef[b] = (float)(loss_scale * 2.0 * w / (mean * numel));
        coef[B + b] = (float)(w / (mean * numel));
    }
}

// Every kernel weight layout is a pure function of the flat master buffer; the list of re-layout jobs is fixed once the plan is
// built, so it lives in a device table and one launch (grid.y = job) refreshes all of them after an optimizer step.
static void train_refresh_weights(cd_unet* h, cudaStream_t st) {
    if (h->repack_tab) { launch_repack_table(h->repack_tab, h->repack_jobs, h->repack_max_n, st); return; }
    std::vector<RepackParams> jobs;
    auto launch_repack_conv = [&jobs](const RepackParams& r, cudaStream_t) { jobs.push_back(r); };
    for (const TapeOp& t : h->tape) {
        if (t.kind != 0) continue;
        const int taps = t.kz * t.kp * t.kr;
        RepackParams r;
        memset(&r, 0, sizeof(r));
        r.master = h->pptr(t.wname); r.taps = taps; r.Cin_real = t.Cin_real; r.Cin_pad = t.Cin_pad; r.Cout = t.Cout; r.transposed = t.transposed;
        r.mode = 0; r.out_f = t.w; r.n = (long long)taps * t.Cin_pad * t.Cout;
        launch_repack_conv(r, st);
        if (t.wmma) {
            r.mode = 1; r.out_h = t.wmma; r.CB = mma_pick_cb(t.Cout, taps);
            launch_repack_conv(r, st);
        }
        if (t.wgat || t.wgat_d) {
            r.pk = t.w; r.kz = t.kz; r.kp = t.kp; r.kr = t.kr;
            if (t.wgat) {       // forward: pack_conv's choice (merged pairs for the transposed conv with kr == 4)
                const bool pairT = t.transposed && t.kr == 4;
                r.mode = pairT ? 5 : 4; r.swap = 0; r.out_h = t.wgat;
                r.n = (long long)(pairT ? gather_packed_elems_pairT(t.Cin_pad, t.Cout, t.kz, t.kp) : gather_packed_elems(t.Cin_pad, t.Cout, taps));
                launch_repack_conv(r, st);
            }
            if (t.wgat_d) {     // data gradient: strided conv <-> transposed conv with the channel roles exchanged, same taps
                const bool pairT = !t.transposed;
                r.mode = pairT ? 5 : 4; r.swap = 1; r.out_h = t.wgat_d;
                r.n = (long long)(pairT ? gather_packed_elems_pairT(t.Cout, t.Cin_pad, t.kz, t.kp) : gather_packed_elems(t.Cout, t.Cin_pad, taps));
                launch_repack_conv(r, st);
            }
        }
        int lo = 0;
        for (size_t s = 0; s < t.srcs.size() && s < 2; ++s) {
            const int Cs = h->bufs[t.srcs[s]].C;
            if (t.wd[s]) {
                r.mode = 2; r.out_f = t.wd[s]; r.ci_lo = lo; r.n_out = Cs; r.n = (long long)taps * t.Cout * Cs;
                launch_repack_conv(r, st);
            }
            if (t.wdm[s]) {
                r.mode = 3; r.out_h = t.wdm[s]; r.ci_lo = lo; r.n_out = Cs; r.CB = mma_pick_cb(Cs, taps); r.n = (long long)taps * t.Cout * Cs;
                launch_repack_conv(r, st);
            }
            lo += Cs;
        }
    }
    long long mx = 1;
    for (const RepackParams& r : jobs) if (r.n > mx) mx = r.n;
    h->repack_tab = h->dalloc<RepackParams>(jobs.size() ? jobs.size() : 1, false);
    if (h->repack_tab && !jobs.empty())
        cudaMemcpyAsync(h->repack_tab, jobs.data(), jobs.size() * sizeof(RepackParams), cudaMemcpyHostToDevice, st);
    cudaStreamSynchronize(st);          // `jobs` is pageable host memory: the copy must be done before it goes out of scope
    h->repack_jobs = (int)jobs.size(); h->repack_max_n = mx;
    if (h->repack_tab) launch_repack_table(h->repack_tab, h->repack_jobs, h->repack_max_n, st);
}

static int train_setup(cd_unet* h) {
    if (h->has_geom || h->dense) return fail("training supports the plain U-Net configurations (no in-loop geometry embedding)");
    const int maxB = h->d.max_batch;
    int gmax = 1;
    for (TapeOp& t : h->tape) {
        if (t.kind == 1) {
            if (t.C > kBwdMaxC || t.C % 8) return fail("training: GroupNorm over %d channels is not supported", t.C);
            if (t.groups > gmax) gmax = t.groups;
        }
        if (t.kind != 0) continue;
        if (!h->pptr(t.wname)) return fail("training: conv weight %s has no master copy", t.wname.c_str());
        const int taps = t.kz * t.kp * t.kr;
        // strided <-> transposed conv as each other's adjoint holds on the circular axis only when the fine grid has exactly twice
        // the coarse phi bins (an odd fine P gets its last bin from output_padding / the wrap, which are not adjoint to each other)
        const int fineP = t.transposed ? h->geom[t.level_out].P : h->geom[t.level_in].P;
        const int coarseP = t.transposed ? h->geom[t.level_in].P : h->geom[t.level_out].P;
        if ((t.stride == 2 || t.transposed) && t.need_dgrad && t.srcs.size() == 1 && gather_conv_supported(t.Cout, t.Cin_pad) && taps <= 64 &&
            fineP == 2 * coarseP && (t.transposed || (t.kr == 4 && t.pr == 1 && 2 * t.Cin_pad <= 256))) {
            const size_t ne = t.transposed ? gather_packed_elems(t.Cout, t.Cin_pad, taps) : gather_packed_elems_pairT(t.Cout, t.Cin_pad, t.kz, t.kp);
            t.wgat_d = h->dalloc<__half>(ne, true);
            if (!t.wgat_d) return fail("out of device memory for the data-gradient weights");
        }
        if (t.stride == 1 && !t.transposed && t.need_dgrad) {
            for (size_t s = 0; s < t.srcs.size() && s < 2; ++s) {
                const int Cs = h->bufs[t.srcs[s]].C;
                t.wd[s] = h->dalloc<float>((size_t)taps * t.Cout * Cs, true);
                if (mma_conv_supported(t.Cout, Cs, taps)) t.wdm[s] = h->dalloc<__half>((size_t)taps * t.Cout * Cs, true);
                if (!t.wd[s]) return fail("out of device memory for the data-gradient weights");
            }
        }
    }
    const int V = h->voxels();
    h->gn_acc = h->dalloc<float>((size_t)maxB * kBwdMaxC * 4, true);
    h->gn_m12 = h->dalloc<float>((size_t)maxB * kBwdMaxC * 4, true);     // m1/m2 per group, then mean/rstd per channel
    h->attn_dcx = h->dalloc<float>((size_t)maxB * 1024, true);
    h->daddvec = h->dalloc<float>((size_t)maxB * (h->addvec_stride > 0 ? h->addvec_stride : 1), true);
    h->x_noisy = h->dalloc<float>((size_t)maxB * V, true);
    h->x0_buf = h->dalloc<float>((size_t)maxB * V, true);
    h->coef = h->dalloc<float>((size_t)maxB * 2, true);
    h->loss_dev = h->dalloc<double>(1, true);
    // per-sample record of the conditioning MLPs' (input, output-gradient) pairs
    const EmbedParams& e = h->embed;
    const int half = e.cond_dim / 2, h2 = half / 2;
    EmbedBwdLayout& L = h->embed_layout;
    int o = 0;
    auto take = [&o](int n) { int r = o; o += (n + 3) / 4 * 4; return r; };
    L.cvec = take(e.cond_dim); L.dt3 = take(half); L.t2 = take(half); L.dc3 = take(half); L.c2 = take(half); L.dt2 = take(half);
    L.t1 = take(h2); L.dt1 = take(h2); L.t0 = take(1); L.dc2 = take(half); L.c1 = take(e.cond_hidden); L.dc1 = take(e.cond_hidden);
    L.c0 = take(e.cond_size);
    L.stride = o;
    h->embed_rec = h->dalloc<float>((size_t)maxB * o, true);
    if (!h->gn_acc || !h->gn_m12 || !h->attn_dcx || !h->daddvec || !h->x_noisy || !h->x0_buf || !h->coef || !h->loss_dev || !h->embed_rec)
        return fail("out of device memory for the training scratch");
    if (e.cond_dim > 256 || half > 128 || e.cond_hidden > 256 || e.cond_size > 256) return fail("training: conditioning MLP too wide");
    train_refresh_weights(h, 0);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) return fail("training setup: %s", cudaGetErrorString(err));
    return 0;
}

// data gradient of one stride-1 convolution into grad[src] (accumulating when that gradient already has contributions)
static int conv_dgrad(cd_unet* h, const TapeOp& t, int s, int B, bool accumulate, __half* tmp, cudaStream_t st) {
    const Geom& g = h->geom[t.level_in];
    const int taps = t.kz * t.kp * t.kr, Cs = h->bufs[t.srcs[s]].C;
    ConvParams p;
    p.gin = g; p.gout = g; p.nsrc = 1; p.Cin_total = t.Cout;
    p.src[0].base = h->bufs[t.dst].grad; p.src[0].C = t.Cout; p.src[1].base = nullptr; p.src[1].C = 0;
    p.Cout = Cs; p.w = t.wd[s]; p.bias = nullptr;
    p.kz = t.kz; p.kp = t.kp; p.kr = t.kr; p.sz = p.sp = p.sr = 1; p.pz = t.pz; p.cp = t.cp; p.pr = t.pr;
    p.transposed = 0; p.write_halo = 0; p.stats = nullptr; p.stats_cpg = 1;
    p.pn_stats = nullptr; p.pn_wsum = nullptr; p.pn_wb = nullptr; p.pn_C = 0; p.pn_count = 1.0; p.B = B;
    __half* out = h->bufs[t.srcs[s]].grad;
    const bool k3 = taps == 27;
    const bool fused_acc = accumulate && k3 && t.wdm[s];
    p.dst = (accumulate && !fused_acc) ? tmp : out;
    p.addsrc = fused_acc ? out : nullptr;
    bool done = false;
    if (t.wdm[s] && h->conv_backend == 0) done = launch_conv_mma(p, t.wdm[s], st) == 0;
    if (!done) {
        if (fused_acc) { p.addsrc = nullptr; p.dst = tmp; }
        launch_conv_generic(p, st);
        if (fused_acc) { launch_plcp_accumulate(g, tmp, out, Cs, 1, B, st); return 0; }
    }
    if (accumulate && !fused_acc) launch_plcp_accumulate(g, tmp, out, Cs, 1, B, st);
    return 0;
}

static int run_backward(cd_unet* h, int B, const float* data, const float* sigma, const float* cond, cudaStream_t st) {
    std::vector<char> ginit(h->bufs.size(), 0);
    launch_fill_zero(h->G, h->n_params, st);
    launch_fill_zero(h->daddvec, (long long)B * (h->addvec_stride > 0 ? h->addvec_stride : 1), st);
    cudaMemsetAsync(h->loss_dev, 0, sizeof(double), st);
    // scratch gradient tensor for data gradients that have to be accumulated outside the conv epilogue: largest (level, C)
    std::vector<__half*>& tmp = h->bwd_tmp;
    if (tmp.empty()) {
        tmp.resize(h->n_levels, nullptr);
        for (int l = 0; l < h->n_levels; ++l) tmp[l] = h->dalloc<__half>((size_t)(kBwdMaxC / 8) * (size_t)h->geom[l].chunk_stride * 8, true);
    }
    // diagnostic per-op timing of the backward pass (option "profile"): events between the ops, read back at the end
    struct Mark { std::string name; double flops; };
    std::vector<Mark> marks;
    size_t nev = 0;
    auto rec = [&]() {
        if (h->prof_evs.size() <= nev) { cudaEvent_t ev; cudaEventCreate(&ev); h->prof_evs.push_back(ev); }
        cudaEventRecord(h->prof_evs[nev++], st);
    };
    auto mark = [&](const char* what, const TapeOp* t, int lvl, int c0, int c1, double flops) {
        if (!h->profile) return;
        char nm[112];
        if (t && t->kind == 0) snprintf(nm, sizeof(nm), "bwd_%s_k%d%s%s_c%d_%d_L%d", what, t->kz * t->kp * t->kr, t->stride > 1 ? "s2" : "", t->transposed ? "T" : "", c0, c1, lvl);
        else snprintf(nm, sizeof(nm), "bwd_%s_c%d_L%d", what, c0, lvl);
        marks.push_back({nm, flops});
        rec();
    };
    if (h->profile) rec();
    for (int i = (int)h->tape.size() - 1; i >= 0; --i) {
        const TapeOp& t = h->tape[i];
        if (t.kind == 3) {
            LossBwdParams p;
            p.g = h->geom[0]; p.h = h->bufs[t.src].ptr; p.C = t.C; p.w = h->final_w; p.bias = h->pptr("model.final_conv.1.conv.bias");
            p.x = h->x_noisy; p.data = data; p.sigma = sigma; p.sigma_data = h->d.sigma_data; p.coef = h->coef; p.lossw = h->coef + B;
            p.dh = h->bufs[t.src].grad; p.dw = h->gptr("model.final_conv.1.conv.weight"); p.db = h->gptr("model.final_conv.1.conv.bias");
            p.loss = h->loss_dev; p.x0_out = nullptr;
            if (!p.dw || !p.db || !p.bias) return fail("training: final conv parameters missing");
            launch_loss_backward(p, B, st);
            ginit[t.src] = 1;
            mark("loss", nullptr, 0, t.C, 0, 0.0);
        } else if (t.kind == 1) {
            if (!ginit[t.dst]) return fail("training: GroupNorm output %d has no gradient (tape order)", t.dst);
            const Geom& g = h->geom[t.level_in];
            GnBwdParams p;
            memset(&p, 0, sizeof(p));
            p.g = g; p.x = h->bufs[t.src].ptr; p.dy = h->bufs[t.dst].grad;
            p.stats = (const double*)(h->scratch + t.stats_off * B); p.gamma = h->pptr(t.gname); p.beta = h->pptr(t.betaname);
            p.C = t.C; p.groups = t.groups; p.act = t.act; p.count_per_channel = (double)g.Z * g.P * g.R;
            p.acc = h->gn_acc; p.m12 = h->gn_m12; p.mr = h->gn_m12 + (size_t)h->d.max_batch * kBwdMaxC * 2; p.dgamma = h->gptr(t.gname); p.dbeta = h->gptr(t.betaname);
            if (t.conv_of_src >= 0 && !h->tape[t.conv_of_src].bname.empty()) p.dbias = h->gptr(h->tape[t.conv_of_src].bname);
            if (t.addvec_off >= 0) { p.daddvec = h->daddvec + t.addvec_off; p.addvec_stride = h->addvec_stride; }
            p.dx = h->bufs[t.src].grad; p.dx_accumulate = ginit[t.src];
            if (t.residual >= 0) {
                p.dres = h->bufs[t.residual].grad; p.dres_accumulate = ginit[t.residual];
                p.res_split = h->bufs[t.residual].C / 8;
                if (t.residual2 >= 0) { p.dres2 = h->bufs[t.residual2].grad; p.dres2_accumulate = ginit[t.residual2]; }
                else p.res_split = t.C / 8;
            }
            if (!p.gamma || !p.beta || !p.dgamma || !p.dbeta) return fail("training: GroupNorm parameters %s missing", t.gname.c_str());
            launch_gn_backward(p, B, st);
            mark(t.residual >= 0 ? "gn_res" : "gn", nullptr, t.level_in, t.C, 0, 0.0);
            ginit[t.src] = 1;
            if (t.residual >= 0) ginit[t.residual] = 1;
            if (t.residual2 >= 0) ginit[t.residual2] = 1;
        } else if (t.kind == 2) {
            if (!ginit[t.y]) return fail("training: attention output has no gradient");
            AttnBwdParams p;
            memset(&p, 0, sizeof(p));
            p.g = h->geom[t.level_in]; p.qkv = h->bufs[t.qkv].ptr; p.dy = h->bufs[t.y].grad;
            p.kmax = (const unsigned*)(h->scratch + t.skm * B); p.ctx = (const float*)(h->scratch + t.sctx * B);
            p.wout = h->pptr(t.woname); p.C = t.C; p.dcx = h->attn_dcx; p.dwout = h->gptr(t.woname); p.dbout = h->gptr(t.boname);
            p.dqkv = h->bufs[t.qkv].grad;
            if (!p.wout || !p.dwout || !p.dbout) return fail("training: attention parameters missing");
            if (launch_attn_backward(p, B, st)) return fail("training: attention backward does not support %d channels", t.C);
            ginit[t.qkv] = 1;
            mark("attn", nullptr, t.level_in, t.C, 0, 0.0);
        } else if (t.kind == 0) {
            if (!ginit[t.dst]) return fail("training: conv output %d (%s) has no gradient (tape order)", t.dst, t.wname.c_str());
            const Geom& gi = h->geom[t.level_in];
            const Geom& go = h->geom[t.level_out];
            const int taps = t.kz * t.kp * t.kr;
            float* dw = h->gptr(t.wname);
            if (!dw) return fail("training: %s has no gradient slot", t.wname.c_str());
            const double mo = (double)go.Z * go.P * go.R, mi = (double)gi.Z * gi.P * gi.R;
            const double cflops = 2.0 * (t.transposed ? mi : mo) * t.Cin_pad * t.Cout * taps * B;     // executed MACs x 2
            if (!t.bias_via_gn && !t.bname.empty()) {
                launch_channel_sum(go, h->bufs[t.dst].grad, t.Cout, h->gptr(t.bname), B, st);
                mark("bias", nullptr, t.level_out, t.Cout, 0, 0.0);
            }
            ConvParams p;       // the forward convolution, for the SIMT adjoints
            p.gin = gi; p.gout = go; p.nsrc = (int)t.srcs.size(); p.Cin_total = t.Cin_pad;
            for (int s = 0; s < p.nsrc; ++s) { p.src[s].base = h->bufs[t.srcs[s]].ptr; p.src[s].C = h->bufs[t.srcs[s]].C; }
            if (p.nsrc == 1) { p.src[1].base = nullptr; p.src[1].C = 0; }
            p.dst = nullptr; p.Cout = t.Cout; p.w = t.w; p.bias = nullptr; p.kz = t.kz; p.kp = t.kp; p.kr = t.kr;
            p.sz = p.sp = p.sr = t.stride; p.pz = t.pz; p.cp = t.cp; p.pr = t.pr; p.transposed = t.transposed; p.write_halo = 0;
            p.stats = nullptr; p.stats_cpg = 1; p.pn_stats = nullptr; p.pn_wsum = nullptr; p.pn_wb = nullptr; p.pn_C = 0; p.pn_count = 1.0; p.B = B;
            const bool unit = t.stride == 1 && !t.transposed;
            bool wdone = false;
            if (unit && (taps == 27 || taps == 1)) {
                WgradParams w;
                memset(&w, 0, sizeof(w));
                w.g = gi; w.x0 = h->bufs[t.srcs[0]].ptr; w.x1 = t.srcs.size() > 1 ? h->bufs[t.srcs[1]].ptr : nullptr;
                w.src_chunks0 = h->bufs[t.srcs[0]].C / 8; w.dy = h->bufs[t.dst].grad; w.dw = dw;
                w.Cin_pad = t.Cin_pad; w.Cin_real = t.Cin_real; w.Cout = t.Cout; w.taps = taps; w.B = B;
                wdone = launch_wgrad_mma(w, st) == 0;
                if (wdone) mark("wgrad_mma", &t, t.level_in, t.Cin_pad, t.Cout, cflops);
            }
            if (!wdone) {
                if (t.srcs.size() != 1) return fail("training: weight gradient of %s (two sources, %d -> %d) has no kernel", t.wname.c_str(), t.Cin_pad, t.Cout);
                launch_conv_strided_backward(p, h->bufs[t.srcs[0]].ptr, h->bufs[t.dst].grad, nullptr, 0, dw, t.Cin_real, t.transposed, st);
                mark("wgrad_simt", &t, t.level_in, t.Cin_pad, t.Cout, cflops);
            }
            if (t.need_dgrad) {
                if (unit) {
                    for (size_t s = 0; s < t.srcs.size(); ++s) {
                        if (conv_dgrad(h, t, (int)s, B, ginit[t.srcs[s]] != 0, tmp[t.level_in], st)) return 1;
                        ginit[t.srcs[s]] = 1;
                        const int Cs = h->bufs[t.srcs[s]].C;
                        mark(t.wdm[s] && h->conv_backend == 0 ? "dgrad_mma" : "dgrad_simt", &t, t.level_in, t.Cout, Cs, cflops * Cs / t.Cin_pad);
                    }
                } else {
                    bool ddone = false;
                    if (t.wgat_d && h->conv_backend == 0) {
                        // the adjoint of a strided conv is the transposed conv with the same taps (and vice versa): the gather
                        // tcgen05 kernel on dY, whose halo rows are re-derived first (its producer may have left them stale)
                        launch_plcp_accumulate(go, h->bufs[t.dst].grad, h->bufs[t.dst].grad, t.Cout, 0, B, st);
                        ConvParams q = p;
                        q.gin = go; q.gout = gi; q.nsrc = 1; q.Cin_total = t.Cout;
                        q.src[0].base = h->bufs[t.dst].grad; q.src[0].C = t.Cout; q.src[1].base = nullptr; q.src[1].C = 0;
                        q.Cout = t.Cin_pad; q.w = nullptr; q.bias = nullptr; q.transposed = !t.transposed; q.write_halo = 1;
                        q.dst = ginit[t.srcs[0]] ? tmp[t.level_in] : h->bufs[t.srcs[0]].grad;
                        ddone = launch_conv_gather(q, t.wgat_d, st) == 0;
                        if (ddone && ginit[t.srcs[0]]) launch_plcp_accumulate(gi, tmp[t.level_in], h->bufs[t.srcs[0]].grad, t.Cin_pad, 1, B, st);
                    }
                    if (!ddone)
                        launch_conv_strided_backward(p, nullptr, h->bufs[t.dst].grad, h->bufs[t.srcs[0]].grad, ginit[t.srcs[0]], nullptr, t.Cin_real,
                                                     t.transposed, st);
                    ginit[t.srcs[0]] = 1;
                    mark(ddone ? "dgrad_gather" : "dgrad_simt", &t, t.level_in, t.Cout, t.Cin_pad, cflops);
                }
            }
        }
    }
    // conditioning MLPs
    const EmbedParams& e = h->embed;
    const EmbedBwdLayout& L = h->embed_layout;
    launch_embed_backward(e, L, sigma, cond, h->daddvec, h->embed_rec, B, st);
    if (!h->linear_tab) {       // the jobs are fixed once the plan is built (all pointers are the handle's own buffers)
        const int half = e.cond_dim / 2, h2 = half / 2;
        const std::string M = "model.";
        struct Lin { const char* name; int dy, x, n_out, n_in; };
        const Lin lins[6] = {{"time_mlp.5", L.dt3, L.t2, half, half}, {"time_mlp.3", L.dt2, L.t1, half, h2}, {"time_mlp.1", L.dt1, L.t0, h2, 1},
                             {"cond_mlp.4", L.dc3, L.c2, half, half}, {"cond_mlp.2", L.dc2, L.c1, half, e.cond_hidden},
                             {"cond_mlp.0", L.dc1, L.c0, e.cond_hidden, e.cond_size}};
        std::vector<LinearJob> jobs;
        for (const Lin& l : lins) {
            float* dW = h->gptr(M + l.name + ".weight");
            float* db = h->gptr(M + l.name + ".bias");
            if (!dW || !db) return fail("training: %s has no gradient slot", l.name);
            jobs.push_back({h->embed_rec + l.dy, h->embed_rec + l.x, dW, db, (long long)L.stride, (long long)L.stride, l.n_out, l.n_in});
        }
        for (int k = 0; k < e.n_proj; ++k) {
            float* dW = h->gptr(h->proj_names[k] + ".weight");
            float* db = h->gptr(h->proj_names[k] + ".bias");
            if (!dW || !db) return fail("training: projection %d has no gradient slot", k);
            jobs.push_back({h->daddvec + e.poff[k], h->embed_rec + L.cvec, dW, db, (long long)h->addvec_stride, (long long)L.stride, e.pC[k], e.cond_dim});
        }
        int mx = 1;
        for (const LinearJob& j : jobs) mx = std::max(mx, j.n_out * (j.n_in + 1));
        h->linear_tab = h->dalloc<LinearJob>(jobs.size(), false);
        if (!h->linear_tab) return fail("out of device memory for the training tables");
        cudaMemcpyAsync(h->linear_tab, jobs.data(), jobs.size() * sizeof(LinearJob), cudaMemcpyHostToDevice, st);
        cudaStreamSynchronize(st);
        h->linear_jobs = (int)jobs.size(); h->linear_max_n = mx;
    }
    launch_linear_wgrad_table(h->linear_tab, h->linear_jobs, h->linear_max_n, B, st);
    mark("embed_mlps", nullptr, 0, e.cond_dim, 0, 0.0);
    if (h->profile && nev > 1) {
        cudaEventSynchronize(h->prof_evs[nev - 1]);
        for (size_t k = 0; k < marks.size(); ++k) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, h->prof_evs[k], h->prof_evs[k + 1]);
            ProfAcc& a = h->prof[marks[k].name];
            a.count += 1; a.ms += ms; a.flops += marks[k].flops;
        }
    }
    cudaError_t err = cudaPeekAtLastError();
    if (err != cudaSuccess) return fail("training backward: %s", cudaGetErrorString(err));
    return 0;
}

}  // namespace cd

extern "C" {

int cd_train_step(cd_unet* h, const float* data_dev, const float* noise_dev, const float* sigma_dev, const float* cond_dev,
                  float loss_scale, int B, void* stream) {
    if (check_call(h, B)) return 1;
    if (!h->train) return fail("the handle was not finalized with option train");
    if (!data_dev || !noise_dev || !sigma_dev || !cond_dev) return fail("null argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int V = h->voxels();
    const long long n = (long long)B * V;
    noisy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(data_dev, noise_dev, sigma_dev, h->x_noisy, n, V);
    loss_coef_kernel<<<1, 256, 0, st>>>(sigma_dev, B, (double)n, loss_scale, h->coef);
    Run r;
    memset(&r.fin, 0, sizeof(r.fin));
    r.B = B; r.st = st; r.x = h->x_noisy; r.sigma = sigma_dev; r.sigma_stride = 1; r.cond = cond_dev; r.scale_input = 1; r.time_is_raw = 0;
    r.fin.x0_out = h->x0_buf;
    if (run_forward(h, r)) return 1;
    return run_backward(h, B, data_dev, sigma_dev, cond_dev, st);
}

int cd_train_loss(cd_unet* h, double* loss_host, void* stream) {
    if (!h || !h->train || !loss_host) return fail("bad argument");
    CD_CUDA(cudaMemcpyAsync(loss_host, h->loss_dev, sizeof(double), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CD_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}

int cd_train_adam(cd_unet* h, float lr, float beta1, float beta2, float eps, float loss_scale, void* stream) {
    if (!h || !h->train) return fail("the handle was not finalized with option train");
    if (!(loss_scale > 0.f)) return fail("loss_scale must be positive");
    h->adam_step += 1;
    launch_adam(h->P, h->M, h->V, h->G, h->n_params, lr, beta1, beta2, eps, h->adam_step, 1.f / loss_scale, (cudaStream_t)stream);
    train_refresh_weights(h, (cudaStream_t)stream);
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) return fail("adam: %s", cudaGetErrorString(e));
    return 0;
}

int cd_train_refresh(cd_unet* h, void* stream) {
    if (!h || !h->train) return fail("the handle was not finalized with option train");
    train_refresh_weights(h, (cudaStream_t)stream);
    return 0;
}

int64_t cd_train_flat_size(const cd_unet* h) { return h && h->train ? h->n_params : 0; }
int cd_train_num_params(const cd_unet* h) { return h && h->train ? (int)h->params.size() : 0; }
int cd_train_param_info(const cd_unet* h, int index, char* name, int cap, int64_t* offset, int64_t* numel) {
    if (!h || !h->train || index < 0 || index >= (int)h->params.size() || !name || cap < 1) return fail("bad argument");
    snprintf(name, (size_t)cap, "%s", h->params[index].name.c_str());
    if (offset) *offset = h->params[index].offset;
    if (numel) *numel = h->params[index].numel;
    return 0;
}
void* cd_train_buffer(cd_unet* h, int which) {
    if (!h || !h->train) return nullptr;
    return which == 0 ? (void*)h->P : which == 1 ? (void*)h->G : which == 2 ? (void*)h->M : which == 3 ? (void*)h->V
         : which == 4 ? (void*)h->x0_buf : nullptr;
}

}  // extern "C"
