#!/usr/bin/env python
"""Benchmark of the CaloDiffusion sampling hot path on B200 (contract: see DESIGN.md "Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload ds2|ds3] [--impl reference]

A "step" is one pass of the hot path over one batch: the whole reverse-diffusion loop (BASELINE config 2:
Dataset-2, DDPM, 400 denoiser evaluations; config 3: Dataset-3, Heun) for B showers per GPU, replayed as one
CUDA graph.  `value` = showers/s with inputs resident in HBM; `e2e` = showers/s through the public API
(`CaloDiffusion.sample`: pinned-host conditioning -> device, showers -> host numpy) inside the timed region.
Synthetic conditioning/noise, random-init weights (no datasets or checkpoints offline).
Under torchrun (N > 1) every rank samples its own B showers (weak scaling, no data-path collective) and one
NCCL all_gather collects the showers at the end of each step, as SURVEY 8(e) prescribes.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# hot-path keys of the reference's shipped configs (configs/config_dataset{2,3}.json)
_COMMON = dict(COND_SIZE_UNET=128, BLOCK_ATTN=True, MID_ATTN=True, COMPRESS_Z=True, CYLINDRICAL=True, R_Z_INPUT=True,
               TRAINING_OBJ="hybrid_weight", LOSS_TYPE="l2", COND_EMBED="id", SHOWER_EMBED="", NSTEPS=400)
WORKLOADS = {
    "ds2": dict(
        cfg=dict(_COMMON, DATASET_NUM=2, SHAPE_ORIG=[-1, 6480], SHAPE_PAD=[-1, 1, 45, 16, 9],
                 SHAPE_FINAL=[-1, 1, 45, 16, 9], LAYER_SIZE_UNET=[32, 32, 64, 32], SHOWERMAP="layer-logit-norm",
                 NOISE_SCHED="log", TIME_EMBED="log", SAMPLER="DDPM", BATCH=128),
        name="CaloChallenge Dataset-2 (45x16x9), cylindrical U-Net, DDPM 400 steps", sampler="DDPM", steps=400,
        evals_per_shower=400, batch=1000, n_layers=46, gflop_per_eval=5.4439),
    "ds3": dict(
        cfg=dict(_COMMON, DATASET_NUM=3, SHAPE_ORIG=[-1, 40500], SHAPE_PAD=[-1, 1, 45, 50, 18],
                 SHAPE_FINAL=[-1, 1, 45, 50, 18], LAYER_SIZE_UNET=[32, 32, 32, 32], SHOWERMAP="logit-norm",
                 NOISE_SCHED="cosine", TIME_EMBED="sigma", SAMPLER="Heun", BATCH=64),
        name="CaloChallenge Dataset-3 (45x50x18), cylindrical U-Net, EDM Heun 50 steps", sampler="Heun", steps=50,
        evals_per_shower=100, batch=256, n_layers=0, gflop_per_eval=30.351),
    # BASELINE config 4: the HGCal pytest-fixture shape (tests/test_execution.py:178-190), pre-embedded showers
    "hgcal": dict(
        cfg=dict(_COMMON, DATASET_NUM=111, HGCAL=True, SHAPE_ORIG=[-1, 28, 1988], SHAPE_PAD=[-1, 1, 28, 12, 21],
                 SHAPE_FINAL=[-1, 1, 28, 12, 21], LAYER_SIZE_UNET=[32, 32, 64, 96], SHOWERMAP="layer-logit-norm",
                 SHOWER_EMBED="NN-pre-embed", PHI_INPUT=True, NOISE_SCHED="log", TIME_EMBED="log", SAMPLER="DDim", BATCH=128,
                 MAX_CELLS=1988),
        name="HGCal test-fixture shape (28x12x21, pre-embedded), cylindrical U-Net, DDIM 200 steps", sampler="DDim", steps=200,
        evals_per_shower=200, batch=128, n_layers=29, gflop_per_eval=6.4713, cond_e=3),
    # BASELINE config 1: Dataset-1 photons, 368 irregular voxels embedded to a (5,10,30) grid inside every evaluation
    "ds1": dict(
        cfg=dict(_COMMON, DATASET_NUM=1, PART_TYPE="photon", SHAPE_ORIG=[-1, 368], SHAPE_PAD=[-1, 1, 368],
                 SHAPE_FINAL=[-1, 1, 5, 10, 30], LAYER_SIZE_UNET=[32, 32, 64, 96], SHOWERMAP="layer-logit-norm",
                 SHOWER_EMBED="orig-NN", PHI_INPUT=True, NOISE_SCHED="log", TIME_EMBED="log", SAMPLER="DDim", BATCH=128),
        name="CaloChallenge Dataset-1 photons (368 voxels, in-loop GeoEmbed to 5x10x30), cylindrical U-Net, DDIM 50 steps",
        sampler="DDim", steps=50, evals_per_shower=50, batch=1000, n_layers=6, gflop_per_eval=1.4708, binning="ds1_meta.json"),
}
# BASELINE config 5: the Dataset-2 training step (train/train_diffusion.py:42-66), 128 showers per GPU, throughput only
WORKLOADS["ds2_train"] = dict(WORKLOADS["ds2"], train=True, batch=128, ref_key="ds2",
                              name="CaloChallenge Dataset-2 (45x16x9) training step: noise -> denoise -> hybrid-weight l2 -> backward -> Adam")
TRAIN_LR = 4e-4      # configs/config_dataset2.json "LR"


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=float(p["hbm_gbs"]), tf=float(p["bf16_tflops"]), tf_sus=float(p["bf16_tflops_sustained"]),
                    src="MEASURED_PEAKS.json")
    except Exception:
        return dict(hbm=6650.0, tf=1590.0, tf_sus=1400.0, src="fallback (B200_PROFILING.md)")


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [c.strip() for c in out.strip().split(",")]
                if len(f) >= 6:
                    self.rows.append(f)
            except Exception:
                pass
            self._halt.wait(0.2)

    def finish(self):
        self._halt.set()
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(int(float(r[0])) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": int(float(self.rows[0][1])), "reasons": reasons,
                "samples": len(sm)}


def _data_shape(cfg):
    """models/diffusion.py:40-43: the caller-facing shower shape."""
    return list(cfg["SHAPE_ORIG"][1:] if "orig" in cfg.get("SHOWER_EMBED", "") else cfg["SHAPE_PAD"][1:])


def _binning(wl, cfg):
    """Dataset-1 binning table (tests/golden/ds1_meta.json, written from the reference's parsed XML) -> temp BIN_FILE."""
    import json
    import tempfile
    from calodiffusion_b200.geometry import write_binning_xml
    with open(os.path.join(ROOT, "tests", "golden", wl["binning"])) as f:
        meta = json.load(f)
    path = os.path.join(tempfile.mkdtemp(prefix="cd_bins_"), "binning.xml")
    cfg["BIN_FILE"] = write_binning_xml(path, meta["xml_r_edges"], meta["xml_a_bins"])
    return meta


HGCAL_REF_OVERRIDES = {  # the reference's pytest hgcal fixture (tests/test_execution.py:178-190)
    "HGCAL": True, "SHAPE_ORIG": [-1, 28, 1988], "DATASET_NUM": 111, "SHAPE_PAD": [-1, 1, 28, 12, 21],
    "SHAPE_FINAL": [-1, 1, 28, 12, 21], "MAX_CELLS": 1988, "LAYER_SIZE_UNET": [32, 32, 64, 96],
    "SHOWER_EMBED": "NN-pre-embed",
}
REF_CONFIGS = {"ds2": ("config_dataset2.json", None), "ds3": ("config_dataset3.json", None),
               "hgcal": ("config_HGCal.json", HGCAL_REF_OVERRIDES),
               "ds1": ("config_dataset1_photon.json", {"BIN_FILE": "{REF_ROOT}/CaloChallenge/code/binning_dataset_1_photons.xml"})}


def config_block(wl, B, steps, world, strong=False):
    """The `config` object of the JSON line: identical on both arms (the driver compares them)."""
    if wl.get("train"):
        return {"workload": wl["name"], "batch_per_gpu": B, "optimizer": f"Adam lr {TRAIN_LR}",
                "parallelism": f"dp{world} (replicated weights, one all-reduce of the flat fp32 gradient per step)",
                "l2": "activations saved for the backward pass (GBs) exceed the 126 MB L2; no explicit flush",
                "weights": "random init (seed 0)"}
    evals = wl["evals_per_shower"] * steps / wl["steps"]
    if strong:      # B = the whole job's showers, split over the GPUs
        return {"workload": wl["name"], "batch_total": B, "sampler": wl["sampler"], "sampling_steps": steps,
                "denoiser_evals_per_shower": evals, "parallelism": f"dp{world} (event-batch shards of ONE batch, final all_gather)",
                "l2": "working set per evaluation (GBs of activations) exceeds the 126 MB L2; no explicit flush",
                "weights": "random init (seed 0)"}
    cfg = {"workload": wl["name"], "batch_per_gpu": B, "sampler": wl["sampler"], "sampling_steps": steps,
           "denoiser_evals_per_shower": evals, "parallelism": f"dp{world} (event-batch shards, final all_gather)",
           "l2": "working set per evaluation (GBs of activations) exceeds the 126 MB L2; no explicit flush",
           "weights": "random init (seed 0)"}
    if wl["sampler"] == "Heun":   # same key on both arms (the driver compares the blocks)
        cfg["sampler_note"] = ("the B200 arm takes the last step (t_next = 0) as an Euler step (SAMPLER_OPTIONS LAST_STEP_EULER: finite "
                               "showers, 2*steps - 1 evaluations); the unmodified reference divides by zero there (2*steps evaluations)")
    return cfg


def reference_leg(key, budget_s=10.0, batch=None):
    """The UNMODIFIED reference (`oracle/_ref`, staged by oracle/build_ref.sh) timed through its own public API,
    `CaloDiffusion.sample` (models/diffusion.py:77-104) with its own sampler classes, on whatever device its
    `utils.get_device()` picks (the host cores when CUDA is hidden, else the GPU with torch/cuDNN eager defaults),
    on a bounded sample of the workload: the last n steps of the real schedule (`sample_offset`) at a reduced batch.
    Returns dict(value showers/s, cores, kind, sample, device, ms_per_eval)."""
    import torch
    from oracle import ref_shim
    wl = WORKLOADS[key]
    if wl.get("train"):
        return reference_train_leg(key, budget_s=budget_s, batch=batch)
    if not ref_shim.available():
        v, cores, sample, _ = port_leg(wl, budget_s=budget_s)
        return dict(value=v, cores=cores, kind="port", sample=sample, device="cpu", ms_per_eval=None)
    cfg_name, overrides = REF_CONFIGS[key]
    overrides = {k: (v.format(REF_ROOT=ref_shim.REF_ROOT) if isinstance(v, str) else v) for k, v in (overrides or {}).items()}
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model, cfg = ref_shim.load_reference_model(cfg_name, wl["sampler"], overrides)
    on_gpu = model.device.type == "cuda"
    B = int(batch or (wl["batch"] if on_gpu else 16))
    g = torch.Generator().manual_seed(1)
    E = torch.rand(B, wl.get("cond_e", 1), generator=g).to(model.device)
    layers = torch.randn(B, wl["n_layers"], generator=g).to(model.device) if wl["n_layers"] else None
    per_step = 2 if wl["sampler"] == "Heun" else 1      # denoiser evaluations per sampler step

    def run(n):   # n sampler steps at the tail of the real schedule
        t0 = time.perf_counter()
        out = model.sample(E, layers, num_steps=wl["steps"], sample_offset=wl["steps"] - n)
        if on_gpu:
            torch.cuda.synchronize()
        assert out.shape[0] == B
        return time.perf_counter() - t0

    run(1)                                                # warm-up (thread pool / cuDNN autotune / allocator)
    evals_done, t_total, n = 0, 0.0, 1
    while t_total < budget_s and evals_done < 4096:
        t_total += run(n)
        evals_done += n * per_step
        n = min(n * 2, 16 if on_gpu else 8)
    value = B * evals_done / t_total / wl["evals_per_shower"]
    sample = (f"unmodified reference CaloDiffusion.sample + {wl['sampler']} on {model.device.type}: {evals_done} denoiser "
              f"evaluations (tail of the {wl['steps']}-step schedule via sample_offset) at batch {B} in {t_total:.1f} s, "
              f"scaled to {wl['evals_per_shower']} evaluations per shower")
    return dict(value=value, cores=cores, kind="reference", sample=sample, device=model.device.type,
                ms_per_eval=1e3 * t_total / evals_done, batch=B)


def reference_train_leg(key, budget_s=10.0, batch=None):
    """The UNMODIFIED reference's training-loop body (train/train_diffusion.py:42-66: zero_grad, randn noise, compute_loss,
    backward, torch.optim.Adam.step of train/train.py:144) on synthetic showers, on whatever device the reference picks.
    Bounded sample: whole optimizer steps at a reduced batch on the CPU (16), the real batch on a GPU."""
    import torch
    from oracle import ref_shim
    wl = WORKLOADS[key]
    if not ref_shim.available():
        raise RuntimeError("oracle/_ref is not staged (run oracle/build_ref.sh in the build container)")
    cfg_name, overrides = REF_CONFIGS[wl["ref_key"]]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model, cfg = ref_shim.load_reference_model(cfg_name, wl["sampler"], overrides)
    model.train()
    on_gpu = model.device.type == "cuda"
    B = int(batch or (wl["batch"] if on_gpu else 16))
    opt = torch.optim.Adam(model.parameters(), lr=float(cfg.get("LR", TRAIN_LR)))
    g = torch.Generator().manual_seed(1)
    E = torch.rand(B, 1, generator=g).to(model.device)
    layers = torch.randn(B, wl["n_layers"], generator=g).to(model.device)
    data = torch.randn([B] + cfg["SHAPE_PAD"][1:], generator=g).to(model.device)

    def step():
        model.zero_grad()
        opt.zero_grad()
        t = torch.randint(0, model.nsteps, (B,), device=model.device).long()
        noise = torch.randn_like(data)
        loss = model.compute_loss(data=data, energy=E, noise=noise, layers=layers, time=t)
        loss.backward()
        opt.step()
        return loss.item()          # the loop's own host read (train_loss += batch_loss.item())

    step()
    n, t_total = 0, 0.0
    while t_total < budget_s and n < 4096:
        t0 = time.perf_counter()
        step()
        if on_gpu:
            torch.cuda.synchronize()
        t_total += time.perf_counter() - t0
        n += 1
    sample = (f"unmodified reference training-loop body (compute_loss -> backward -> torch.optim.Adam) on {model.device.type}: "
              f"{n} optimizer steps at batch {B} in {t_total:.1f} s")
    return dict(value=B * n / t_total, cores=cores, kind="reference", sample=sample, device=model.device.type,
                ms_per_eval=1e3 * t_total / n, batch=B)


def reference_leg_subprocess(key, device, budget_s=10.0, timeout=900):
    """Run `reference_leg` in a fresh interpreter (CUDA hidden for the CPU leg) and parse its JSON line."""
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    if device == "cpu":
        env["CUDA_VISIBLE_DEVICES"] = ""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference-leg", "--workload", key, "--budget", str(budget_s)]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)
    for ln in reversed(out.stdout.strip().splitlines()):
        if ln.startswith("{"):
            return json.loads(ln)
    raise RuntimeError("reference leg failed: " + out.stderr[-2000:])


def port_leg(wl, budget_s=12.0, batch=16):
    """Fallback when oracle/_ref is absent: the oracle port (plain torch CPU fp32 ops, all host threads) on a
    bounded sample of the same workload.  Returns (showers_per_s, cores, sample description, seconds)."""
    import torch
    from oracle.unet_oracle import DenoiserOracle, sample_ddim, sample_edm, synth_state_dict
    from calodiffusion_b200.calodiffusion import unet_param_shapes
    cfg = wl["cfg"]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    zpr = cfg["SHAPE_FINAL"][2:]
    cond_size = wl.get("cond_e", 1) + wl["n_layers"]
    in_ch = 1 + (2 if cfg.get("R_Z_INPUT") else 0) + (1 if cfg.get("PHI_INPUT") else 0)
    shapes = {"model." + k: v for k, v in unet_param_shapes(cfg["LAYER_SIZE_UNET"], in_ch, cond_size, cfg["COND_SIZE_UNET"], zpr).items()}
    sd = synth_state_dict(shapes, seed=0)
    geom = None
    if wl.get("binning"):   # in-loop NNConverter with the area-weighted maps and their pseudo-inverses (utils/utils.py:588-603)
        import json
        from oracle.geom_embed import GeomEmbedOracle, geom_weight_mats
        with open(os.path.join(ROOT, "tests", "golden", wl["binning"])) as f:
            meta = json.load(f)
        mats, _ = geom_weight_mats(meta["lay_r_edges"])
        rel = [i for i, e in enumerate(meta["xml_r_edges"]) if len(e) > 1]
        alphas = [meta["xml_a_bins"][i] for i in rel]
        bounds = [0]
        for i, w in enumerate(mats):
            bounds.append(bounds[-1] + alphas[i] * w.shape[1])
        geom = GeomEmbedOracle.nn_converter(mats, [torch.linalg.pinv(w) for w in mats], bounds, alphas, max(alphas))
    o = DenoiserOracle(cfg, sd, geom=geom)
    g = torch.Generator().manual_seed(1)
    E = torch.rand(batch, wl.get("cond_e", 1), generator=g)
    layers = torch.randn(batch, wl["n_layers"], generator=g) if wl["n_layers"] else None
    start = torch.randn([batch] + _data_shape(cfg), generator=torch.Generator().manual_seed(2))
    evals_done, t_total, n = 0, 0.0, 1
    with torch.no_grad():
        o.denoise(start, E, torch.ones(batch), layers)   # warm-up (thread pool, allocator)
        while t_total < budget_s and evals_done < 4096:
            t0 = time.perf_counter()
            if wl["sampler"] in ("DDPM", "DDim"):
                sample_ddim(o, start, E, layers, wl["steps"], 1.0 if wl["sampler"] == "DDPM" else 0.0, None,
                            sample_offset=wl["steps"] - n)
                evals = n
            else:
                sample_edm(o, start, E, layers, max(n, 2), 2 if wl["sampler"] == "Heun" else 1, None, last_step_euler=False)
                evals = 2 * max(n, 2) if wl["sampler"] == "Heun" else max(n, 2)
            t_total += time.perf_counter() - t0
            evals_done += evals
            n = min(n * 2, 8)
    showers_per_s = batch * evals_done / t_total / wl["evals_per_shower"]
    sample = (f"{evals_done} denoiser evaluations of the {wl['sampler']} loop at batch {batch} "
              f"({t_total:.1f} s), scaled to {wl['evals_per_shower']} evaluations per shower")
    return showers_per_s, cores, sample, t_total


def _metric(wl):
    return ("training samples/sec", "samples/s") if wl.get("train") else ("showers/sec sampled", "showers/s")


def run_reference_arm(args, wl):
    """--impl reference: the reference's own implementation of the path on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B, steps = int(args.batch or wl["batch"]), int(args.sampling_steps or wl["steps"])
    vals, last = [], None
    for _ in range(max(args.warmup, 0)):
        reference_leg(args.workload, budget_s=2.0)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        last = reference_leg(args.workload, budget_s=args.budget)
        vals.append(last["value"])
    dt = time.perf_counter() - t0
    v = sum(vals) / len(vals)
    metric, unit = _metric(wl)
    line = {
        "impl": "reference", "metric": metric, "value": v, "unit": unit, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_block(wl, B, steps, args.gpus),
        "cpu_baseline": {"value": v, "unit": unit, "cores": last["cores"], "kind": last["kind"], "sample": last["sample"],
                         "device": last["device"]},
        "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "each step = a bounded sample of the workload (see cpu_baseline.sample) on the host CPU, CUDA hidden",
    }
    print(json.dumps(line), flush=True)


def _op_profile(model, x, E, layers, sigma, reps=3):
    """Per-op CUDA-event timing of single denoiser evaluations at the bench batch (eager diagnostic mode)."""
    import torch
    model.denoise(x, E=E, sigma=sigma, layers=layers)
    torch.cuda.synchronize()
    model.set_profile(2)
    for _ in range(reps):
        model.denoise(x, E=E, sigma=sigma, layers=layers)
    torch.cuda.synchronize()
    rep = model.profile_report()
    model.set_profile(0)
    return rep


def _roofline(rep, peaks, B, workload):
    """Roofline entries of the dominant kernel launch shape (largest share of the evaluation time) and of the dominant
    HBM-bound one; the per-class table goes to the op-profile file."""
    total_ms = sum(v["ms"] for v in rep.values())
    # group the per-shape op names into kernel classes
    classes = {}
    for name, v in rep.items():
        if name.startswith("conv_") and name.endswith("_mma"):
            cls = "conv_mma_kernel"
        elif name.startswith("conv_"):
            cls = "conv_generic_kernel"
        elif name.startswith("gn_apply"):
            cls = "gn_apply_kernel"
        else:
            cls = name.split("_L")[0]
        c = classes.setdefault(cls, dict(ms=0.0, flops=0.0, bytes=0.0, count=0))
        for k in ("ms", "flops", "bytes", "count"):
            c[k] += v[k]
    def entry(name):
        c = rep[name]
        if c["flops"] > 0 and name.startswith("conv"):
            achieved = c["flops"] / (c["ms"] * 1e-3) / 1e12
            roof = {"bound": "tensor", "achieved": achieved, "peak": peaks["tf_sus"], "unit": "TFLOP/s",
                    "frac": achieved / peaks["tf_sus"], "traffic": None}
        else:
            achieved = c["bytes"] / (c["ms"] * 1e-3) / 1e9
            roof = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm"], "unit": "GB/s",
                    "frac": achieved / peaks["hbm"], "traffic": None}
        roof.update(kernel=name, launches=c["count"], avg_launch_ms=c["ms"] / max(c["count"], 1),
                    share_of_eval=c["ms"] / total_ms,
                    peak_source=peaks["src"] + (" (sustained bf16/fp16 dense)" if roof["bound"] == "tensor" else ""))
        # DRAM bytes per launch of this kernel shape from the committed `ncu --set full` capture (profiles/), if recorded
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                tab = json.load(f)
                t = tab.get(f"{name}@{workload}") or tab.get(name)
            if t and t.get("batch") == B and t.get("workload") == workload:
                roof["traffic"] = t["dram_bytes_per_launch"]
                roof["traffic_source"] = t["source"]
        except Exception:
            pass
        return roof

    top = max(rep, key=lambda k: rep[k]["ms"])
    roof = entry(top)
    # the largest HBM-bound launch shape as well (north_star quotes both rooflines)
    hbm_names = [k for k in rep if not k.startswith("conv") and not k.startswith("attention") and rep[k]["bytes"] > 0]
    second = entry(max(hbm_names, key=lambda k: rep[k]["ms"])) if hbm_names else None
    table = {k: {kk: v[kk] for kk in ("ms", "flops", "bytes", "count")}
             for k, v in sorted(classes.items(), key=lambda kv: -kv[1]["ms"])}
    return roof, second, table, total_ms


def run_gpu_arm(args, wl):
    import numpy as np
    import torch
    import torch.distributed as dist
    from calodiffusion_b200.calodiffusion import CaloDiffusion

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    from calodiffusion_b200 import distributed as D
    cfg, B, steps = dict(wl["cfg"]), int(args.batch or wl["batch"]), int(args.sampling_steps or wl["steps"])
    strong = args.scaling == "strong"
    n_total = B if strong else B * world                   # showers of the whole job per step
    lo, hi = D.shard_range(n_total, rank, world)
    B = hi - lo                                            # this rank's slice
    if wl.get("binning"):
        _binning(wl, cfg)
    torch.manual_seed(0)                                   # identical random-init weights on every rank
    model = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=B)
    if os.environ.get("CD_NO_GRAPH"):
        model.use_graph = False                            # diagnostics: eager launches instead of one CUDA graph
    sampler = model.sampler_algorithm
    sampler.shower_offset = lo                             # global shower index keys the Philox stream
    if wl["sampler"] == "Heun":
        sampler.last_step_euler = True                     # finite showers; 2 evals on all other steps
    vox = int(np.prod(_data_shape(cfg)))
    g = torch.Generator().manual_seed(1)                   # the job's full conditioning, identical on every rank
    E_all = torch.rand(n_total, wl.get("cond_e", 1), generator=g).pin_memory()
    L_all = torch.randn(n_total, wl["n_layers"], generator=g).pin_memory() if wl["n_layers"] else None
    E_host, L_host = E_all[lo:hi], (L_all[lo:hi] if L_all is not None else None)
    E_dev = E_host.to(dev)
    L_dev = L_host.to(dev) if L_host is not None else None
    start_dev = torch.randn([B] + _data_shape(cfg), device=dev, generator=torch.Generator(device=dev).manual_seed(2 + rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step():
        x, _, _ = sampler(model, start_dev, E_dev, L_dev, steps, 0, False)
        if world > 1:
            D.gather_showers(x.reshape(B, vox), n_total)        # the one collective of the path
        return x

    seeds = iter(range(1000, 100000))

    def api_step():
        if world == 1:
            return model.sample(E_host, L_host, num_steps=steps)    # numpy on host (models/diffusion.py:77-104)
        # several GPUs: the public entry is distributed.sample_sharded - full conditioning from pinned host memory on every
        # rank, each samples its slice, one all-gather; rank 0 reads the whole job's showers back to the host
        e = E_all.to(dev, non_blocking=True)
        l = L_all.to(dev, non_blocking=True) if L_all is not None else None
        out = D.sample_sharded(model, e, l, num_steps=steps, seed=next(seeds))
        return out.cpu().numpy() if rank == 0 else out[lo:hi].cpu().numpy()

    for _ in range(args.warmup):
        x = device_step()
    barrier()
    assert bool(torch.isfinite(x).all()), "non-finite showers"
    launches0 = model.launch_count()
    clk = ClockSampler(local_rank)
    if rank == 0:
        clk.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        device_step()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = clk.finish() if rank == 0 else None
    launches = model.launch_count() - launches0

    # end to end through the public API, host buffers in, host numpy out
    api_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = api_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert out.shape[0] == (B if world == 1 or rank else n_total) and np.isfinite(out).all()

    t = torch.tensor([ms, e2e_s * 1e3], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0]), float(t[1])
    total_showers = n_total * args.steps
    value = total_showers / (ms * 1e-3)
    e2e = total_showers / (e2e_ms * 1e-3)
    h2d = 4 * (B if world == 1 else n_total) * (wl.get("cond_e", 1) + wl["n_layers"])    # per rank
    d2h = 4 * (B if world == 1 else n_total) * vox                                          # rank 0

    if rank == 0:
        peaks = _peaks()
        sig = torch.full((B,), 1.5, device=dev)
        rep = _op_profile(model, start_dev, E_dev, L_dev, sig)
        roof, roof_hbm, table, eval_ms = _roofline(rep, peaks, B, args.workload)
        reps = 3
        cpu = gpu_ref = None
        if world == 1 and not args.no_cpu_baseline:
            # the unmodified reference (oracle/_ref) in fresh interpreters: on the host cores (CUDA hidden), and eager on
            # this same B200 with torch/cuDNN defaults (TF32 convolutions) - SURVEY 8(d) "GPU reference baseline"
            r = reference_leg_subprocess(args.workload, "cpu", budget_s=12.0)
            cpu = {"value": r["value"], "unit": "showers/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
            try:
                r = reference_leg_subprocess(args.workload, "cuda", budget_s=6.0)
                gpu_ref = {"value": r["value"], "unit": "showers/s", "ms_per_eval": r["ms_per_eval"], "batch": r.get("batch"),
                           "kind": r["kind"], "device": r["device"], "sample": r["sample"]}
            except Exception as exc:   # reported, never fatal: this leg is context for the headline, not the headline
                gpu_ref = {"unavailable": str(exc)[-300:]}
        evals = wl["evals_per_shower"] * steps / wl["steps"]
        line = {
            "metric": "showers/sec sampled", "value": value, "unit": "showers/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": config_block(wl, n_total if strong else B, steps, world, strong),
            "e2e": {"value": e2e, "unit": "showers/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roof,
            "roofline_hbm": roof_hbm,
            "cpu_baseline": cpu,
            "gpu_reference": gpu_ref,
            "shower_evals_per_s": value * evals,
            "model_tflops": value * evals * wl["gflop_per_eval"] / 1e3,
            "eval_ms_profiled": eval_ms / reps,
        }
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", f"op_profile_{args.workload}_n{world}.json"), "w") as f:
            json.dump({"batch": B, "reps": reps, "classes": table, "ops": rep}, f, indent=1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_train_arm(args, wl):
    """--workload ds2_train (BASELINE config 5): one step = one optimizer step on 128 synthetic showers per GPU."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    from calodiffusion_b200.distributed import DataParallelTrainer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    cfg, B = dict(wl["cfg"]), int(args.batch or wl["batch"])
    torch.manual_seed(0)
    model = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=B)
    trainer = DataParallelTrainer(model, lr=TRAIN_LR)       # broadcasts rank 0's weights, hooks the gradient all-reduce
    shape = _data_shape(cfg)
    vox = int(np.prod(shape))
    g = torch.Generator().manual_seed(1 + rank)             # every rank trains on its own showers
    data_host = torch.randn([B] + shape, generator=g).pin_memory()
    E_host = torch.rand(B, 1, generator=g).pin_memory()
    L_host = torch.randn(B, wl["n_layers"], generator=g).pin_memory()
    data_dev, E_dev, L_dev = data_host.to(dev), E_host.to(dev), L_host.to(dev)
    gen = torch.Generator(device=dev).manual_seed(2 + rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step():
        # the loop body of train/train_diffusion.py:42-66 with the batch already on the device; fresh noise and noise levels
        # every step (torch's device RNG, as the reference draws them), no host read of the loss
        noise = torch.randn(data_dev.shape, device=dev, generator=gen)
        rnd = torch.randn(B, device=dev, generator=gen)
        trainer.step(data_dev, E_dev, L_dev, noise=noise, rnd_normal=rnd, read_loss=False)

    def api_step():
        # through the public API from pinned host buffers, loss read back like the reference's `batch_loss.item()`
        d = data_host.to(dev, non_blocking=True)
        e = E_host.to(dev, non_blocking=True)
        l = L_host.to(dev, non_blocking=True)
        return float(trainer.step(d, e, l))

    for _ in range(max(args.warmup, 3)):
        device_step()
    barrier()
    clk = ClockSampler(local_rank)
    if rank == 0:
        clk.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        device_step()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = clk.finish() if rank == 0 else None
    loss0 = api_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        loss = api_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert np.isfinite(loss) and np.isfinite(loss0), "non-finite training loss"
    t = torch.tensor([ms, e2e_s * 1e3], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0]), float(t[1])
    total = world * B * args.steps
    value, e2e = total / (ms * 1e-3), total / (e2e_ms * 1e-3)

    if rank == 0:
        peaks = _peaks()
        model.grad_sync = None                              # per-op profile of one rank's step (no collective inside)
        model.set_train_profile(2)
        reps = 3
        for _ in range(reps):
            noise = torch.randn(data_dev.shape, device=dev, generator=gen)
            rnd = torch.randn(B, device=dev, generator=gen)
            model.loss_function(model, data_dev, E_dev, noise=noise, layers=L_dev, rnd_normal=rnd)
        torch.cuda.synchronize()
        rep = model.train_profile_report()
        model.set_train_profile(0)
        step_ms = sum(v["ms"] for v in rep.values()) / reps
        ops_per_step = sum(v["count"] for v in rep.values()) // reps
        top = max((k for k in rep if rep[k]["flops"] > 0), key=lambda k: rep[k]["ms"])
        tv = rep[top]
        tf = tv["flops"] / (tv["ms"] * 1e-3) / 1e12
        roof = {"bound": "tensor", "kernel": top, "achieved": tf, "peak": peaks["tf_sus"], "unit": "TFLOP/s", "frac": tf / peaks["tf_sus"],
                "traffic": None, "share_of_step": tv["ms"] / (step_ms * reps), "launches_per_step": tv["count"] // reps,
                "ms_per_launch": tv["ms"] / tv["count"], "peak_source": peaks["src"] + " (sustained)",
                "flops_model": "2 x executed MACs of the convolution whose weight/data gradient the kernel computes"}
        table = sorted(((k, v["ms"] / reps, v["count"] // reps, (v["flops"] / (v["ms"] * 1e-3) / 1e12) if v["flops"] else None)
                        for k, v in rep.items()), key=lambda r: -r[1])
        cpu = gpu_ref = None
        if world == 1 and not args.no_cpu_baseline:
            r = reference_leg_subprocess(args.workload, "cpu", budget_s=12.0)
            cpu = {"value": r["value"], "unit": "samples/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
            try:
                r = reference_leg_subprocess(args.workload, "cuda", budget_s=6.0)
                gpu_ref = {"value": r["value"], "unit": "samples/s", "ms_per_step": r["ms_per_eval"], "batch": r.get("batch"),
                           "kind": r["kind"], "device": r["device"], "sample": r["sample"]}
            except Exception as exc:
                gpu_ref = {"unavailable": str(exc)[-300:]}
        line = {
            "metric": "training samples/sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": config_block(wl, B, 0, world),
            "e2e": {"value": e2e, "unit": "samples/s", "h2d_bytes_per_step": 4 * B * (vox + 1 + wl["n_layers"]), "d2h_bytes_per_step": 8},
            "gpu_launches": int(ops_per_step * args.steps),
            "gpu_launches_note": "one per engine op (lower bound: several ops launch 2-3 kernels), forward + backward + Adam re-layout excluded",
            "clocks": clocks, "roofline": roof, "cpu_baseline": cpu, "gpu_reference": gpu_ref,
            "step_ms_profiled": step_ms, "grad_allreduce_bytes": int(model.grad_buffer().numel() * 4), "loss": loss,
        }
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", f"op_profile_{args.workload}_n{world}.txt"), "w") as f:
            f.write(f"# {wl['name']}; B={B}; per optimizer step, eager per-op CUDA events (forward + backward), {reps} reps\n")
            for k, m_, c, tfl in table:
                f.write(f"{k:44s} {m_:8.3f} ms  x{c:3d}  {'' if tfl is None else f'{tfl:7.1f} TFLOP/s'}\n")
            f.write(f"total {step_ms:.3f} ms\n")
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "reference-leg"])
    ap.add_argument("--budget", type=float, default=10.0, help="seconds of reference work per step (--impl reference)")
    ap.add_argument("--workload", default="ds2", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="override showers per GPU (diagnostics only)")
    ap.add_argument("--sampling-steps", type=int, default=0, help="override sampler steps (diagnostics only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: the workload's batch on EVERY GPU; strong: the workload's batch in total, split over the GPUs")
    args = ap.parse_args()
    if args.impl == "reference":
        os.environ["CUDA_VISIBLE_DEVICES"] = ""      # the reference picks cuda when it sees one (utils.get_device)
    wl = WORKLOADS[args.workload]
    if args.impl == "reference-leg":      # internal: one bounded reference measurement on the visible device
        print(json.dumps(reference_leg(args.workload, budget_s=args.budget)), flush=True)
    elif args.impl == "reference":
        run_reference_arm(args, wl)
    elif wl.get("train"):
        run_train_arm(args, wl)
    else:
        run_gpu_arm(args, wl)


if __name__ == "__main__":
    main()
