#!/usr/bin/env python
"""Informational GPU baseline: the reference's op-by-op algorithm (the oracle's plain torch restatement: F.conv3d with
explicit circular padding, F.group_norm, einsum attention, ...) executed eagerly by torch/cuDNN on the same B200, i.e. the
library kernels the reference itself would run on Blackwell (SURVEY 8d "GPU reference baseline").  Not part of bench.py's
contract; prints one JSON line per precision setting.
    python scripts/torch_gpu_baseline.py [--workload ds2|ds3] [--batch N] [--reps K]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from bench import WORKLOADS  # noqa: E402
from calodiffusion_b200.calodiffusion import unet_param_shapes  # noqa: E402
from oracle.unet_oracle import DenoiserOracle, synth_state_dict  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="ds2")
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
wl = WORKLOADS[a.workload]
cfg, B = wl["cfg"], a.batch or wl["batch"]
zpr = cfg["SHAPE_FINAL"][2:]
in_ch = 1 + (2 if cfg.get("R_Z_INPUT") else 0) + (1 if cfg.get("PHI_INPUT") else 0)
shapes = {"model." + k: v for k, v in unet_param_shapes(cfg["LAYER_SIZE_UNET"], in_ch, wl.get("cond_e", 1) + wl["n_layers"], cfg["COND_SIZE_UNET"], zpr).items()}
sd = synth_state_dict(shapes, seed=0)
g = torch.Generator().manual_seed(1)
E = torch.rand(B, wl.get("cond_e", 1), generator=g).cuda()
L = torch.randn(B, wl["n_layers"], generator=g).cuda() if wl["n_layers"] else None
x = torch.randn([B] + cfg["SHAPE_PAD"][1:], device="cuda")
sig = torch.full((B,), 1.5, device="cuda")
for name, tf32, dtype in (("tf32 conv (torch default), fp32 tensors", True, torch.float32), ("bf16 tensors", True, torch.bfloat16)):
    torch.backends.cudnn.allow_tf32 = tf32
    torch.backends.cudnn.benchmark = True
    o = DenoiserOracle(cfg, sd, dtype=dtype, device="cuda")
    try:
        with torch.no_grad():
            for _ in range(2):
                o.denoise(x, E, sig, L)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.reps):
                o.denoise(x, E, sig, L)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.reps
        print(json.dumps({"baseline": "torch/cuDNN eager restatement of the reference on this GPU", "precision": name,
                          "workload": wl["name"], "batch": B, "ms_per_denoiser_eval": ms, "shower_evals_per_s": B / ms * 1e3,
                          "showers_per_s_at_%d_evals" % wl["evals_per_shower"]: B / ms * 1e3 / wl["evals_per_shower"]}), flush=True)
    except Exception as ex:   # e.g. out of memory at the full batch
        print(json.dumps({"baseline": "torch/cuDNN eager", "precision": name, "error": str(ex)[:200]}), flush=True)
    del o
    torch.cuda.empty_cache()
