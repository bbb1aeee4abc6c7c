#!/usr/bin/env python
"""Target program for `ncu --set full`: a few eager denoiser evaluations of the bench workload (no graph), so a
kernel-name filter plus a launch count picks the dominant kernels at the bench batch."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from bench import WORKLOADS  # noqa: E402
from calodiffusion_b200.calodiffusion import CaloDiffusion  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="ds2")
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--evals", type=int, default=2)
a = ap.parse_args()
wl = WORKLOADS[a.workload]
cfg, B = dict(wl["cfg"]), a.batch or wl["batch"]
torch.manual_seed(0)
m = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=B)
g = torch.Generator().manual_seed(1)
E = torch.rand(B, 1, generator=g).cuda()
L = torch.randn(B, wl["n_layers"], generator=g).cuda() if wl["n_layers"] else None
x = torch.randn([B] + cfg["SHAPE_PAD"][1:], device="cuda")
sig = torch.full((B,), 1.5, device="cuda")
if os.environ.get("CD_DEBUG_SYNC"):
    m.denoise(x[:1], E=E[:1], sigma=sig[:1], layers=None if L is None else L[:1])
    from calodiffusion_b200 import _lib
    _lib.check(m._lib.cd_unet_set_option(m._handle, b"debug_sync", 1), "set_option")
for _ in range(a.evals):
    y = m.denoise(x, E=E, sigma=sig, layers=L)
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
