#!/usr/bin/env python
"""Two training steps of the Dataset-2 workload at B showers (for an ncu launch list of the backward pass):
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python scripts/train_profile.py 128
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from calodiffusion_b200.calodiffusion import CaloDiffusion  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    wl = bench.WORKLOADS["ds2_train"]
    cfg = dict(wl["cfg"])
    torch.manual_seed(0)
    m = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=B)
    g = torch.Generator().manual_seed(1)
    data = torch.randn([B] + cfg["SHAPE_PAD"][1:], generator=g).cuda()
    E = torch.rand(B, 1, generator=g).cuda()
    L = torch.randn(B, 46, generator=g).cuda()
    for _ in range(steps):
        loss = m.loss_function(m, data, E, noise=torch.randn_like(data), layers=L, rnd_normal=torch.randn(B, device="cuda"))
        m.fused_adam_step(4e-4)
    torch.cuda.synchronize()
    print("loss", float(loss))


if __name__ == "__main__":
    main()
