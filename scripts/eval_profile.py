"""Diagnostic: per-op profile of one denoiser evaluation at the bench batch (prints the total and the top ops)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import WORKLOADS, _op_profile
from calodiffusion_b200.calodiffusion import CaloDiffusion
key = sys.argv[1] if len(sys.argv) > 1 else "ds2"
top = int(sys.argv[2]) if len(sys.argv) > 2 else 8
wl = WORKLOADS[key]
B = int(sys.argv[3]) if len(sys.argv) > 3 else wl["batch"]
cfg = dict(wl["cfg"])
torch.manual_seed(0)
m = CaloDiffusion(cfg, n_steps=400, loss_type="l2", max_batch=B)
g = torch.Generator().manual_seed(1)
E = torch.rand(B, wl.get("cond_e", 1), generator=g).cuda()
L = torch.randn(B, wl["n_layers"], generator=g).cuda() if wl["n_layers"] else None
x = torch.randn([B] + cfg["SHAPE_PAD"][1:], device="cuda")
sig = torch.full((B,), 1.5, device="cuda")
rep = _op_profile(m, x, E, L, sig, reps=5)
tot = sum(v["ms"] for v in rep.values()) / 5
print(f"{key} B={B}: total {tot:.3f} ms/eval; " + "  ".join(f"{k} {v['ms']/v['count']:.3f}x{v['count']//5}" for k, v in sorted(rep.items(), key=lambda kv: -kv[1]['ms'])[:top]))
