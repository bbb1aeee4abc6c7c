"""Diagnostic: time the tcgen05 conv launches of one denoiser evaluation with parts of the kernel switched off
(library built with CD_EXTRA=-DCD_MMA_DEBUG; results are garbage in those modes, only the timing is meaningful).
    CD_EXTRA=-DCD_MMA_DEBUG bash calodiffusion_b200/csrc/build.sh && python scripts/conv_modes.py [workload] [batch]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import WORKLOADS, _op_profile
from calodiffusion_b200.calodiffusion import CaloDiffusion
key = sys.argv[1] if len(sys.argv) > 1 else "ds2"
wl = WORKLOADS[key]
B = int(sys.argv[2]) if len(sys.argv) > 2 else wl["batch"]
cfg = dict(wl["cfg"])
torch.manual_seed(0)
m = CaloDiffusion(cfg, n_steps=400, loss_type="l2", max_batch=B)
g = torch.Generator().manual_seed(1)
E = torch.rand(B, 1, generator=g).cuda()
L = torch.randn(B, wl["n_layers"], generator=g).cuda() if wl["n_layers"] else None
x = torch.randn([B] + cfg["SHAPE_PAD"][1:], device="cuda")
sig = torch.full((B,), 1.5, device="cuda")
names = {0: "baseline", 1: "epilogue off", 3: "epilogue off + A loaded once", 2: "A loaded once", 4: "no exchange/shuffles", 8: "no stores",
         16: "no statistics", 28: "no shuffles/stores/stats", 32: "no TMEM loads", 60: "TMEM-wait + barrier only"}
names.update({129: "epilogue off, no MMAs", 131: "epilogue off, no MMAs, A once", 128: "no MMAs"})
modes = [int(a) for a in sys.argv[3].split(",")] if len(sys.argv) > 3 else (0, 1, 3, 2, 4, 8, 16, 28, 32, 60)
for mode in modes:
    os.environ["CD_DBG_MODE"] = str(mode)
    rep = _op_profile(m, x, E, L, sig, reps=3)
    sel = {k: v["ms"] / v["count"] for k, v in rep.items() if k.startswith("conv_k27") and "_L0_" in k}
    print(f"mode {mode:2d} ({names.get(mode, '?'):30s}): " + "  ".join(f"{k} {v:.3f} ms" for k, v in sorted(sel.items())), flush=True)
