import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from conftest import load_case, noise_bank, rel_l2
from calodiffusion_b200.calodiffusion import CaloDiffusion
from calodiffusion_b200 import sample as S
c = load_case("ds3")
cfg = dict(c["cfg"], SAMPLER="Heun")
m = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=4)
m.load_state_dict(dict(c["sd"]))
for graph in (False, True):
    m.use_graph = graph
    s = S.Heun(cfg)
    s.noise_bank = torch.stack(noise_bank(c["shape"], c["B"], 3)).cuda()
    x, xs, x0s = s(m, c["x"].cuda(), c["E"].cuda(), None, 3, 0, True)
    for i in range(3):
        a = x0s[i].cpu().numpy(); g = c["gold"]["Heun_x0s"][i]
        print("graph", graph, "step", i, "finite", np.isfinite(a).all(), "nbad", (~np.isfinite(a)).sum(), "maxabs", np.nanmax(np.abs(a)),
              "err", rel_l2(np.nan_to_num(a), g))
        b = xs[i].cpu().numpy(); print("   xs finite", np.isfinite(b).all(), rel_l2(np.nan_to_num(b), c["gold"]["Heun_xs"][i]))
