"""Golden vectors for the layer-energy model of LayerDiffusion from the LIVE reference (build container only).

    python oracle/make_golden_layer.py        # writes tests/golden/layer.npz + layer_meta.json

The UNMODIFIED reference `LayerDiffusion` (config_dataset2.json) is built on CPU, its `layer_model` (dense ResNet) gets
seeded synthetic weights, and in the layer state (`set_layer_state(True)`) its `denoise` at two noise levels, its raw
`layer_forward`, and a 3-step DDim run over (B, 46) vectors with a replayed noise bank are stored.
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_shim  # noqa: E402
from oracle.make_golden import ReplayNoise  # noqa: E402
from oracle.unet_oracle import synth_state_dict  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def main():
    ref_shim.install()
    from calodiffusion.utils import utils as ref_utils
    from calodiffusion.models.layerdiffusion import LayerDiffusion
    cfg = ref_utils.LoadJson(f"{ref_shim.REF_ROOT}/calodiffusion/configs/config_dataset2.json")
    cfg["SAMPLER"] = "DDim"
    cfg["LAYER_SAMPLER"] = "DDim"
    model = LayerDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"])
    model.eval()
    shapes = {k: list(v.shape) for k, v in model.layer_model.state_dict().items()}
    sd = synth_state_dict(shapes, seed=4)
    model.layer_model.load_state_dict(sd)
    model.set_layer_state(True)
    B, D = 5, cfg["SHAPE_FINAL"][2] + 1
    g = torch.Generator().manual_seed(1)
    E = torch.rand(B, 1, generator=g)
    x = torch.randn(B, D, generator=torch.Generator().manual_seed(2))
    out = {}
    with torch.no_grad():
        for tag, sig in (("a", [1.5, 0.3, 4.0, 0.05, 20.0]), ("b", [40.0, 0.01, 0.5, 2.0, 80.0])):
            sigma = torch.tensor(sig).view(B, 1)
            out[f"denoise_{tag}_sigma"] = np.array(sig, np.float32)
            out[f"denoise_{tag}"] = model.denoise(x * (1 + sigma ** 2).sqrt(), E=E, sigma=sigma, layers=None).numpy()
        t = torch.tensor([0.2, -0.7, 1.1, 0.0, -2.0])
        out["forward"] = model.layer_forward(x, E, t).numpy()
        gb = torch.Generator().manual_seed(3)
        bank = [torch.randn(B, D, generator=gb) for _ in range(3)]
        sampler = ref_utils.load_attr("sampler", "DDim")(cfg)
        with ReplayNoise(bank):
            xf, xs, x0s = sampler(model, x, E, None, 3, 0, True)
        out["DDim_x"] = xf.numpy()
        out["DDim_x0s"] = np.stack([v.numpy() for v in x0s])
        out["DDim_xs"] = np.stack([v.numpy() for v in xs])
    np.savez_compressed(os.path.join(OUT, "layer.npz"), **out)
    with open(os.path.join(OUT, "layer_meta.json"), "w") as f:
        json.dump({"config": {k: v for k, v in cfg.items() if k != "flags"}, "shapes": shapes, "B": B, "weight_seed": 4}, f, indent=0)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
