"""Generate golden vectors from the LIVE reference (run in the build container only).

    python oracle/make_golden.py            # writes tests/golden/*.npz

For each case: seeded synthetic weights (oracle.unet_oracle.synth_state_dict over the reference's own
state-dict name→shape table) are loaded into the UNMODIFIED reference `CaloDiffusion`
(/root/reference/calodiffusion/models/calodiffusion.py) on CPU fp32; seeded inputs go through the
reference's own `denoise` and sampler classes (models/sample.py) with `torch.randn`/`randn_like`
patched to replay a seeded noise bank.  Inputs are regenerated from the seeds by the tests; only the
reference OUTPUTS (plus the name→shape table) are stored.
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_shim  # noqa: E402
from oracle.unet_oracle import synth_state_dict  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")

HGCAL_OVERRIDES = {  # /root/reference/tests/test_execution.py:178-190 (pytest hgcal fixture shape)
    "HGCAL": True, "SHAPE_ORIG": [-1, 28, 1988], "DATASET_NUM": 111, "SHAPE_PAD": [-1, 1, 28, 12, 21],
    "SHAPE_FINAL": [-1, 1, 28, 12, 21], "MAX_CELLS": 1988, "LAYER_SIZE_UNET": [32, 32, 64, 96],
    "SHOWER_EMBED": "NN-pre-embed",
}

CASES = {
    "ds2": dict(config="config_dataset2.json", overrides=None, cond_e=1, cond_layers=46),
    "ds3": dict(config="config_dataset3.json", overrides=None, cond_e=1, cond_layers=0),
    "hgcal": dict(config="config_HGCal.json", overrides=HGCAL_OVERRIDES, cond_e=3, cond_layers=29),
}


def make_inputs(shape, B, cond_e, cond_layers, seed=1):
    """The recipe the tests repeat: E~U[0,1) (B,cond_e); layers~N(0,1); x~N(0,1)."""
    g = torch.Generator().manual_seed(seed)
    E = torch.rand(B, cond_e, generator=g)
    layers = torch.randn(B, cond_layers, generator=g) if cond_layers else None
    x = torch.randn([B] + list(shape), generator=torch.Generator().manual_seed(seed + 1))
    return E, layers, x


def noise_bank(shape, B, n, seed=3):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn([B] + list(shape), generator=g) for _ in range(n)]


class ReplayNoise:
    """Patch torch.randn / torch.randn_like so the reference samplers consume a fixed bank."""

    def __init__(self, bank):
        self.bank, self.i = bank, 0

    def __enter__(self):
        self._randn, self._randn_like = torch.randn, torch.randn_like

        def randn(*a, **k):
            v = self.bank[self.i]
            self.i += 1
            return v.clone()

        torch.randn = randn
        torch.randn_like = lambda t, *a, **k: randn()
        return self

    def __exit__(self, *exc):
        torch.randn, torch.randn_like = self._randn, self._randn_like


def run_sampler(model, name, start, E, layers, steps, bank):
    from calodiffusion.utils import utils as ref_utils
    sampler = ref_utils.load_attr("sampler", name)(model.config)
    with ReplayNoise(bank):
        x, xs, x0s = sampler(model, start, E, layers, steps, 0, True)
    return x, xs, x0s


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(os.cpu_count())
    for case, spec in CASES.items():
        model, cfg = ref_shim.load_reference_model(spec["config"], "DDPM", spec["overrides"])
        ref_sd = model.state_dict()
        shapes = {k: list(v.shape) for k, v in ref_sd.items()}
        sd = synth_state_dict(shapes, seed=0)
        model.load_state_dict(sd)
        shape = cfg["SHAPE_PAD"][1:]
        B = 2
        E, layers, x = make_inputs(shape, B, spec["cond_e"], spec["cond_layers"])
        out = {}
        with torch.no_grad():
            for tag, sig in (("a", [1.5, 0.3]), ("b", [40.0, 0.01])):
                sigma = torch.tensor(sig).view(B, 1, 1, 1, 1)
                scale = (1 + sigma ** 2).sqrt()  # x is at the noise level sigma
                y = model.denoise(x * scale, E=E, sigma=sigma, layers=layers)
                out[f"denoise_{tag}_sigma"] = np.array(sig, np.float32)
                out[f"denoise_{tag}"] = y.numpy()
            start = x
            if case == "ds2":
                for name, steps in (("DDPM", 4), ("DDim", 3)):
                    bank = noise_bank(shape, B, steps)
                    xf, xs, x0s = run_sampler(model, name, start, E, layers, steps, bank)
                    out[f"{name}_x"] = xf.numpy()
                    out[f"{name}_x0s"] = np.stack([t.numpy() for t in x0s])
                    out[f"{name}_xs"] = np.stack([t.numpy() for t in xs])
            if case == "ds3":
                for name, steps in (("Euler", 3), ("Heun", 3)):
                    bank = noise_bank(shape, B, steps)
                    xf, xs, x0s = run_sampler(model, name, start, E, layers, steps, bank)
                    out[f"{name}_x"] = xf.numpy()
                    out[f"{name}_x0s"] = np.stack([t.numpy() for t in x0s])
                    out[f"{name}_xs"] = np.stack([t.numpy() for t in xs])
            if case == "hgcal":
                bank = noise_bank(shape, B, 2)
                xf, xs, x0s = run_sampler(model, "DDim", start, E, layers, 2, bank)
                out["DDim_x"] = xf.numpy()
                out["DDim_x0s"] = np.stack([t.numpy() for t in x0s])
        np.savez_compressed(os.path.join(OUT, f"{case}.npz"), **out)
        cfg_out = {k: v for k, v in cfg.items() if k != "flags"}
        with open(os.path.join(OUT, f"{case}_meta.json"), "w") as f:
            json.dump({"config": cfg_out, "shapes": shapes, "cond_e": spec["cond_e"],
                       "cond_layers": spec["cond_layers"], "B": B, "weight_seed": 0, "input_seed": 1,
                       "noise_seed": 3}, f, indent=0)
        print(case, {k: v.shape for k, v in out.items()})

    # cylindrical-conv known answer, reference tests/test_cyl_conv.py:7-25 run through the live module
    ref_shim.install()
    from calodiffusion.models.models import CylindricalConv
    conv = CylindricalConv(1, 1, kernel_size=(1, 3, 3), padding=(0, 1, 1), bias=False)
    with torch.no_grad():
        conv.conv.weight.fill_(1.0)
        xin = torch.tensor([1.0, 2.0, 3.0]).view(1, 1, 1, 1, 3).expand(1, 1, 1, 4, 3).contiguous()
        y = conv(xin)
    np.savez(os.path.join(OUT, "cylconv_kat.npz"), x=xin.numpy(), y=y.numpy())
    print("cylconv", y.flatten().tolist())


if __name__ == "__main__":
    main()
