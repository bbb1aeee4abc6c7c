"""Golden vectors for BASELINE config 1 (CaloChallenge Dataset 1 photons, in-loop NNConverter) from the LIVE reference.

    python oracle/make_golden_ds1.py        # writes tests/golden/ds1.npz + ds1_meta.json  (build container only)

The UNMODIFIED reference `CaloDiffusion` (config_dataset1_photon.json, BIN_FILE -> the in-tree binning XML) is built on
CPU; its U-Net gets the seeded synthetic weights of oracle.unet_oracle.synth_state_dict, its `NN_embed` keeps the
reference's own initialisation (area-weighted GeomConverter maps + 1e-5 noise, pseudo-inverse decoders), which is
stored.  Outputs stored: GeomConverter.weight_mats, NNConverter.enc / dec of seeded inputs, `denoise` at two noise
levels, and a 3-step DDim run (x, x0s, xs) with a replayed noise bank.  Inputs are regenerated from seeds by the tests.
Also stores Embeder / Decoder (HGCal_utils.py:296-349) outputs for seeded sparse matrices of the HGCal fixture shape.
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_shim  # noqa: E402
from oracle.make_golden import make_inputs, noise_bank, run_sampler  # noqa: E402
from oracle.unet_oracle import synth_state_dict  # noqa: E402
from oracle.geom_embed import hgcal_matrices  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
XML = "/root/reference/CaloChallenge/code/binning_dataset_1_photons.xml"


def main():
    torch.set_num_threads(os.cpu_count())
    model, cfg = ref_shim.load_reference_model("config_dataset1_photon.json", "DDim", {"BIN_FILE": XML})
    assert model.do_embed
    ref_sd = model.state_dict()
    shapes = {k: list(v.shape) for k, v in ref_sd.items()}
    unet_shapes = {k: v for k, v in shapes.items() if k.startswith("model.")}
    sd = dict(ref_sd)
    sd.update(synth_state_dict(unet_shapes, seed=0))
    model.load_state_dict(sd)
    gc = model.NN_embed.gc
    from calodiffusion.utils.XMLHandler import XMLHandler
    bins = XMLHandler("photon", XML)
    out = {}
    for i in range(gc.num_layers):
        out[f"gc_weight_{i}"] = gc.weight_mats[i].numpy()
        out[f"enc_w_{i}"] = model.NN_embed.encs[i].weight.detach().numpy().copy()
        out[f"dec_w_{i}"] = model.NN_embed.decs[i].weight.detach().numpy().copy()
    out["layer_boundaries"] = np.asarray(gc.layer_boundaries, np.int64)
    out["lay_alphas"] = np.asarray(gc.lay_alphas, np.int64)
    out["all_r_edges"] = gc.all_r_edges.numpy()
    B = 3
    shape = cfg["SHAPE_ORIG"][1:]
    E, layers, x = make_inputs(shape, B, 1, 6)
    with torch.no_grad():
        grid = model.NN_embed.enc(x)
        out["enc_out"] = grid.numpy()
        gy = torch.randn(grid.shape, generator=torch.Generator().manual_seed(7))
        out["dec_out"] = model.NN_embed.dec(gy).numpy()
        for tag, sig in (("a", [1.5, 0.3, 4.0]), ("b", [40.0, 0.01, 0.5])):
            sigma = torch.tensor(sig).view(B, 1)
            y = model.denoise(x * (1 + sigma ** 2).sqrt(), E=E, sigma=sigma, layers=layers)
            out[f"denoise_{tag}_sigma"] = np.array(sig, np.float32)
            out[f"denoise_{tag}"] = y.numpy()
        bank = noise_bank(shape, B, 3)
        xf, xs, x0s = run_sampler(model, "DDim", x, E, layers, 3, bank)
        out["DDim_x"] = xf.numpy()
        out["DDim_x0s"] = np.stack([t.numpy() for t in x0s])
        out["DDim_xs"] = np.stack([t.numpy() for t in xs])

        # HGCal Embeder / Decoder (non-trainable: the mask is not applied, HGCal_utils.py:316,341)
        from calodiffusion.utils.HGCal_utils import Embeder, Decoder
        enc_mat, dec_mat = hgcal_matrices()
        xh = torch.randn(2, 1, 28, 1988, generator=torch.Generator().manual_seed(11))
        gh = Embeder(12, 21, enc_mat, enc_mat != 0)(xh)
        out["hgcal_enc_out"] = gh.numpy()
        yh = torch.randn(2, 1, 28, 12, 21, generator=torch.Generator().manual_seed(12))
        out["hgcal_dec_out"] = Decoder(12, 21, dec_mat, dec_mat != 0)(yh).numpy()
    np.savez_compressed(os.path.join(OUT, "ds1.npz"), **out)
    cfg_out = {k: v for k, v in cfg.items() if k not in ("flags", "BIN_FILE")}
    with open(os.path.join(OUT, "ds1_meta.json"), "w") as f:
        json.dump({"config": cfg_out, "shapes": unet_shapes, "cond_e": 1, "cond_layers": 6, "B": B, "weight_seed": 0,
                   "input_seed": 1, "noise_seed": 3, "lay_r_edges": [list(map(float, e)) for e in gc.lay_r_edges],
                   # the binning file as the reference's XMLHandler parsed it (all layers), so tests can re-emit it
                   "xml_r_edges": [list(map(float, e)) for e in bins.r_edges], "xml_a_bins": [int(a) for a in bins.a_bins]},
                  f, indent=0)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
