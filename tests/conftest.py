import json
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def load_case(case):
    """Golden outputs of the live reference + the seeded inputs/weights that produced them."""
    from oracle.unet_oracle import synth_state_dict
    with open(os.path.join(GOLDEN, f"{case}_meta.json")) as f:
        meta = json.load(f)
    gold = dict(np.load(os.path.join(GOLDEN, f"{case}.npz")))
    cfg = meta["config"]
    sd = synth_state_dict(meta["shapes"], seed=meta["weight_seed"])
    shape = cfg["SHAPE_ORIG"][1:] if "orig" in cfg.get("SHOWER_EMBED", "") else cfg["SHAPE_PAD"][1:]  # diffusion.py:40-43
    B = meta["B"]
    g = torch.Generator().manual_seed(meta["input_seed"])
    E = torch.rand(B, meta["cond_e"], generator=g)
    layers = torch.randn(B, meta["cond_layers"], generator=g) if meta["cond_layers"] else None
    x = torch.randn([B] + list(shape), generator=torch.Generator().manual_seed(meta["input_seed"] + 1))
    return dict(cfg=cfg, sd=sd, gold=gold, E=E, layers=layers, x=x, shape=shape, B=B, meta=meta)


def noise_bank(shape, B, n, seed=3):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn([B] + list(shape), generator=g) for _ in range(n)]


@pytest.fixture(scope="session", params=["ds2", "ds3", "hgcal"])
def case(request):
    return request.param, load_case(request.param)


def write_binning_xml(path, meta, particle="photon"):
    """Re-emit the CaloChallenge binning file (as the reference's XMLHandler parsed it when the golden was made)."""
    from calodiffusion_b200.geometry import write_binning_xml as emit
    return emit(path, meta["xml_r_edges"], meta["xml_a_bins"], particle)


def ds1_geom_oracle(c):
    """The oracle's NNConverter enc/dec with the reference's stored converter weights."""
    from oracle.geom_embed import GeomEmbedOracle
    g = c["gold"]
    n = len(g["lay_alphas"])
    enc_w = [torch.from_numpy(g[f"enc_w_{i}"]) for i in range(n)]
    dec_w = [torch.from_numpy(g[f"dec_w_{i}"]) for i in range(n)]
    return GeomEmbedOracle.nn_converter(enc_w, dec_w, g["layer_boundaries"], g["lay_alphas"], int(g["lay_alphas"].max()))
