"""CPU: the layer-energy model oracle (dense ResNet of LayerDiffusion) vs the live reference's golden outputs, and the
product's parameter table vs the reference's state-dict names."""
import json
import os

import numpy as np
import torch

from conftest import GOLDEN, rel_l2
from oracle.layer_oracle import LayerDenoiserOracle
from oracle.unet_oracle import sample_ddim, synth_state_dict


def load_layer_case():
    with open(os.path.join(GOLDEN, "layer_meta.json")) as f:
        meta = json.load(f)
    gold = dict(np.load(os.path.join(GOLDEN, "layer.npz")))
    B, D = meta["B"], meta["config"]["SHAPE_FINAL"][2] + 1
    E = torch.rand(B, 1, generator=torch.Generator().manual_seed(1))
    x = torch.randn(B, D, generator=torch.Generator().manual_seed(2))
    gb = torch.Generator().manual_seed(3)
    bank = [torch.randn(B, D, generator=gb) for _ in range(3)]
    return dict(cfg=meta["config"], sd=synth_state_dict(meta["shapes"], seed=meta["weight_seed"]), gold=gold, E=E, x=x, bank=bank,
                shapes=meta["shapes"], B=B, D=D)


def test_layer_oracle_matches_reference():
    c = load_layer_case()
    o = LayerDenoiserOracle(c["cfg"], c["sd"])
    with torch.no_grad():
        for tag in ("a", "b"):
            sig = torch.from_numpy(c["gold"][f"denoise_{tag}_sigma"])
            y = o.denoise(c["x"] * (1 + sig.view(-1, 1) ** 2).sqrt(), c["E"], sig).numpy()
            assert rel_l2(y, c["gold"][f"denoise_{tag}"]) < 2e-5, tag
        t = torch.tensor([0.2, -0.7, 1.1, 0.0, -2.0])
        assert rel_l2(o.forward(c["x"], c["E"], t).numpy(), c["gold"]["forward"]) < 2e-5
        x, xs, x0s = sample_ddim(o, c["x"], c["E"], None, 3, 0.0, c["bank"])
    assert rel_l2(torch.stack(x0s).numpy(), c["gold"]["DDim_x0s"]) < 2e-5
    assert rel_l2(torch.stack(xs).numpy(), c["gold"]["DDim_xs"]) < 2e-5
    assert rel_l2(x.numpy(), c["gold"]["DDim_x"]) < 2e-5


def test_resnet_param_table_matches_reference_names():
    from calodiffusion_b200.layerdiffusion import resnet_param_shapes
    c = load_layer_case()
    mine = resnet_param_shapes(c["D"], 5, 256, 128, 1)
    assert list(mine) == list(c["shapes"])
    assert {k: list(v) for k, v in mine.items()} == c["shapes"]
