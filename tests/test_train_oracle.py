"""The training oracle (oracle/train_oracle.py: loss + autograd through the functional denoiser) against fingerprints of the LIVE
reference's `compute_loss(...).backward()` (tests/golden/train_grads.npz, oracle/make_golden_train.py)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, load_case
from oracle.train_oracle import loss_and_grads, sigma_for


def train_inputs(shape, B, cond_e, cond_layers, seed=51):
    g = torch.Generator().manual_seed(seed)
    E = torch.rand(B, cond_e, generator=g)
    layers = torch.randn(B, cond_layers, generator=g) if cond_layers else None
    data = torch.randn([B] + list(shape), generator=g)
    noise = torch.randn([B] + list(shape), generator=g)
    rnd = torch.randn(B, generator=g)
    time = torch.randint(0, 400, (B,), generator=g)      # used by the discrete noise schedules (dataset 3: 'cosine')
    return E, layers, data, noise, rnd, time


@pytest.mark.parametrize("case", ["ds2", "ds3"])
def test_oracle_gradients_match_reference_fingerprints(case):
    gold = np.load(os.path.join(GOLDEN, "train_grads.npz"))
    c = load_case(case)
    E, layers, data, noise, rnd, time = train_inputs(c["shape"], 2, c["meta"]["cond_e"], c["meta"]["cond_layers"])
    torch.set_num_threads(os.cpu_count() or 1)
    loss, grads = loss_and_grads(c["cfg"], c["sd"], data, E, layers, noise, sigma_for(c["cfg"], rnd, time))
    assert abs(loss - float(gold[f"{case}_loss"])) < 1e-5 * abs(float(gold[f"{case}_loss"]))
    worst = 0.0
    for name, norm, seed, proj in zip(gold[f"{case}_names"], gold[f"{case}_norm"], gold[f"{case}_seed"], gold[f"{case}_proj"]):
        g = grads[str(name)].double().reshape(-1)
        v = torch.randn(g.numel(), generator=torch.Generator().manual_seed(int(seed)), dtype=torch.float64)
        assert abs(float(g.norm()) - norm) <= 2e-4 * norm + 1e-12, name
        worst = max(worst, abs(float(g @ v) - proj) / (norm * float(v.norm()) + 1e-30))
    assert worst < 2e-4, worst
