"""Pin of the training oracle to the LIVE reference (build container only):  python oracle/make_golden_train.py

The unmodified reference CaloDiffusion (config_dataset2.json, seeded synthetic weights) runs
`compute_loss(data, E, noise=noise, layers=layers, rnd_normal=rnd)` and `.backward()` on CPU fp32 (models/diffusion.py:106-110,
models/loss.py:118-179); stored per parameter: the gradient's L2 norm and its product with a seeded random vector, plus the loss.
The full gradients (2.2 M values) are recomputed by the oracle in the tests; these fingerprints check that the oracle's are
the reference's."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_shim  # noqa: E402
from oracle.make_golden import CASES, OUT  # noqa: E402
from oracle.unet_oracle import synth_state_dict  # noqa: E402


def train_inputs(shape, B, cond_e, cond_layers, seed=51):
    g = torch.Generator().manual_seed(seed)
    E = torch.rand(B, cond_e, generator=g)
    layers = torch.randn(B, cond_layers, generator=g) if cond_layers else None
    data = torch.randn([B] + list(shape), generator=g)
    noise = torch.randn([B] + list(shape), generator=g)
    rnd = torch.randn(B, generator=g)
    time = torch.randint(0, 400, (B,), generator=g)      # used by the discrete noise schedules (dataset 3: 'cosine')
    return E, layers, data, noise, rnd, time


def main():
    torch.set_num_threads(os.cpu_count())
    out = {}
    for case in ("ds2", "ds3"):
        spec = CASES[case]
        model, cfg = ref_shim.load_reference_model(spec["config"], "DDPM", spec["overrides"])
        shapes = {k: list(v.shape) for k, v in model.state_dict().items()}
        model.load_state_dict(synth_state_dict(shapes, seed=0))
        model.train()
        E, layers, data, noise, rnd, time = train_inputs(cfg["SHAPE_PAD"][1:], 2, spec["cond_e"], spec["cond_layers"])
        # Diffusion.compute_loss (models/diffusion.py:106-110) does not forward `time`, so a discrete schedule would draw it from
        # the global RNG: call the loss object directly (Loss.__call__, models/loss.py:118-142) with both given
        loss = model.loss_function(model, data, E, noise=noise, time=time, layers=layers, rnd_normal=rnd)
        loss.backward()
        out[f"{case}_loss"] = np.float64(loss.item())
        names, norms, projs = [], [], []
        for name, p in model.named_parameters():
            g = p.grad.detach().double().reshape(-1)
            v = torch.randn(g.numel(), generator=torch.Generator().manual_seed(hash(name) % (2 ** 31)), dtype=torch.float64)
            names.append(name); norms.append(float(g.norm())); projs.append(float(g @ v))
        out[f"{case}_names"] = np.array(names)
        out[f"{case}_norm"] = np.array(norms)
        print(case, "loss", loss.item(), "params", len(names))
        # python's hash() of str is salted per process: store the projection seeds explicitly
        out[f"{case}_seed"] = np.array([hash(n) % (2 ** 31) for n in names], np.int64)
        out[f"{case}_proj"] = np.array(projs)
    np.savez_compressed(os.path.join(OUT, "train_grads.npz"), **out)


if __name__ == "__main__":
    main()
