"""Golden vectors at the BASELINE horizons, from the LIVE reference (build container; several minutes of CPU).

    python oracle/make_golden_long.py [case ...]      # writes tests/golden/long_<case>.npz

The short goldens of make_golden.py pin single evaluations and 2-8 step loops at B = 2; these pin the loops the
benchmark actually runs (BASELINE.json configs): Dataset-2 DDPM 400 steps, Dataset-3 Heun 50 steps (99 evaluations;
last step Euler-guarded by a harness subclass of the reference's `Heun`, since the unguarded reference ends in 1/0),
the HGCal fixture shape DDim 200 and Dataset-1 photons DDim 50 through the in-loop converter.  Same recipe as
make_golden.py: seeded synthetic weights loaded into the UNMODIFIED reference `CaloDiffusion`, seeded inputs, the
reference's own sampler classes with torch.randn / randn_like replaying a seeded noise bank.  Stored: the final
showers and x0 / x snapshots every `stride` steps (inputs and noise are regenerated from the seeds by the tests).

`hist_ds2` / `hist_ds3` hold per-shower FEATURES (tests/features.py) of N showers sampled by the reference (DDim 50 /
guarded Heun 12) for two independent seeds: set A is re-sampled by the CUDA path from the same start noise (per-shower
agreement), set B calibrates what "two independent samples of the same distribution" looks like for the KS and
separation-power statistics.
"""
import json
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_shim  # noqa: E402
from oracle.make_golden import CASES, ReplayNoise, make_inputs, noise_bank  # noqa: E402
from oracle.unet_oracle import synth_state_dict  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")

LONG = {
    # name: (short-golden case supplying config + weights recipe, sampler, steps, B, snapshot stride)
    "ds2_ddpm400": ("ds2", "DDPM", 400, 4, 25),
    "ds3_heun50": ("ds3", "HeunGuarded", 50, 2, 10),
    "hgcal_ddim200": ("hgcal", "DDim", 200, 2, 25),
}


def guarded_heun_class():
    """Reference `Heun` with the final step (t_next = 0) taken as an Euler step (the guard of utils/sampling.py:232);
    every other step is the reference's own code.  Harness-side subclass: nothing under the reference tree changes."""
    from calodiffusion.models import sample as ref_sample

    class HeunGuarded(ref_sample.Heun):
        def in_loop_sampler(self):
            if float(self.t_next) == 0.0:
                d_cur = (self.x_hat - self.denoised) / self.t_hat
                return self.x_hat + (self.t_next - self.t_hat) * d_cur
            return super().in_loop_sampler()

    return HeunGuarded


def build_model(case, sampler="DDPM"):
    spec = CASES[case]
    model, cfg = ref_shim.load_reference_model(spec["config"], sampler, spec["overrides"])
    shapes = {k: list(v.shape) for k, v in model.state_dict().items()}
    model.load_state_dict(synth_state_dict(shapes, seed=0))
    return model, cfg, spec


def run(model, sampler_name, start, E, layers, steps, bank):
    from calodiffusion.utils import utils as ref_utils
    cls = guarded_heun_class() if sampler_name == "HeunGuarded" else ref_utils.load_attr("sampler", sampler_name)
    sampler = cls(model.config)
    with ReplayNoise(bank):
        return sampler(model, start, E, layers, steps, 0, True)


def long_case(name):
    case, sampler, steps, B, stride = LONG[name]
    model, cfg, spec = build_model(case)
    shape = cfg["SHAPE_PAD"][1:]
    E, layers, x = make_inputs(shape, B, spec["cond_e"], spec["cond_layers"], seed=11)
    bank = noise_bank(shape, B, steps, seed=13)
    t0 = time.time()
    with torch.no_grad():
        xf, xs, x0s = run(model, sampler, x, E, layers, steps, bank)
    idx = sorted(set(list(range(stride - 1, steps, stride)) + [0, steps - 1]))
    out = dict(x=xf.numpy(), snap_idx=np.array(idx, np.int32),
               x0s=np.stack([x0s[i].numpy() for i in idx]), xs=np.stack([xs[i].numpy() for i in idx]))
    np.savez_compressed(os.path.join(OUT, f"long_{name}.npz"), **out)
    with open(os.path.join(OUT, f"long_{name}_meta.json"), "w") as f:
        json.dump({"case": case, "sampler": sampler, "steps": steps, "B": B, "input_seed": 11, "noise_seed": 13,
                   "stride": stride}, f)
    print(name, f"{time.time() - t0:.0f}s", {k: v.shape for k, v in out.items()}, "finite:", bool(np.isfinite(out["x"]).all()))


def long_ds1():
    """Dataset-1 photons, DDim 50 steps through the reference's in-loop NNConverter (BASELINE config 1)."""
    xml = f"{ref_shim.REF_ROOT}/CaloChallenge/code/binning_dataset_1_photons.xml"
    model, cfg = ref_shim.load_reference_model("config_dataset1_photon.json", "DDim", {"BIN_FILE": xml})
    shapes = {k: list(v.shape) for k, v in model.state_dict().items() if k.startswith("model.")}
    sd = dict(model.state_dict())
    sd.update(synth_state_dict(shapes, seed=0))
    model.load_state_dict(sd)
    gold = np.load(os.path.join(OUT, "ds1.npz"))      # the converter weights the short golden stored (random 1e-5 init noise)
    with torch.no_grad():
        for i in range(model.NN_embed.gc.num_layers):
            model.NN_embed.encs[i].weight.copy_(torch.from_numpy(gold[f"enc_w_{i}"]))
            model.NN_embed.decs[i].weight.copy_(torch.from_numpy(gold[f"dec_w_{i}"]))
    B, steps = 4, 50
    g = torch.Generator().manual_seed(11)
    E = torch.rand(B, 1, generator=g)
    layers = torch.randn(B, 6, generator=g)
    x = torch.randn(B, 368, generator=torch.Generator().manual_seed(12))
    bank = noise_bank([368], B, steps, seed=13)
    with torch.no_grad():
        xf, xs, x0s = run(model, "DDim", x, E, layers, steps, bank)
    idx = [0, 9, 19, 29, 39, 49]
    np.savez_compressed(os.path.join(OUT, "long_ds1_ddim50.npz"), x=xf.numpy(), snap_idx=np.array(idx, np.int32),
                        x0s=np.stack([x0s[i].numpy() for i in idx]), xs=np.stack([xs[i].numpy() for i in idx]))
    print("ds1_ddim50", xf.shape, "finite:", bool(np.isfinite(xf.numpy()).all()))


HIST = {
    # name: (case, sampler, steps, N, batch)
    "hist_ds2": ("ds2", "DDim", 50, 512, 64),
    "hist_ds3": ("ds3", "HeunGuarded", 12, 128, 16),
}


def hist_inputs(shape, n, cond_e, cond_layers, seed):
    g = torch.Generator().manual_seed(seed)
    E = torch.rand(n, cond_e, generator=g)
    layers = torch.randn(n, cond_layers, generator=g) if cond_layers else None
    start = torch.randn([n] + list(shape), generator=g)
    return E, layers, start


def hist_case(name):
    from features import shower_features
    from oracle.reverse_norm import reverse_norm
    case, sampler, steps, N, bs = HIST[name]
    model, cfg, spec = build_model(case)
    shape = cfg["SHAPE_PAD"][1:]
    ds = 2 if case == "ds2" else 3
    smap = cfg["SHOWERMAP"]
    out = {}
    for tag, seed in (("A", 31), ("B", 32)):
        E, layers, start = hist_inputs(shape, N, spec["cond_e"], spec["cond_layers"], seed)
        xs = []
        t0 = time.time()
        with torch.no_grad():
            for lo in range(0, N, bs):
                sl = slice(lo, lo + bs)
                # DDim (eta 0) and Heun/Euler without churn multiply their per-step draw by zero: any finite bank will do
                bank = [torch.zeros([len(range(*sl.indices(N)))] + list(shape))] * (steps + 1)
                xf, _, _ = run(model, sampler, start[sl], E[sl], None if layers is None else layers[sl], steps, bank)
                xs.append(xf)
                print(name, tag, lo, f"{time.time() - t0:.0f}s", flush=True)
        x = torch.cat(xs).numpy()
        data, e = reverse_norm(x, E.numpy(), ds, smap, layerE=None if layers is None else layers.numpy(),
                               emax=1000.0, emin=1.0, max_deposit=2.0)
        f = shower_features(data, e, shape[1:])
        for k, v in f.items():
            out[f"{tag}_{k}"] = v
        out[f"{tag}_x_first8"] = x[:8]          # a few raw showers for a direct per-voxel check
    np.savez_compressed(os.path.join(OUT, f"{name}.npz"), **out)
    with open(os.path.join(OUT, f"{name}_meta.json"), "w") as f:
        json.dump({"case": case, "sampler": sampler, "steps": steps, "N": N, "seeds": {"A": 31, "B": 32},
                   "reverse_norm": dict(emax=1000.0, emin=1.0, max_deposit=2.0)}, f)
    print(name, "done")


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(os.cpu_count())
    want = sys.argv[1:] or list(LONG) + ["ds1_ddim50"] + list(HIST)
    for name in want:
        if name in LONG:
            long_case(name)
        elif name == "ds1_ddim50":
            long_ds1()
        elif name in HIST:
            hist_case(name)
        else:
            raise SystemExit(f"unknown case {name}")


if __name__ == "__main__":
    main()
