"""GPU parity at the BASELINE horizons (VERDICT r01 "pin parity at the real configurations").

Goldens: tests/golden/long_*.npz and hist_*.npz, written by oracle/make_golden_long.py from the unmodified reference
on CPU fp32 (DDPM 400 steps on Dataset 2, guarded Heun 50 steps = 99 evaluations on Dataset 3, DDim 200 on the HGCal
fixture shape, DDim 50 on Dataset-1 photons through the in-loop converter; 512 / 128-shower feature sets).
Every gate is 2-3x the error measured on B200 (the measured curve is written to gpurun_out/ and committed under
profiles/r02_error_vs_step.txt); fp16 activations re-rounded at every layer make the error grow slowly with the number of
steps, which is exactly what these tests watch.
"""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, ROOT, load_case, noise_bank, rel_l2

pytestmark = pytest.mark.gpu

# measured on B200 (profiles/r02_error_vs_step.txt): worst recorded step / final = 1.62e-3 / 7.3e-4 (ds2 DDPM 400),
# 5.7e-4 / 2.5e-4 (ds3 Heun 50), 9.4e-4 / 3.6e-4 (hgcal DDim 200), 6.9e-4 / 4.8e-4 (ds1 DDim 50).  The error does NOT grow
# with the number of steps (the x0 error is largest at the high-noise start and the sampler contracts it); gates ~2.5x measured.
TOL_LONG = {"ds2_ddpm400": 4e-3, "ds3_heun50": 1.5e-3, "hgcal_ddim200": 2.5e-3, "ds1_ddim50": 2e-3}
TOL_FINAL = {"ds2_ddpm400": 2e-3, "ds3_heun50": 8e-4, "hgcal_ddim200": 1e-3, "ds1_ddim50": 1.3e-3}
TOL_ROWWISE = 3e-3     # two fp16 runs with different tile alignment: measured worst per-shower 1.22e-3


def _model(case, sampler, max_batch):
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    c = load_case(case)
    cfg = dict(c["cfg"], SAMPLER=sampler)
    m = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=max_batch)
    m.load_state_dict(dict(c["sd"]))
    return m, c


def _long_inputs(c, B, seed):
    g = torch.Generator().manual_seed(seed)
    E = torch.rand(B, c["meta"]["cond_e"], generator=g)
    layers = torch.randn(B, c["meta"]["cond_layers"], generator=g) if c["meta"]["cond_layers"] else None
    x = torch.randn([B] + list(c["shape"]), generator=torch.Generator().manual_seed(seed + 1))
    return E, layers, x


def _curve(name, idx, e_x0, e_x, e_final):
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "error_vs_step.txt"), "a") as f:
        f.write(f"# {name}: rel-L2 vs the live-reference golden (CPU fp32) per recorded step\n")
        for i, a, b in zip(idx, e_x0, e_x):
            f.write(f"{name} step {int(i):4d}  x0 {a:.3e}  x {b:.3e}\n")
        f.write(f"{name} final {e_final:.3e}\n")


@pytest.mark.parametrize("name", ["ds2_ddpm400", "ds3_heun50", "hgcal_ddim200"])
def test_long_horizon_sampling(name):
    from calodiffusion_b200 import sample as S
    with open(os.path.join(GOLDEN, f"long_{name}_meta.json")) as f:
        meta = json.load(f)
    gold = np.load(os.path.join(GOLDEN, f"long_{name}.npz"))
    sampler = "Heun" if meta["sampler"] == "HeunGuarded" else meta["sampler"]
    B, steps = meta["B"], meta["steps"]
    m, c = _model(meta["case"], sampler, B)
    E, layers, x = _long_inputs(c, B, meta["input_seed"])
    s = getattr(S, sampler)(c["cfg"])
    if meta["sampler"] == "HeunGuarded":
        s.last_step_euler = True
    s.noise_bank = torch.stack(noise_bank(c["shape"], B, steps, seed=meta["noise_seed"])).cuda()
    xf, xs, x0s = s(m, x.cuda(), E.cuda(), None if layers is None else layers.cuda(), steps, 0, True)
    assert len(xs) == steps and len(x0s) == steps
    idx = gold["snap_idx"]
    e_x0 = [rel_l2(x0s[i].cpu().numpy(), gold["x0s"][k]) for k, i in enumerate(idx)]
    e_x = [rel_l2(xs[i].cpu().numpy(), gold["xs"][k]) for k, i in enumerate(idx)]
    e_f = rel_l2(xf.cpu().numpy(), gold["x"])
    _curve(name, idx, e_x0, e_x, e_f)
    print(name, "x0 errors", [f"{e:.2e}" for e in e_x0], "final", f"{e_f:.2e}")
    assert np.isfinite(xf.cpu().numpy()).all()
    assert max(e_x0 + e_x + [e_f]) < TOL_LONG[name], (e_x0, e_x, e_f)
    assert e_f < TOL_FINAL[name], e_f


def test_long_horizon_ds1_in_loop_converter():
    """Dataset-1 photons, DDim 50 steps, the geometry maps applied inside every evaluation (BASELINE config 1)."""
    import tempfile
    from conftest import write_binning_xml
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    from calodiffusion_b200 import sample as S
    c = load_case("ds1")
    gold = np.load(os.path.join(GOLDEN, "long_ds1_ddim50.npz"))
    path = os.path.join(tempfile.mkdtemp(prefix="cd_bins_"), "binning.xml")
    cfg = dict(c["cfg"], SAMPLER="DDim", BIN_FILE=write_binning_xml(path, c["meta"]))
    B, steps = 4, 50
    m = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=B)
    sd = {k: v for k, v in m.state_dict().items()}
    sd.update(c["sd"])
    g = c["gold"]
    for i in range(len(g["lay_alphas"])):
        sd[f"NN_embed.encs.{i}.weight"] = torch.from_numpy(g[f"enc_w_{i}"])
        sd[f"NN_embed.decs.{i}.weight"] = torch.from_numpy(g[f"dec_w_{i}"])
    m.load_state_dict(sd)
    gen = torch.Generator().manual_seed(11)
    E = torch.rand(B, 1, generator=gen)
    layers = torch.randn(B, 6, generator=gen)
    x = torch.randn(B, 368, generator=torch.Generator().manual_seed(12))
    s = S.DDim(cfg)
    s.noise_bank = torch.stack(noise_bank([368], B, steps, seed=13)).cuda()
    xf, xs, x0s = s(m, x.cuda(), E.cuda(), layers.cuda(), steps, 0, True)
    idx = gold["snap_idx"]
    e_x0 = [rel_l2(x0s[i].cpu().numpy(), gold["x0s"][k]) for k, i in enumerate(idx)]
    e_x = [rel_l2(xs[i].cpu().numpy(), gold["xs"][k]) for k, i in enumerate(idx)]
    e_f = rel_l2(xf.cpu().numpy(), gold["x"])
    _curve("ds1_ddim50", idx, e_x0, e_x, e_f)
    print("ds1_ddim50 x0 errors", [f"{e:.2e}" for e in e_x0], "final", f"{e_f:.2e}")
    assert max(e_x0 + e_x + [e_f]) < TOL_LONG["ds1_ddim50"] and e_f < TOL_FINAL["ds1_ddim50"]


def test_bench_batch_rowwise():
    """One Dataset-2 evaluation at the bench batch (B = 1000: every persistent CTA wraps over many tiles and sample
    boundaries) equals, shower by shower, the same inputs run as 250 batches of 4."""
    m, c = _model("ds2", "DDPM", 1000)
    B = 1000
    g = torch.Generator().manual_seed(5)
    x = torch.randn([B] + list(c["shape"]), generator=g).cuda()
    E = torch.rand(B, 1, generator=g).cuda()
    L = torch.randn(B, 46, generator=g).cuda()
    sig = torch.exp(torch.randn(B, generator=g) * 1.2 - 1.2).cuda()
    x = x * (1 + sig.view(-1, 1, 1, 1, 1) ** 2).sqrt()
    full = m.denoise(x, E=E, sigma=sig, layers=L)
    assert torch.isfinite(full).all()
    worst = 0.0
    for lo in range(0, B, 4):
        part = m.denoise(x[lo:lo + 4], E=E[lo:lo + 4], sigma=sig[lo:lo + 4], layers=L[lo:lo + 4])
        d = (part - full[lo:lo + 4]).flatten(1).norm(dim=1) / full[lo:lo + 4].flatten(1).norm(dim=1)
        worst = max(worst, float(d.max()))
    print(f"B=1000 vs 250 x B=4: worst per-shower rel-L2 {worst:.2e}")
    assert worst < TOL_ROWWISE
    # and a handful of rows against the CPU oracle (first, a CTA-wrap in the middle, last)
    from oracle.unet_oracle import DenoiserOracle
    o = DenoiserOracle(c["cfg"], c["sd"])
    rows = [0, 1, 499, 500, 998, 999]
    with torch.no_grad():
        ref = o.denoise(x[rows].cpu(), E[rows].cpu(), sig[rows].cpu(), L[rows].cpu())
    e = rel_l2(full[rows].cpu().numpy(), ref.numpy())
    print(f"B=1000 rows {rows} vs oracle: {e:.2e}")
    assert e < 2e-3


def _hist_inputs(shape, n, cond_e, cond_layers, seed):
    g = torch.Generator().manual_seed(seed)
    E = torch.rand(n, cond_e, generator=g)
    layers = torch.randn(n, cond_layers, generator=g) if cond_layers else None
    start = torch.randn([n] + list(shape), generator=g)
    return E, layers, start


@pytest.mark.parametrize("name", ["hist_ds2", "hist_ds3"])
def test_feature_histograms_vs_reference(name):
    """north_star check 3 at a size that can reject: N showers (512 on Dataset 2 / DDim 50, 128 on Dataset 3 / guarded
    Heun 12) sampled by the CUDA path vs the reference's (features stored in tests/golden/hist_*.npz):
      (a) same start noise and conditioning: per-shower features agree, histograms coincide;
      (b) independent start noise and conditioning: no feature's KS test rejects, and the separation power is within
          the spread of two independent reference samples (sets A and B of the golden)."""
    from calodiffusion_b200 import sample as S
    from calodiffusion_b200.postprocess import reverse_norm_device
    from features import shower_features, separation_power, ks_pvalue
    with open(os.path.join(GOLDEN, f"{name}_meta.json")) as f:
        meta = json.load(f)
    gold = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    N, steps = meta["N"], meta["steps"]
    sampler = "Heun" if meta["sampler"] == "HeunGuarded" else meta["sampler"]
    m, c = _model(meta["case"], sampler, N)
    ds = 2 if meta["case"] == "ds2" else 3
    smap = c["cfg"]["SHOWERMAP"]
    s = getattr(S, sampler)(c["cfg"])
    if meta["sampler"] == "HeunGuarded":
        s.last_step_euler = True

    def sample_features(seed):
        E, layers, start = _hist_inputs(c["shape"], N, c["meta"]["cond_e"], c["meta"]["cond_layers"], seed)
        Lc = None if layers is None else layers.cuda()
        x, _, _ = s(m, start.cuda(), E.cuda(), Lc, steps, 0, False)
        data, e = reverse_norm_device(x, E.cuda(), layerE=Lc, showerMap=smap, dataset_num=ds, **meta["reverse_norm"])
        return x, shower_features(data.cpu().numpy(), e.cpu().numpy(), c["shape"][1:])

    x_same, f_same = sample_features(meta["seeds"]["A"])
    e8 = rel_l2(x_same[:8].cpu().numpy(), gold["A_x_first8"])
    print(name, f"first 8 showers vs reference rel-L2 {e8:.2e}")
    assert e8 < 6e-3
    _, f_ind = sample_features(77)
    feats = [k[2:] for k in gold.files if k.startswith("A_") and k != "A_x_first8"]
    worst_same, worst_p, worst_ratio = 0.0, 1.0, 0.0
    lines = []
    for k in feats:
        fa, fb = gold[f"A_{k}"], gold[f"B_{k}"]
        sep_same = separation_power(fa, f_same[k])
        sep_ref = separation_power(fa, fb)              # two independent reference samples
        sep_ind = separation_power(fa, f_ind[k])
        p_ind, p_ref = ks_pvalue(fa, f_ind[k]), ks_pvalue(fa, fb)
        worst_same, worst_p = max(worst_same, sep_same), min(worst_p, p_ind)
        worst_ratio = max(worst_ratio, sep_ind / max(sep_ref, 5.0 / N))
        lines.append(f"{name} {k:18s} same-start sep {sep_same:.2e} | independent: KS p {p_ind:.3f} (ref-vs-ref {p_ref:.3f}) "
                     f"sep {sep_ind:.3e} (ref-vs-ref {sep_ref:.3e})")
    print("\n".join(lines))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "feature_histograms.txt"), "a") as f:
        f.write("\n".join(lines) + "\n")
    # (a) same start: a handful of the N showers may cross a bin edge (each crossing of a ~N/20-entry bin costs ~1/N)
    assert worst_same < 40.0 / N, worst_same
    # (b) independent: 13 features, p > 1e-3 each; separation no worse than 4x the reference-vs-reference spread
    assert worst_p > 1e-3, worst_p
    assert worst_ratio < 4.0, worst_ratio


@pytest.mark.parametrize("name,steps", [("DDim", 3), ("DDPM", 4), ("Euler", 3)])
def test_reference_sampler_objects_drive_this_model(name, steps):
    """The drop-in claim of SURVEY 8(b): the REFERENCE's own sampler classes (oracle/_ref/calodiffusion/models/sample.py,
    unmodified) drive this backend's model through `model(x, sigma=, E=, layers=)` / `model.denoise(...)`, and the result
    matches what the same sampler produced with the reference model (tests/golden/ds2.npz / ds3.npz)."""
    from oracle import ref_shim
    if not ref_shim.available():
        pytest.skip("oracle/_ref not staged (run oracle/build_ref.sh in the build container)")
    ref_shim.install()
    from calodiffusion.models import sample as ref_sample
    from oracle.make_golden import ReplayNoise
    case = "ds3" if name == "Euler" else "ds2"
    m, c = _model(case, name, 2)
    s = getattr(ref_sample, name)(c["cfg"])
    bank = [b.cuda() for b in noise_bank(c["shape"], c["B"], steps)]
    Lc = None if c["layers"] is None else c["layers"].cuda()
    with ReplayNoise(bank):
        x, xs, x0s = s(m, c["x"].cuda(), c["E"].cuda(), Lc, steps, 0, True)
    e0 = rel_l2(torch.stack(x0s).cpu().numpy(), c["gold"][f"{name}_x0s"])
    e2 = rel_l2(x.cpu().numpy(), c["gold"][f"{name}_x"])
    print(f"reference {name} object on the B200 model: x0s {e0:.2e} final {e2:.2e}")
    assert max(e0, e2) < 4e-3
