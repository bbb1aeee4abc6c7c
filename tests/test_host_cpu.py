"""CPU tests of the host logic: C-ABI exports, parameter naming, and the sampler -> cd_eval programs
(interpreted on CPU with the oracle denoiser and compared with the live reference's golden outputs)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, load_case, noise_bank, rel_l2

from calodiffusion_b200 import _lib
from calodiffusion_b200.calodiffusion import unet_param_shapes
from calodiffusion_b200 import sample as b200_sample
from calodiffusion_b200 import utils as b200_utils
from oracle.unet_oracle import DenoiserOracle, r_bin_centres


def test_library_exports_every_declared_symbol():
    """The built .so must export every function include/calodiffusion_b200.h declares (no compute calls)."""
    header = open(os.path.join(ROOT, "include", "calodiffusion_b200.h")).read()
    declared = set(re.findall(r"\b(cd_[a-z_0-9]+)\s*\(", header))
    declared -= {"cd_unet", "cd_eval", "cd_unet_desc"}
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    assert os.path.exists(_lib.LIB_PATH), "build the library first: calodiffusion_b200/csrc/build.sh"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert _lib.load().cd_version() == 1


def test_struct_sizes_match_header(tmp_path):
    """ctypes mirrors vs the real header, measured by compiling it with gcc (plain C, so the header stays C)."""
    import subprocess
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "calodiffusion_b200.h"\n'
                   'int main(){printf("%zu %zu %zu %zu", sizeof(cd_eval), sizeof(cd_unet_desc), '
                   'offsetof(cd_eval, pre_a), offsetof(cd_unet_desc, r_centres));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    assert got == [ctypes.sizeof(_lib.Eval), ctypes.sizeof(_lib.UnetDesc), _lib.Eval.pre_a.offset,
                   _lib.UnetDesc.r_centres.offset]


@pytest.mark.parametrize("case", ["ds2", "ds3", "hgcal"])
def test_param_names_match_reference_state_dict(case):
    c = load_case(case)
    cfg = c["cfg"]
    inch = 1 + 2 * bool(cfg.get("R_Z_INPUT")) + bool(cfg.get("PHI_INPUT"))
    cs = (2 + cfg["SHAPE_FINAL"][2] if "layer" in cfg.get("SHOWERMAP", "") else 1) + (2 if cfg.get("HGCAL") else 0)
    sp = unet_param_shapes(cfg["LAYER_SIZE_UNET"], inch, cs, cfg["COND_SIZE_UNET"], cfg["SHAPE_FINAL"][2:])
    mine = {"model." + k: list(v) for k, v in sp.items()}
    assert list(mine.keys()) == list(c["meta"]["shapes"].keys())
    assert mine == c["meta"]["shapes"]


def test_r_centres_and_registry():
    for ds, n in ((2, 9), (3, 18), (1, 30), (111, 21)):
        want = np.asarray(r_bin_centres(ds, n), np.float32)
        assert np.allclose(b200_utils.scaled_r_centres(ds, n), want / want[-1], rtol=1e-7)
    with pytest.raises(ValueError):
        b200_utils.scaled_r_centres(2, 10)
    with pytest.raises(ValueError):
        b200_utils.load_attr("sampler", "NoSuchSampler")
    assert b200_utils.load_attr("sampler", "DDPM") is b200_sample.DDPM


class _InterpretedModel:
    """Executes a cd_eval program on CPU with the oracle denoiser (same semantics as cd_sample_run)."""

    def __init__(self, oracle):
        self.o = oracle

    def denoise(self, x, E=None, sigma=None, layers=None):
        return self.o.denoise(x, E, sigma, layers)

    def run_program(self, start, E, layers, evals, start_scale, final_buf, noise_bank=None, seed=0, shower_offset=0,
                    n_rec_x=0, n_rec_x0=0):
        buf = [None] * 8
        buf[0] = start * np.float32(start_scale)
        xs, x0s = [None] * n_rec_x, [None] * n_rec_x0
        B = start.shape[0]
        for e in evals:
            recorded = False
            if e.pre_b != 0.0 or e.pre_src >= 0:
                src = e.pre_src if e.pre_src >= 0 else e.x_in
                if e.rec_x >= 0 and e.rec_x_buf == e.x_in and src == e.x_in:
                    xs[e.rec_x], recorded = buf[e.rec_x_buf], True      # recorded before the churn noise
                noise = noise_bank[e.pre_noise_index] if e.pre_b != 0.0 else 0.0
                buf[e.x_in] = e.pre_a * buf[src] + e.pre_b * noise
            x0 = self.o.denoise(buf[e.x_in], E, torch.full((B,), e.sigma), layers)
            if e.x0_buf >= 0:
                buf[e.x0_buf] = x0
            if e.rec_x0 >= 0:
                x0s[e.rec_x0] = x0
            if e.d_out >= 0:
                buf[e.d_out] = e.d_scale * (buf[e.x_in] - x0)
            if e.x_out >= 0:
                with np.errstate(all="ignore"):
                    out = e.k[0] * x0 + e.k[1] * buf[e.x_in]
                    if e.a1 >= 0:
                        out = out + e.k[2] * buf[e.a1]
                    elif e.a1 == -2 and e.k[2] != 0.0:
                        out = out + e.k[2] * noise_bank[e.noise_index]
                    if e.a2 >= 0:
                        out = out + e.k[3] * buf[e.a2]
                    if e.a3 >= 0:
                        out = out + e.k[4] * buf[e.a3]
                buf[e.x_out] = out
            if e.rec_x >= 0 and not recorded:
                xs[e.rec_x] = buf[e.rec_x_buf]
        return buf[final_buf], xs, x0s


@pytest.mark.parametrize("name,steps", [("DDPM", 4), ("DDim", 3)])
def test_ddpm_program_reproduces_reference(name, steps):
    c = load_case("ds2")
    model = _InterpretedModel(DenoiserOracle(c["cfg"], c["sd"]))
    sampler = getattr(b200_sample, name)(c["cfg"])
    sampler.noise_bank = noise_bank(c["shape"], c["B"], steps)
    with torch.no_grad():
        x, xs, x0s = sampler(model, c["x"], c["E"], c["layers"], steps, 0, True)
    assert rel_l2(torch.stack(x0s).numpy(), c["gold"][f"{name}_x0s"]) < 5e-5
    assert rel_l2(torch.stack(xs).numpy(), c["gold"][f"{name}_xs"]) < 5e-5
    assert rel_l2(x.numpy(), c["gold"][f"{name}_x"]) < 5e-5


@pytest.mark.parametrize("name", ["Euler", "Heun"])
def test_edm_program_reproduces_reference(name):
    c = load_case("ds3")
    model = _InterpretedModel(DenoiserOracle(c["cfg"], c["sd"]))
    sampler = getattr(b200_sample, name)(c["cfg"])
    sampler.noise_bank = noise_bank(c["shape"], c["B"], 3)
    with torch.no_grad():
        x, xs, x0s = sampler(model, c["x"], c["E"], None, 3, 0, True)
    assert rel_l2(torch.stack(x0s).numpy(), c["gold"][f"{name}_x0s"]) < 1e-4
    assert rel_l2(torch.stack(xs).numpy(), c["gold"][f"{name}_xs"]) < 1e-4
    if name == "Euler":
        assert rel_l2(x.numpy(), c["gold"]["Euler_x"]) < 1e-4
    else:
        assert not np.isfinite(x.numpy()).all()  # reference defect reproduced


@pytest.mark.parametrize("order", [4, 2])
def test_lms_program_reproduces_reference(order):
    """The LMS class (a cd_eval program with a derivative ring) with the oracle denoiser vs the live reference's LMS."""
    import os
    from conftest import GOLDEN
    gold = np.load(os.path.join(GOLDEN, "lms.npz"))
    c = load_case("ds2")
    model = _InterpretedModel(DenoiserOracle(c["cfg"], c["sd"]))
    sampler = b200_utils.load_attr("sampler", "LMS")(dict(c["cfg"], SAMPLER_OPTIONS={"ORDER": order}))
    assert sampler.order == order
    with torch.no_grad():
        x, xs, x0s = sampler(model, c["x"], c["E"], c["layers"], 6, 0, True)
    assert xs == [] and x0s == []
    assert rel_l2(x.numpy(), gold[f"LMS_order{order}_x"]) < 5e-5
    with pytest.raises(ValueError):
        b200_sample.linear_multistep_coeff(3, [3.0, 2.0, 1.0, 0.0], 1, 0)   # utils/sampling.py:78-79


@pytest.mark.parametrize("name", ["Euler", "Heun", "DPM2"])
def test_churn_programs_reproduce_reference(name):
    """NOISY_SAMPLE churn (models/sample.py:639-662): x_hat = x_cur + sqrt(t_hat^2 - t_cur^2)*S_noise*noise before the
    evaluation, for the first- and second-order samplers, vs the live reference (tests/golden/churn.npz).  Heun's d' reads
    the un-noised x_cur (:820), the recorded xs are the un-noised states (:660)."""
    import os
    from conftest import GOLDEN
    gold = np.load(os.path.join(GOLDEN, "churn.npz"))
    c = load_case("ds2")
    cfg = dict(c["cfg"], NOISY_SAMPLE=True)
    model = _InterpretedModel(DenoiserOracle(cfg, c["sd"]))
    sampler = b200_utils.load_attr("sampler", name)(cfg)
    assert sampler.S_churn == 40
    sampler.noise_bank = noise_bank(c["shape"], c["B"], 5)
    with torch.no_grad():
        x, xs, x0s = sampler(model, c["x"], c["E"], c["layers"], 5, 0, True)
    ok = [i for i in range(5) if np.isfinite(gold[f"{name}_xs"][i]).all() and np.isfinite(gold[f"{name}_x0s"][i]).all()]
    assert len(ok) >= 4
    assert rel_l2(torch.stack([x0s[i] for i in ok]).numpy(), gold[f"{name}_x0s"][ok]) < 1e-4
    assert rel_l2(torch.stack([xs[i] for i in ok]).numpy(), gold[f"{name}_xs"][ok]) < 1e-4
    if name == "Euler":
        assert rel_l2(x.numpy(), gold["Euler_x"]) < 1e-4


def test_restart_program_reproduces_reference():
    import os
    from conftest import GOLDEN
    gold = np.load(os.path.join(GOLDEN, "lms.npz"))
    c = load_case("ds2")
    model = _InterpretedModel(DenoiserOracle(c["cfg"], c["sd"]))
    sampler = b200_utils.load_attr("sampler", "Restart")(c["cfg"])
    sampler.noise_bank = noise_bank(c["shape"], c["B"], 4)
    with torch.no_grad():
        x, xs, x0s = sampler(model, c["x"], c["E"], c["layers"], 4, 0, True)
    assert rel_l2(torch.stack(x0s).numpy(), gold["Restart_x0s"]) < 1e-4
    assert rel_l2(x.numpy(), gold["Restart_x"]) < 1e-4
    active = b200_sample.Restart(dict(c["cfg"], SAMPLER_OPTIONS={"RESTART_LIST": {1: [4, 1, 19.35, 40.79]}}))
    with pytest.raises(ValueError):
        active(model, c["x"], c["E"], c["layers"], 4, 0, False)


def test_dpm2_program_reproduces_reference():
    import os
    from conftest import GOLDEN
    gold = np.load(os.path.join(GOLDEN, "lms.npz"))
    c = load_case("ds2")
    model = _InterpretedModel(DenoiserOracle(c["cfg"], c["sd"]))
    sampler = b200_utils.load_attr("sampler", "DPM2")(c["cfg"])
    sampler.noise_bank = noise_bank(c["shape"], c["B"], 4)
    with torch.no_grad():
        x, xs, x0s = sampler(model, c["x"], c["E"], c["layers"], 4, 0, True)
    assert rel_l2(torch.stack(x0s).numpy(), gold["DPM2_x0s"]) < 1e-4
    assert rel_l2(torch.stack(xs).numpy(), gold["DPM2_xs"]) < 1e-4
    assert not np.isfinite(x.numpy()).all()  # reference defect reproduced: last step divides by t_mid = 0
    sampler.last_step_euler = True
    with torch.no_grad():
        xg, _, _ = sampler(model, c["x"], c["E"], c["layers"], 4, 0, False)
    assert np.isfinite(xg.numpy()).all()


@pytest.mark.parametrize("steps", [5, 1])
def test_consistency_program_reproduces_reference(steps):
    import os
    from conftest import GOLDEN
    from oracle.unet_oracle import sample_consistency
    gold = np.load(os.path.join(GOLDEN, "lms.npz"))
    c = load_case("ds2")
    o = DenoiserOracle(c["cfg"], c["sd"])
    sampler = b200_utils.load_attr("sampler", "Consistency")(c["cfg"])
    sampler.noise_bank = noise_bank(c["shape"], c["B"], 5)
    with torch.no_grad():
        x, xs, x0 = sampler(_InterpretedModel(o), c["x"], c["E"], c["layers"], steps, 0, True)
    assert len(xs) == steps
    if steps == 5:
        assert rel_l2(torch.stack(xs).numpy(), gold["Consistency_xs"]) < 1e-4
        assert rel_l2(x0.numpy(), gold["Consistency_x0_last"]) < 1e-4
        assert rel_l2(x.numpy(), gold["Consistency_x"]) < 1e-4
    else:   # single step: x = x0 at the top noise level, no draw
        with torch.no_grad():
            xo, _, _ = sample_consistency(o, c["x"], c["E"], c["layers"], 1, 100, None)
        assert rel_l2(x.numpy(), xo.numpy()) < 1e-4 and torch.equal(x, x0)


def test_no_cpu_fallback_without_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    c = load_case("ds2")
    cfg = dict(c["cfg"], SAMPLER="DDPM")
    with pytest.raises(RuntimeError):
        CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"])
