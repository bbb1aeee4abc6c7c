"""CPU: the oracle restatement must reproduce the live reference's golden outputs (fp32, tol 2e-5)."""
import numpy as np
import torch

from conftest import load_case, noise_bank, rel_l2
from oracle.unet_oracle import DenoiserOracle, cylindrical_conv, lms_coeff, sample_consistency, sample_ddim, sample_edm, sample_lms

TOL = 2e-5  # fp32 CPU vs fp32 CPU, different op order only


def test_cylconv_known_answer():
    """reference calodiffusion/tests/test_cyl_conv.py:7-25: rows [9,18,15] on every phi row."""
    import os
    from conftest import GOLDEN
    kat = np.load(os.path.join(GOLDEN, "cylconv_kat.npz"))
    x = torch.from_numpy(kat["x"])
    w = torch.ones(1, 1, 1, 3, 3)
    y = cylindrical_conv(x, w, None, padding=(0, 1, 1)).numpy()
    assert np.array_equal(y, kat["y"])
    assert y.reshape(4, 3).tolist() == [[9.0, 18.0, 15.0]] * 4


def test_denoise_matches_reference(case):
    name, c = case
    m = DenoiserOracle(c["cfg"], c["sd"])
    for tag in ("a", "b"):
        sig = torch.from_numpy(c["gold"][f"denoise_{tag}_sigma"])
        scale = (1 + sig.view(-1, 1, 1, 1, 1) ** 2).sqrt()
        with torch.no_grad():
            y = m.denoise(c["x"] * scale, c["E"], sig, c["layers"]).numpy()
        assert rel_l2(y, c["gold"][f"denoise_{tag}"]) < TOL, (name, tag)


def test_ddpm_ddim_match_reference():
    c = load_case("ds2")
    m = DenoiserOracle(c["cfg"], c["sd"])
    for name, steps, eta in (("DDPM", 4, 1.0), ("DDim", 3, 0.0)):
        bank = noise_bank(c["shape"], c["B"], steps)
        with torch.no_grad():
            x, xs, x0s = sample_ddim(m, c["x"], c["E"], c["layers"], steps, eta, bank)
        assert rel_l2(torch.stack(x0s).numpy(), c["gold"][f"{name}_x0s"]) < TOL
        assert rel_l2(torch.stack(xs).numpy(), c["gold"][f"{name}_xs"]) < TOL
        assert rel_l2(x.numpy(), c["gold"][f"{name}_x"]) < TOL


def test_edm_euler_heun_match_reference():
    c = load_case("ds3")
    m = DenoiserOracle(c["cfg"], c["sd"])
    for name, order in (("Euler", 1), ("Heun", 2)):
        bank = noise_bank(c["shape"], c["B"], 3)
        with torch.no_grad():
            x, xs, x0s = sample_edm(m, c["x"], c["E"], None, 3, order, bank)
        assert rel_l2(torch.stack(x0s).numpy(), c["gold"][f"{name}_x0s"]) < 5e-5
        assert rel_l2(torch.stack(xs).numpy(), c["gold"][f"{name}_xs"]) < 5e-5
        if name == "Euler":
            assert rel_l2(x.numpy(), c["gold"]["Euler_x"]) < 5e-5
        else:  # reference defect: Heun's last step divides by t_next = 0
            assert not np.isfinite(c["gold"]["Heun_x"]).any()
            assert not np.isfinite(x.numpy()).any()


def test_hgcal_ddim_matches_reference():
    c = load_case("hgcal")
    m = DenoiserOracle(c["cfg"], c["sd"])
    bank = noise_bank(c["shape"], c["B"], 2)
    with torch.no_grad():
        x, xs, x0s = sample_ddim(m, c["x"], c["E"], c["layers"], 2, 0.0, bank)
    assert rel_l2(torch.stack(x0s).numpy(), c["gold"]["DDim_x0s"]) < TOL
    assert rel_l2(x.numpy(), c["gold"]["DDim_x"]) < TOL


def test_lms_matches_reference():
    """tests/golden/lms.npz (oracle/make_golden_lms.py): the live reference's LMS, 6 steps, ORDER 4 (default) and 2."""
    import os
    from conftest import GOLDEN
    gold = np.load(os.path.join(GOLDEN, "lms.npz"))
    c = load_case("ds2")
    m = DenoiserOracle(c["cfg"], c["sd"])
    for order in (4, 2):
        with torch.no_grad():
            x, xs, x0s = sample_lms(m, c["x"], c["E"], c["layers"], 6, order)
        assert xs == [] and x0s == []
        assert rel_l2(x.numpy(), gold[f"LMS_order{order}_x"]) < 5e-5, order


def test_lms_coefficients():
    """Closed forms: order 1 is the step h; order 2 is Adams-Bashforth on a non-uniform grid; the basis sums to h."""
    t = np.array([80.0, 30.0, 9.0, 2.0, 0.3, 0.0])
    for i in range(1, 5):
        h = t[i + 1] - t[i]
        assert abs(lms_coeff(1, t, i, 0) - h) < 1e-12
        hp = t[i] - t[i - 1]
        assert abs(lms_coeff(2, t, i, 0) - (h + h * h / (2 * hp))) < 1e-10
        assert abs(lms_coeff(2, t, i, 1) + h * h / (2 * hp)) < 1e-10
    for order in (3, 4):
        assert abs(sum(lms_coeff(order, t, 4, j) for j in range(order)) - (t[5] - t[4])) < 1e-10
    import pytest
    with pytest.raises(ValueError):
        lms_coeff(3, t, 1, 0)


def test_restart_default_is_euler():
    """The live reference's Restart with its default (string-keyed) RESTART_LIST never restarts (sample.py:877): its
    showers are the Euler loop's."""
    import os
    from conftest import GOLDEN
    gold = np.load(os.path.join(GOLDEN, "lms.npz"))
    c = load_case("ds2")
    m = DenoiserOracle(c["cfg"], c["sd"])
    with torch.no_grad():
        x, xs, x0s = sample_edm(m, c["x"], c["E"], c["layers"], 4, 1, noise_bank(c["shape"], c["B"], 4))
    assert rel_l2(torch.stack(x0s).numpy(), gold["Restart_x0s"]) < 5e-5
    assert rel_l2(x.numpy(), gold["Restart_x"]) < 5e-5


def test_dpm2_matches_reference():
    import os
    from conftest import GOLDEN
    gold = np.load(os.path.join(GOLDEN, "lms.npz"))
    c = load_case("ds2")
    m = DenoiserOracle(c["cfg"], c["sd"])
    with torch.no_grad():
        x, xs, x0s = sample_edm(m, c["x"], c["E"], c["layers"], 4, "dpm2", noise_bank(c["shape"], c["B"], 4))
    assert rel_l2(torch.stack(x0s).numpy(), gold["DPM2_x0s"]) < 5e-5
    assert rel_l2(torch.stack(xs).numpy(), gold["DPM2_xs"]) < 5e-5
    # reference defect: the last step's t_mid is 0
    assert not np.isfinite(gold["DPM2_x"]).all() and not np.isfinite(x.numpy()).all()


def test_consistency_matches_reference():
    import os
    from conftest import GOLDEN
    gold = np.load(os.path.join(GOLDEN, "lms.npz"))
    c = load_case("ds2")
    m = DenoiserOracle(c["cfg"], c["sd"])
    with torch.no_grad():
        x, xs, x0 = sample_consistency(m, c["x"], c["E"], c["layers"], 5, 100, noise_bank(c["shape"], c["B"], 5))
    assert rel_l2(torch.stack(xs).numpy(), gold["Consistency_xs"]) < 5e-5
    assert rel_l2(x0.numpy(), gold["Consistency_x0_last"]) < 5e-5
    assert rel_l2(x.numpy(), gold["Consistency_x"]) < 5e-5
