"""GPU parity: LayerDiffusion's layer-energy denoiser (dense ResNet) through the C ABI, against the live
reference's goldens (fp32 kernel: rel-L2 <= 1e-4), and the two-stage `sample()` against the oracle chain."""
import numpy as np
import pytest
import torch

from conftest import load_case, noise_bank, rel_l2
from test_layer_oracle import load_layer_case

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-4


@pytest.fixture(scope="module")
def ld():
    from calodiffusion_b200.layerdiffusion import LayerDiffusion
    c = load_layer_case()
    base = load_case("ds2")
    cfg = dict(base["cfg"], SAMPLER="DDim", LAYER_SAMPLER="DDim", LAYER_STEPS=3)
    m = LayerDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=8)
    m.layer_model.load_state_dict(c["sd"])
    m.base_model.load_state_dict({k.removeprefix("model."): v for k, v in base["sd"].items()})
    return m, c, base, cfg


def test_layer_denoise_and_forward_match_reference(ld):
    m, c, _, _ = ld
    m.set_layer_state(True)
    try:
        errs = []
        for tag in ("a", "b"):
            sig = torch.from_numpy(c["gold"][f"denoise_{tag}_sigma"])
            y = m.denoise((c["x"] * (1 + sig.view(-1, 1) ** 2).sqrt()).cuda(), E=c["E"].cuda(), sigma=sig.cuda().view(-1, 1), layers=None)
            assert tuple(y.shape) == (5, 46)
            errs.append(rel_l2(y.cpu().numpy(), c["gold"][f"denoise_{tag}"]))
        t = torch.tensor([0.2, -0.7, 1.1, 0.0, -2.0])
        ef = rel_l2(m.layer_forward(c["x"].cuda(), c["E"].cuda(), t.cuda()).cpu().numpy(), c["gold"]["forward"])
    finally:
        m.set_layer_state(False)
    print("layer denoise", [f"{e:.2e}" for e in errs], "forward", f"{ef:.2e}")
    assert max(errs + [ef]) < TOL_FP32


@pytest.mark.parametrize("graph", [False, True])
def test_layer_ddim_sampling(ld, graph):
    from calodiffusion_b200 import sample as S
    m, c, _, cfg = ld
    m.use_graph = graph
    m.set_layer_state(True)
    try:
        s = S.DDim(cfg)
        s.noise_bank = torch.stack(c["bank"]).cuda()
        x, xs, x0s = s(m, c["x"].cuda(), c["E"].cuda(), None, 3, 0, True)
    finally:
        m.set_layer_state(False)
    e0 = rel_l2(torch.stack(x0s).cpu().numpy(), c["gold"]["DDim_x0s"])
    e1 = rel_l2(torch.stack(xs).cpu().numpy(), c["gold"]["DDim_xs"])
    e2 = rel_l2(x.cpu().numpy(), c["gold"]["DDim_x"])
    print("layer DDim", "graph" if graph else "eager", f"x0s {e0:.2e} xs {e1:.2e} final {e2:.2e}")
    assert max(e0, e1, e2) < TOL_FP32


def test_two_stage_sample_matches_oracle_chain(ld):
    """LayerDiffusion.sample: layer energies from the layer model, then showers conditioned on them (both on the device),
    against the same chain on the oracle with shared noise banks."""
    from calodiffusion_b200 import sample as S
    from oracle.layer_oracle import LayerDenoiserOracle
    from oracle.unet_oracle import DenoiserOracle, sample_ddim
    m, c, base, cfg = ld
    m.use_graph = True
    B = 2
    E = base["E"][:B]
    g = torch.Generator().manual_seed(31)
    lstart = torch.randn(B, 46, generator=g)
    lbank = [torch.randn(B, 46, generator=g) for _ in range(3)]
    bank = noise_bank(base["shape"], B, 2, seed=17)
    # device: the pieces LayerDiffusion.sample chains, with replayed noise
    m.set_layer_state(True)
    try:
        m.layer_sampler.noise_bank = torch.stack(lbank).cuda()
        layers, _, _ = m.layer_sampler(m, lstart.cuda(), E.cuda(), None, 3, 0, False)
    finally:
        m.set_layer_state(False)
        m.layer_sampler.noise_bank = None
    s = S.DDim(cfg)
    s.noise_bank = torch.stack(bank).cuda()
    x, _, _ = s(m, base["x"][:B].cuda(), E.cuda(), layers, 2, 0, False)
    # oracle chain
    lo = LayerDenoiserOracle(cfg, c["sd"])
    uo = DenoiserOracle(base["cfg"], base["sd"])
    with torch.no_grad():
        lref, _, _ = sample_ddim(lo, lstart, E, None, 3, 0.0, lbank)
        xref, _, _ = sample_ddim(uo, base["x"][:B], E, lref, 2, 0.0, bank)
    el, ex = rel_l2(layers.cpu().numpy(), lref.numpy()), rel_l2(x.cpu().numpy(), xref.numpy())
    print(f"two-stage: layers {el:.2e} showers {ex:.2e}")
    assert el < TOL_FP32 and ex < 1e-2
    # public API: dict result with the reference's keys, finite showers, both engines alive side by side
    out = m.sample(E.cuda(), None, num_steps=2, return_layers=True)
    assert set(out) == {"x", "layers"} and out["x"].shape == (B, 1, 45, 16, 9) and np.isfinite(out["x"]).all()
    assert tuple(out["layers"].shape) == (B, 46)


def test_growing_batch_rebuilds_each_engine(ld):
    """a small warm-up batch followed by a larger one.  Each of the two engines (layer model, U-Net) is rebuilt for
    the larger batch on its own (`_max_batch` is per handle), so the second call must not fail with 'batch outside max_batch'."""
    m, c, base, cfg = ld
    for B in (4, 12):      # the fixture was built with max_batch = 8: 12 forces a rebuild of both handles
        E = torch.rand(B, 1, generator=torch.Generator().manual_seed(B)).cuda()
        out = m.sample(E, None, num_steps=2, return_layers=True)
        assert out["x"].shape[0] == B and np.isfinite(out["x"]).all() and tuple(out["layers"].shape) == (B, 46)
