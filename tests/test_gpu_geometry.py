"""GPU parity: in-loop geometry embedding (BASELINE config 1) through the C ABI.

Dataset 1 photons: x is (B, 368); every denoiser evaluation runs NNConverter.enc -> cylindrical U-Net on (5,10,30) ->
NNConverter.dec on the device.  Checked against the live reference's goldens (tests/golden/ds1.npz) and the oracle.
Tolerance as in test_gpu_parity.py (fp16 activations, fp32 accumulation): rel-L2 gate 2.5e-3 (3x measured); the maps alone are fp32
(<= 1e-5)."""
import numpy as np
import pytest
import torch

from conftest import ds1_geom_oracle, load_case, noise_bank, rel_l2, write_binning_xml

pytestmark = pytest.mark.gpu

TOL = 2.5e-3   # 3x the largest error measured for these checks (8.0e-4, profiles/r01_pytest_gpu_geometry.log)


@pytest.fixture(scope="module")
def ds1(tmp_path_factory):
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    c = load_case("ds1")
    xml = write_binning_xml(tmp_path_factory.mktemp("ds1") / "binning.xml", c["meta"])
    cfg = dict(c["cfg"], SAMPLER="DDim", BIN_FILE=xml)
    m = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=4)
    sd = {k: v for k, v in c["sd"].items()}
    for i in range(5):
        sd[f"NN_embed.encs.{i}.weight"] = torch.from_numpy(c["gold"][f"enc_w_{i}"])
        sd[f"NN_embed.decs.{i}.weight"] = torch.from_numpy(c["gold"][f"dec_w_{i}"])
    assert sorted(sd) == sorted(m.state_dict())      # reference state-dict names incl. NN_embed.*
    m.load_state_dict(sd)
    return m, c, cfg


def test_nnconverter_enc_dec_device(ds1):
    m, c, _ = ds1
    assert m.do_embed and m._voxels == 368
    y = m.NN_embed.enc(c["x"].cuda())
    assert tuple(y.shape) == (3, 1, 5, 10, 30)
    e_enc = rel_l2(y.cpu().numpy(), c["gold"]["enc_out"])
    gy = torch.randn(c["gold"]["enc_out"].shape, generator=torch.Generator().manual_seed(7))
    e_dec = rel_l2(m.NN_embed.dec(gy.cuda()).cpu().numpy(), c["gold"]["dec_out"])
    print(f"NNConverter enc {e_enc:.2e} dec {e_dec:.2e}")
    assert max(e_enc, e_dec) < 1e-5
    with pytest.raises(RuntimeError):
        m.NN_embed.enc(c["x"])     # no CPU path


def test_ds1_denoise_matches_reference_golden(ds1):
    m, c, _ = ds1
    errs = []
    for tag in ("a", "b"):
        sig = torch.from_numpy(c["gold"][f"denoise_{tag}_sigma"])
        y = m.denoise((c["x"] * (1 + sig.view(-1, 1) ** 2).sqrt()).cuda(), E=c["E"].cuda(), sigma=sig.cuda().view(-1, 1),
                      layers=c["layers"].cuda())
        assert tuple(y.shape) == (3, 368)
        errs.append(rel_l2(y.cpu().numpy(), c["gold"][f"denoise_{tag}"]))
    print("ds1 denoise rel-L2 vs reference:", [f"{e:.2e}" for e in errs])
    assert max(errs) < TOL


def test_ds1_forward_matches_oracle(ds1):
    from oracle.unet_oracle import DenoiserOracle
    m, c, _ = ds1
    o = DenoiserOracle(c["cfg"], c["sd"], geom=ds1_geom_oracle(c))
    t = torch.tensor([0.2, -0.7, 1.1])
    with torch.no_grad():
        ref = o.unet(c["x"], c["E"], t, c["layers"]).numpy()
    y = m.forward(c["x"].cuda(), c["E"].cuda(), t.cuda(), c["layers"].cuda()).cpu().numpy()
    e = rel_l2(y, ref)
    print("ds1 forward rel-L2:", f"{e:.2e}")
    assert e < TOL


@pytest.mark.parametrize("graph", [False, True])
def test_ds1_ddim_sampling(ds1, graph):
    from calodiffusion_b200 import sample as S
    m, c, cfg = ds1
    m.use_graph = graph
    s = S.DDim(cfg)
    s.noise_bank = torch.stack(noise_bank(c["shape"], c["B"], 3)).cuda()
    x, xs, x0s = s(m, c["x"].cuda(), c["E"].cuda(), c["layers"].cuda(), 3, 0, True)
    e0 = rel_l2(torch.stack(x0s).cpu().numpy(), c["gold"]["DDim_x0s"])
    e1 = rel_l2(torch.stack(xs).cpu().numpy(), c["gold"]["DDim_xs"])
    e2 = rel_l2(x.cpu().numpy(), c["gold"]["DDim_x"])
    print("ds1 DDim", "graph" if graph else "eager", f"x0s {e0:.2e} xs {e1:.2e} final {e2:.2e}")
    assert max(e0, e1, e2) < TOL


def test_ds1_config1_sample_api(ds1):
    """BASELINE config 1 through the public API: model.sample(E, layers, 50 DDim steps) on a ragged last batch, Philox noise;
    finite output of the caller-facing shape, and a second engine fed the same noise bank agrees with the oracle."""
    from oracle.unet_oracle import DenoiserOracle, sample_ddim
    m, c, cfg = ds1
    m.use_graph = True
    g = torch.Generator().manual_seed(21)
    E, layers = torch.rand(4, 1, generator=g), torch.randn(4, 6, generator=g)
    out = m.sample(E.cuda(), layers.cuda(), num_steps=50)
    assert out.shape == (4, 368) and np.isfinite(out).all()
    # 6 steps against the oracle with a shared noise bank (B = 3: ragged vs max_batch = 4)
    from calodiffusion_b200 import sample as S
    s = S.DDim(cfg)
    bank = noise_bank(c["shape"], 3, 6, seed=9)
    s.noise_bank = torch.stack(bank).cuda()
    x, _, _ = s(m, c["x"].cuda(), E[:3].cuda(), layers[:3].cuda(), 6, 0, False)
    o = DenoiserOracle(c["cfg"], c["sd"], geom=ds1_geom_oracle(c))
    with torch.no_grad():
        xo, _, _ = sample_ddim(o, c["x"], E[:3], layers[:3], 6, 0.0, bank)
    e = rel_l2(x.cpu().numpy(), xo.numpy())
    print("ds1 DDim 6 steps vs oracle", f"{e:.2e}")
    assert e < TOL


def test_hgcal_matrix_converter_in_loop():
    """HGCal fixture shape with the geometry embedding inside the loop: x is (B,1,28,1988); MatrixConverter maps (seeded
    sparse stand-ins at the probe's density) run in every evaluation.  Against the oracle with the same matrices."""
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    from oracle.geom_embed import GeomEmbedOracle, hgcal_matrices
    from oracle.unet_oracle import DenoiserOracle
    c = load_case("hgcal")
    enc_mat, dec_mat = hgcal_matrices()
    cfg = dict(c["cfg"], SAMPLER="DDim", SHOWER_EMBED="NN", SHAPE_PAD=[-1, 1, 28, 1988],
               GEOM_EMBED=dict(enc_mat=enc_mat, dec_mat=dec_mat))
    m = CaloDiffusion(cfg, n_steps=cfg["NSTEPS"], loss_type=cfg["LOSS_TYPE"], max_batch=2)
    m.load_state_dict({k: v for k, v in c["sd"].items()}, strict=False)
    assert m.do_embed and m._voxels == 28 * 1988
    # converter alone vs the live reference's Embeder / Decoder goldens
    gold = load_case("ds1")["gold"]
    xh = torch.randn(2, 1, 28, 1988, generator=torch.Generator().manual_seed(11))
    yh = torch.randn(2, 1, 28, 12, 21, generator=torch.Generator().manual_seed(12))
    e_enc = rel_l2(m.NN_embed.enc(xh.cuda()).cpu().numpy(), gold["hgcal_enc_out"])
    e_dec = rel_l2(m.NN_embed.dec(yh.cuda()).cpu().numpy(), gold["hgcal_dec_out"])
    print(f"MatrixConverter enc {e_enc:.2e} dec {e_dec:.2e}")
    assert max(e_enc, e_dec) < 1e-5
    # denoise through the in-loop maps vs the oracle
    ocfg = {k: v for k, v in cfg.items() if k != "GEOM_EMBED"}
    o = DenoiserOracle(ocfg, c["sd"], geom=GeomEmbedOracle.matrices(enc_mat, dec_mat, 12, 21))
    sig = torch.tensor([1.5, 0.3])
    x = xh * (1 + sig.view(-1, 1, 1, 1) ** 2).sqrt()
    with torch.no_grad():
        ref = o.denoise(x, c["E"], sig, c["layers"]).numpy()
    y = m.denoise(x.cuda(), E=c["E"].cuda(), sigma=sig.cuda().view(-1, 1, 1, 1), layers=c["layers"].cuda())
    assert tuple(y.shape) == (2, 1, 28, 1988)
    e = rel_l2(y.cpu().numpy(), ref)
    print("hgcal in-loop denoise rel-L2:", f"{e:.2e}")
    assert e < TOL


def test_geom_map_errors():
    import ctypes as C
    from calodiffusion_b200 import _lib
    from calodiffusion_b200.calodiffusion import CaloDiffusion
    c = load_case("ds2")
    m = CaloDiffusion(dict(c["cfg"], SAMPLER="DDPM"), n_steps=400, loss_type="l2", max_batch=2)
    lib = m._lib
    h = lib.cd_unet_create(C.byref(m._desc(2)))
    rp = np.array([0, 1, 2], np.int32)
    col = np.array([0, 5], np.int32)
    val = np.ones(2, np.float32)
    assert lib.cd_unet_set_geom_map(h, 0, 2, 3, rp.ctypes.data, col.ctypes.data, val.ctypes.data) != 0   # column 5 >= 3
    assert b"column" in lib.cd_last_error()
    assert lib.cd_unet_set_geom_map(h, 2, 2, 8, rp.ctypes.data, col.ctypes.data, val.ctypes.data) != 0   # bad `which`
    assert lib.cd_unet_set_geom_map(h, 0, 2, 8, rp.ctypes.data, col.ctypes.data, val.ctypes.data) == 0
    for name, p in m.model.named_parameters():
        t = p.detach().cpu().contiguous()
        shp = (C.c_int64 * t.dim())(*t.shape)
        assert lib.cd_unet_set_weight(h, ("model." + name).encode(), t.data_ptr(), shp, t.dim()) == 0
    assert lib.cd_unet_finalize(h) != 0       # enc without dec
    assert b"both" in lib.cd_last_error()
    lib.cd_unet_destroy(h)
