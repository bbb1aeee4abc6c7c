"""CPU: geometry-embedding oracle vs the live reference's golden outputs (BASELINE config 1, Dataset 1 photons), and the
product's host-side converter logic (binning parser, GeomConverter matrices, CSR export) vs that oracle."""
import numpy as np
import pytest
import scipy.sparse as sp
import torch

from conftest import ds1_geom_oracle, load_case, noise_bank, rel_l2, write_binning_xml
from oracle import geom_embed
from oracle.unet_oracle import DenoiserOracle, sample_ddim

TOL = 2e-5


def _csr(m):
    return sp.csr_matrix((m["val"], m["col"], m["rowptr"]), shape=(m["n_rows"], m["n_cols"]))


def test_oracle_weight_mats_match_reference():
    c = load_case("ds1")
    mats, edges = geom_embed.geom_weight_mats(c["meta"]["lay_r_edges"])
    assert np.array_equal(edges.numpy(), c["gold"]["all_r_edges"])
    for i, w in enumerate(mats):
        assert np.array_equal(w.numpy(), c["gold"][f"gc_weight_{i}"])


def test_oracle_enc_dec_match_reference():
    c = load_case("ds1")
    g = ds1_geom_oracle(c)
    assert rel_l2(g.enc(c["x"]).numpy(), c["gold"]["enc_out"]) < 1e-6
    gy = torch.randn(c["gold"]["enc_out"].shape, generator=torch.Generator().manual_seed(7))
    assert rel_l2(g.dec(gy).numpy(), c["gold"]["dec_out"]) < 1e-6


def test_oracle_ds1_denoise_and_ddim_match_reference():
    c = load_case("ds1")
    m = DenoiserOracle(c["cfg"], c["sd"], geom=ds1_geom_oracle(c))
    with torch.no_grad():
        for tag in ("a", "b"):
            sig = torch.from_numpy(c["gold"][f"denoise_{tag}_sigma"])
            y = m.denoise(c["x"] * (1 + sig.view(-1, 1) ** 2).sqrt(), c["E"], sig, c["layers"]).numpy()
            assert rel_l2(y, c["gold"][f"denoise_{tag}"]) < TOL, tag
        x, xs, x0s = sample_ddim(m, c["x"], c["E"], c["layers"], 3, 0.0, noise_bank(c["shape"], c["B"], 3))
    assert rel_l2(torch.stack(x0s).numpy(), c["gold"]["DDim_x0s"]) < TOL
    assert rel_l2(torch.stack(xs).numpy(), c["gold"]["DDim_xs"]) < TOL
    assert rel_l2(x.numpy(), c["gold"]["DDim_x"]) < TOL


def test_oracle_hgcal_matrices_match_reference():
    c = load_case("ds1")
    enc_mat, dec_mat = geom_embed.hgcal_matrices()
    xh = torch.randn(2, 1, 28, 1988, generator=torch.Generator().manual_seed(11))
    yh = torch.randn(2, 1, 28, 12, 21, generator=torch.Generator().manual_seed(12))
    assert rel_l2(geom_embed.matrix_enc(xh, enc_mat, 12, 21).numpy(), c["gold"]["hgcal_enc_out"]) < 1e-6
    assert rel_l2(geom_embed.matrix_dec(yh, dec_mat).numpy(), c["gold"]["hgcal_dec_out"]) < 1e-6


def test_host_binning_and_geomconverter(tmp_path):
    from calodiffusion_b200 import geometry
    c = load_case("ds1")
    bins = geometry.BinningXML("photon", write_binning_xml(tmp_path / "binning.xml", c["meta"]))
    gc = geometry.GeomConverter(bins)
    assert list(gc.layer_boundaries) == list(c["gold"]["layer_boundaries"])
    assert list(gc.lay_alphas) == list(c["gold"]["lay_alphas"])
    assert gc.grid_shape() == (5, 10, 30) and gc.flat_size() == 368
    for i, w in enumerate(gc.weight_mats):
        assert np.array_equal(w.numpy(), c["gold"][f"gc_weight_{i}"])
    try:
        geometry.BinningXML("pion", str(tmp_path / "binning.xml"))
        assert False, "unknown particle must raise"
    except ValueError:
        pass


def test_host_nnconverter_csr_matches_oracle(tmp_path):
    from calodiffusion_b200 import geometry
    c = load_case("ds1")
    bins = geometry.BinningXML("photon", write_binning_xml(tmp_path / "binning.xml", c["meta"]))
    nn = geometry.NNConverter(bins=bins)
    assert [k for k, _ in nn.named_parameters()] == [f"encs.{i}.weight" for i in range(5)] + [f"decs.{i}.weight" for i in range(5)]
    # the reference's initialisation: area map + 1e-5 noise, pseudo-inverse decoder
    for i in range(5):
        assert np.abs(nn.encs[i].weight.detach().numpy() - c["gold"][f"gc_weight_{i}"]).max() < 1e-4
    nn.load_state_dict({k: torch.from_numpy(c["gold"][k.replace("encs.", "enc_w_").replace("decs.", "dec_w_").replace(".weight", "")])
                        for k in nn.state_dict()})
    enc, dec = nn.csr_maps()
    g = ds1_geom_oracle(c)
    x = c["x"].numpy()
    assert rel_l2((_csr(enc) @ x.T).T.reshape(c["gold"]["enc_out"].shape), c["gold"]["enc_out"]) < 1e-6
    gy = torch.randn(c["gold"]["enc_out"].shape, generator=torch.Generator().manual_seed(7))
    assert rel_l2((_csr(dec) @ gy.reshape(3, -1).numpy().T).T, g.dec(gy).numpy()) < 1e-6
    assert enc["rowptr"].dtype == np.int32 and enc["col"].dtype == np.int32 and enc["val"].dtype == np.float32


def test_host_matrixconverter_csr_matches_oracle():
    from calodiffusion_b200 import geometry
    enc_mat, dec_mat = geom_embed.hgcal_matrices()
    mc = geometry.MatrixConverter([-1, 1, 28, 12, 21], enc_mat, dec_mat)
    enc, dec = mc.csr_maps()
    assert enc["n_rows"] == 28 * 252 and enc["n_cols"] == 28 * 1988 and len(enc["val"]) == int((enc_mat != 0).sum())
    xh = torch.randn(2, 1, 28, 1988, generator=torch.Generator().manual_seed(11))
    yh = torch.randn(2, 1, 28, 12, 21, generator=torch.Generator().manual_seed(12))
    assert rel_l2((_csr(enc) @ xh.reshape(2, -1).numpy().T).T, geom_embed.matrix_enc(xh, enc_mat, 12, 21).reshape(2, -1).numpy()) < 1e-6
    assert rel_l2((_csr(dec) @ yh.reshape(2, -1).numpy().T).T, geom_embed.matrix_dec(yh, dec_mat).reshape(2, -1).numpy()) < 1e-6
    try:
        geometry.MatrixConverter([-1, 1, 28, 12, 20], enc_mat, dec_mat)
        assert False
    except ValueError:
        pass


def test_hgcal_converter_init_matches_reference_fingerprints():
    """HGCalConverter.init (GLaM maps from geom_william.pkl; HGCal_utils.py:412-486, 590-634) against fingerprints of the live
    reference's matrices (tests/golden/hgcal_geom.npz).  Host-side construction only: no compute call on the device."""
    import os
    from conftest import GOLDEN, ROOT
    pk = os.path.join(ROOT, "oracle", "_ref", "HGCalShowers", "geoms", "geom_william.pkl")
    if not os.path.exists(pk):
        pk = "/root/reference/HGCalShowers/geoms/geom_william.pkl"
    if not os.path.exists(pk):
        pytest.skip("HGCal geometry pickle not staged")
    from calodiffusion_b200.geometry import HGCalConverter
    gold = np.load(os.path.join(GOLDEN, "hgcal_geom.npz"))
    c = HGCalConverter(bins=[-1, 1, 28, 12, 21], geom_file=pk)
    c.init(norm=False, dataset_num=111)
    e, d = c.enc_mat, c.dec_mat
    assert np.array_equal((e != 0).sum((1, 2)).numpy(), gold["enc_nnz"])
    assert np.allclose(e.double().sum((1, 2)).numpy(), gold["enc_sum"], rtol=0, atol=1e-9)
    assert np.array_equal(c.enc_mask.sum((1, 2)).numpy(), gold["enc_mask_nnz"])
    g = torch.Generator().manual_seed(21)
    ve, vd = torch.randn(e.shape[2], generator=g), torch.randn(d.shape[2], generator=g)
    assert np.allclose((e @ ve).numpy(), gold["enc_proj"], rtol=1e-6, atol=1e-6)
    assert np.allclose((d @ vd).numpy(), gold["dec_proj"], rtol=1e-4, atol=1e-5)
    enc_csr, dec_csr = c.csr_maps()
    assert len(enc_csr["val"]) == int(gold["enc_nnz"].sum())
    assert abs(len(dec_csr["val"]) - int(gold["dec_nnz_eps"].sum())) <= 4      # entries within round-off of the 1e-6 threshold
    from calodiffusion_b200.postprocess import DATASET_PARAMS
    assert DATASET_PARAMS[101]["embed_mean"] == 0.0835 and DATASET_PARAMS[101]["embed_std"] == 3.1083   # consts.py:171-172
