"""Golden vectors for BASELINE config 4 with the REAL HGCal geometry, from the LIVE reference (build container only).

    python oracle/make_golden_hgcal.py        # writes tests/golden/hgcal_geom.npz

The unmodified reference builds `HGCalConverter` from HGCalShowers/geoms/geom_william.pkl (the pickle config 4 names;
utils/HGCal_utils.py:412-690) and its `CaloDiffusion` with the converter inside every evaluation (SHOWER_EMBED 'NN',
SHAPE_PAD (-1,1,28,1988)).  Stored (inputs are regenerated from seeds by the tests):
  * fingerprints of enc_mat / dec_mat: per-layer nnz, sums, and their products with seeded vectors;
  * enc / dec outputs (averaging decode), and the sampled ("sparse") decode with torch.rand_like replaying a seeded tensor
    (per-shower and per-batch draws);
  * ReverseNormHGCal on cell-space showers (embed=False) and on grid-space showers through the converter (embed=True,
    converter normalisation of dataset 101), 'layer-logit-norm';
  * ReverseNormCaloChall for Dataset 1 in original shape (dataset 11: ragged layer boundaries of the photon binning file);
  * `denoise` at two noise levels and a 2-step DDim run of the in-loop model.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_shim  # noqa: E402
from oracle.make_golden import HGCAL_OVERRIDES, ReplayNoise, make_inputs, noise_bank, run_sampler  # noqa: E402
from oracle.unet_oracle import synth_state_dict  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
BINS = [-1, 1, 28, 12, 21]


class ReplayRand:
    """torch.rand_like -> a seeded uniform tensor of the requested shape (generate_sparse_mat's only draw)."""

    def __init__(self, seed):
        self.seed = seed

    def __enter__(self):
        self._orig = torch.rand_like
        torch.rand_like = lambda t, *a, **k: torch.rand(t.shape, generator=torch.Generator().manual_seed(self.seed))
        return self

    def __exit__(self, *exc):
        torch.rand_like = self._orig


def main():
    torch.set_num_threads(os.cpu_count())
    ref_shim.install()
    import calodiffusion.utils.HGCal_utils as H
    from calodiffusion.utils import utils as U
    pk = f"{ref_shim.REF_ROOT}/HGCalShowers/geoms/geom_william.pkl"
    conv = H.HGCalConverter(bins=BINS, geom_file=pk, device=torch.device("cpu"))
    conv.init(norm=False, dataset_num=111)
    out = {}
    e, d = conv.enc_mat, conv.dec_mat
    g = torch.Generator().manual_seed(21)
    ve, vd = torch.randn(e.shape[2], generator=g), torch.randn(d.shape[2], generator=g)
    out["enc_nnz"] = (e != 0).sum((1, 2)).numpy()
    out["enc_sum"] = e.double().sum((1, 2)).numpy()
    out["enc_proj"] = (e @ ve).numpy()
    out["dec_nnz_eps"] = (d.abs() > 1e-6).sum((1, 2)).numpy()
    out["dec_proj"] = (d @ vd).numpy()
    out["enc_mask_nnz"] = conv.enc_mask.sum((1, 2)).numpy()
    # the decoder's positive entries (generate_sparse_mat's mask): the pseudo-inverse is not bit-reproducible across CPUs and
    # the sampled decoding is discrete, so the exact-parity test installs these instead of the locally computed ones
    idx = np.flatnonzero(d.numpy() > 1e-6).astype(np.int32)
    out["dec_pos_idx"], out["dec_pos_val"] = idx, d.numpy().reshape(-1)[idx]
    xh = torch.rand(2, 1, 28, 1988, generator=torch.Generator().manual_seed(11))
    yh = torch.rand(2, 1, 28, 12, 21, generator=torch.Generator().manual_seed(12))
    with torch.no_grad():
        out["enc_out"] = conv.enc(xh).numpy()
        out["dec_out"] = conv.dec(yh).numpy()
        with ReplayRand(33):
            out["dec_sparse"] = conv.dec(yh, sparse_decoding=True, sparse_per_batch=False).numpy()
        with ReplayRand(34):
            out["dec_sparse_per_batch"] = conv.dec(yh, sparse_decoding=True, sparse_per_batch=True).numpy()
    # ReverseNormHGCal
    gen = torch.Generator().manual_seed(41)
    vox_cells = torch.randn(2, 1, 28, 1988, generator=gen).numpy()
    vox_grid = torch.randn(2, 1, 28, 12, 21, generator=gen).numpy()
    einfo = torch.rand(2, 3, generator=gen).numpy()
    layerE = torch.randn(2, 29, generator=gen).numpy()
    kw = dict(emax=1000.0, emin=1.0, max_deposit=2.0, layerE=layerE, showerMap="layer-logit-norm")
    data, gen_out = H.ReverseNormHGCal(vox_cells.copy(), einfo, dataset_num=111, embed=False, **kw)
    out["rn_cells_data"], out["rn_gen_out"] = data.astype(np.float32), gen_out.astype(np.float32)
    conv_n = H.HGCalConverter(bins=BINS, geom_file=pk, device=torch.device("cpu"))
    conv_n.init(norm=True, dataset_num=101)
    data, _ = H.ReverseNormHGCal(vox_grid.copy(), einfo, dataset_num=101, embed=True, NN_embed=conv_n, **kw)
    out["rn_embed_data"] = data.astype(np.float32)
    # (without a 'layer' map the reference's final `data * max_deposit * energy.reshape(-1, 1, 1)` broadcasts (N,1,L,C) against
    #  (N,1,1) into (N,N,L,C) - a shape quirk nothing downstream relies on; not pinned)
    # ReverseNormCaloChall, Dataset 1 photons in original shape (dataset 11)
    xml = f"{ref_shim.REF_ROOT}/CaloChallenge/code/binning_dataset_1_photons.xml"
    v1 = torch.randn(3, 368, generator=gen).numpy()
    e1 = torch.rand(3, 1, generator=gen).numpy()
    l1 = torch.randn(3, 6, generator=gen).numpy()
    data, energy = U.ReverseNormCaloChall(v1.copy(), e1, emax=4194.304, emin=0.256, binning_file=xml, max_deposit=3.1, logE=True,
                                          layerE=l1, showerMap="layer-logit-norm", dataset_num=1, orig_shape=True, ecut=0.0000001)
    out["rn_ds1_data"], out["rn_ds1_energy"] = data, energy
    # in-loop model with the real converter
    over = dict(HGCAL_OVERRIDES, SHOWER_EMBED="NN", SHAPE_PAD=[-1, 1, 28, 1988], BIN_FILE=pk)
    model, cfg = ref_shim.load_reference_model("config_HGCal.json", "DDim", over)
    assert model.do_embed
    shapes = {k: list(v.shape) for k, v in model.state_dict().items()}
    model.load_state_dict(synth_state_dict(shapes, seed=0))
    E, layers, _ = make_inputs([1, 28, 12, 21], 2, 3, 29)
    x = torch.randn(2, 1, 28, 1988, generator=torch.Generator().manual_seed(2))
    with torch.no_grad():
        for tag, sig in (("a", [1.5, 0.3]), ("b", [40.0, 0.01])):
            sigma = torch.tensor(sig).view(2, 1, 1, 1)
            y = model.denoise(x * (1 + sigma ** 2).sqrt(), E=E, sigma=sigma, layers=layers)
            out[f"denoise_{tag}_sigma"] = np.array(sig, np.float32)
            out[f"denoise_{tag}"] = y.numpy()
        bank = noise_bank([1, 28, 1988], 2, 2)
        xf, xs, x0s = run_sampler(model, "DDim", x, E, layers, 2, bank)
        out["DDim_x"] = xf.numpy()
        out["DDim_x0s"] = np.stack([t.numpy() for t in x0s])
    np.savez_compressed(os.path.join(OUT, "hgcal_geom.npz"), **out)
    print({k: (v.shape, v.dtype) for k, v in out.items()})


if __name__ == "__main__":
    main()
