"""The ReverseNormCaloChall oracle against golden outputs of the live reference (utils/utils.py:446-573)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle.reverse_norm import golden_inputs, reverse_norm

CASES = [(2, "layer-logit-norm", True), (2, "logit-norm", False), (3, "logit-norm", False)]


@pytest.mark.parametrize("ds,smap,use_layers", CASES)
@pytest.mark.parametrize("ecut", [0.0, 1.515e-5])
def test_reverse_norm_matches_reference(ds, smap, use_layers, ecut):
    gold = np.load(os.path.join(GOLDEN, "reverse_norm.npz"))
    vox, e, layerE = golden_inputs(ds)
    data, energy = reverse_norm(vox, e, ds, smap, layerE=layerE if use_layers else None, ecut=ecut)
    key = f"ds{ds}_{smap}_ecut{int(ecut > 0)}"
    np.testing.assert_allclose(energy, gold[key + "_energy"], rtol=1e-6)
    ref = gold[key + "_data"]
    # voxels within float rounding of the threshold may flip; everything else must agree to fp32 accuracy
    near_cut = np.abs(np.abs(ref) - ecut) < 1e-3 * ecut if ecut > 0 else np.zeros_like(ref, dtype=bool)
    np.testing.assert_allclose(data[~near_cut], ref[~near_cut], rtol=2e-5, atol=1e-12)
