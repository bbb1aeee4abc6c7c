"""Golden vectors for the ReverseNormCaloChall oracle, from the LIVE reference (build container only).
    python oracle/make_golden_reverse_norm.py      # writes tests/golden/reverse_norm.npz"""
import contextlib
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref_shim  # noqa: E402

ref_shim.install()
from calodiffusion.utils import utils as ru  # noqa: E402


from oracle.reverse_norm import golden_inputs as inputs  # noqa: E402


def main():
    out = {}
    for ds, smap, use_layers in ((2, "layer-logit-norm", True), (2, "logit-norm", False), (3, "logit-norm", False)):
        vox, e, layerE = inputs(ds)
        for ecut in (0.0, 1.515e-5):
            with contextlib.redirect_stdout(io.StringIO()):
                data, energy = ru.ReverseNormCaloChall(vox.copy(), e.copy(), shape=None, config=None, emax=1000.0, emin=1.0,
                                                       max_deposit=2, logE=True, layerE=layerE.copy() if use_layers else None,
                                                       showerMap=smap, dataset_num=ds, ecut=ecut)
            key = f"ds{ds}_{smap}_ecut{int(ecut > 0)}"
            out[key + "_data"] = np.asarray(data, dtype=np.float32)
            out[key + "_energy"] = np.asarray(energy, dtype=np.float32)
    np.savez_compressed(os.path.join(os.path.dirname(HERE), "tests", "golden", "reverse_norm.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
