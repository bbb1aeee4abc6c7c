"""Host utilities mirroring the bits of calodiffusion/utils/utils.py the hot path touches."""
import importlib

import yaml

# utils/utils.py:44-133 — R bin edges per DATASET_NUM
_R_EDGES = {
    0: [0.0, 1.0, 4.0, 5.0, 7.0, 10.0, 15.0, 20.0, 30.0, 50.0, 80.0, 90.0, 100.0, 130.0, 150.0, 160.0, 200.0, 250.0,
        300.0, 350.0, 400.0, 600.0, 1000.0, 2000.0],
    1: [0.0, 2.0, 4.0, 5.0, 6.0, 8.0, 10.0, 12.0, 15.0, 20.0, 25.0, 30.0, 40.0, 50.0, 60.0, 70.0, 80.0, 90.0, 100.0,
        120.0, 130.0, 150.0, 160.0, 200.0, 250.0, 300.0, 350.0, 400.0, 600.0, 1000.0, 2000.0],
    2: [0, 4.65, 9.3, 13.95, 18.6, 23.25, 27.9, 32.55, 37.2, 41.85],
    3: [0, 2.325, 4.65, 6.975, 9.3, 11.625, 13.95, 16.275, 18.6, 20.925, 23.25, 25.575, 27.9, 30.225, 32.55, 34.875,
        37.2, 39.525, 41.85],
}


def LoadJson(file_name):
    """utils/utils.py:439-443: the shipped 'json' configs are YAML flow mappings (single quotes, trailing commas)."""
    with open(file_name) as f:
        return yaml.safe_load(f)


def scaled_r_centres(dataset_num, n_r):
    """Scaled R image values (utils/utils.py:134-150): bin centre / last bin centre.  ValueError on mismatch."""
    if dataset_num >= 100:
        edges = [float(i) for i in range(n_r + 1)]
    elif dataset_num in _R_EDGES:
        edges = _R_EDGES[dataset_num]
    else:
        raise ValueError(f"RZ binning missing for dataset num {dataset_num}")
    centres = [(edges[i] + edges[i + 1]) / 2.0 for i in range(len(edges) - 1)]
    if len(centres) != n_r:
        raise ValueError(f"Mismatch for dataset num {dataset_num} - expecting dataset with final dim {len(centres)}")
    import numpy as np
    c = np.asarray(centres, dtype=np.float32)
    return (c / c[-1]).astype(np.float32)


def load_attr(type_, algo_name):
    """utils/utils.py:1047-1061: the registry is the module namespace (ValueError on unknown names)."""
    modname = {"sampler": "calodiffusion_b200.sample", "loss": "calodiffusion_b200.loss"}.get(type_)
    if modname is None:
        raise ValueError(f"Unknown attribute type {type_}")
    module = importlib.import_module(modname)
    if not hasattr(module, algo_name):
        raise ValueError(f"{type_} '{algo_name}' is not supported by the B200 backend ({modname})")
    return getattr(module, algo_name)
