"""CPU oracle (TEST INFRASTRUCTURE ONLY) for `ReverseNormCaloChall` (utils/utils.py:446-573), the post-processing
that turns sampled voxels back into energies — SURVEY 8(f) row 1.  Plain numpy restatement for the shipped
CaloChallenge maps ('logit-norm', 'layer-logit-norm'; dataset_num 2 / 3; logE; no quantile transform).
Pinned to the live reference by tests/golden/reverse_norm.npz (oracle/make_golden_reverse_norm.py)."""
import numpy as np

# utils/consts.py:82-117 (dataset2_params / dataset3_params)
DATASET_PARAMS = {
    2: dict(logit_mean=-12.8564, logit_std=1.9123, totalE_mean=0.3926, totalE_std=0.05546, layers_mean=-6.35551,
            layers_std=3.90699),
    3: dict(logit_mean=-13.4753, logit_std=1.1070, totalE_mean=0.0, totalE_std=1.0, layers_mean=0.0, layers_std=1.0),
}


def reverse_logit(x, alpha=1e-6):
    """utils/utils.py:233-237."""
    e = np.exp(x)
    o = e / (1 + e)
    return (o - alpha) / (1 - 2 * alpha)


def reverse_norm(voxels, e, dataset_num, shower_map, layerE=None, emax=1000.0, emin=1.0, max_deposit=2.0, ecut=0.0):
    """voxels (B,1,Z,P,R) network-space showers, e (B,1) scaled incident energy -> (data (B, Z*P*R) [GeV units of the
    dataset], energy (B,1)).  utils/utils.py:446-573 for dataset_num > 1, logE=True, orig_shape=False."""
    c = DATASET_PARAMS[dataset_num]
    voxels = np.asarray(voxels)
    energy = emin * (emax / emin) ** e                                        # :477-478
    if "logit" not in shower_map or "norm" not in shower_map:
        raise ValueError("oracle covers the '*logit-norm' shower maps only")
    v = voxels * c["logit_std"] + c["logit_mean"]                             # :492-493
    data = reverse_logit(v)                                                   # :502
    if "layer" in shower_map:                                                 # :519-546
        assert layerE is not None
        totalE, layers = layerE[:, :1], layerE[:, 1:]
        totalE = totalE * c["totalE_std"] + c["totalE_mean"]
        layers = layers * c["layers_std"] + c["layers_mean"]
        layers = reverse_logit(layers)
        layers = layers / np.sum(layers, axis=1, keepdims=True)
        layers = layers * totalE
        data = np.squeeze(data).copy()
        if data.ndim == 3:
            data = data[None]
        eps = 1e-6
        data[data < 0] = 0
        prev_layers = np.sum(data, (2, 3), keepdims=True)
        layers = layers.reshape((-1, data.shape[1], 1, 1))
        rescale = layers / (prev_layers + 1e-10)
        rescale[layers < eps] = 1.0
        rescale[prev_layers < eps] = 1.0
        data = data * rescale
    data = data.reshape(voxels.shape[0], -1) * max_deposit * energy.reshape(-1, 1)   # :562-563
    if ecut > 0:
        data = data.copy()
        data[data < ecut] = 0                                                 # :573
    return data, energy


def golden_inputs(dataset_num, B=3, seed=21):
    """Seeded inputs of tests/golden/reverse_norm.npz (shared by the generator and the tests)."""
    shape = (B, 1, 45, 16, 9) if dataset_num == 2 else (B, 1, 45, 50, 18)
    rng = np.random.default_rng(seed + dataset_num)
    vox = rng.standard_normal(shape).astype(np.float32) * 1.5
    vox[0, 0, 3] = -40.0            # a layer that reverse-logits to (slightly negative) zero: exercises the eps guards
    e = rng.random((B, 1)).astype(np.float32)
    layerE = rng.standard_normal((B, 46)).astype(np.float32)
    layerE[1, 5] = -30.0            # a conditioning layer energy below eps
    return vox, e, layerE
