"""Shower features for the 'histograms indistinguishable' check (restated from the numpy closures in
utils/plots.py and utils/HighLevelFeatures.py of the reference) and the two-sample distances used to compare them."""
import numpy as np


def shower_features(data, e_inc, shape_zpr):
    """data (N, Z*P*R) deposited energies, e_inc (N,1) -> dict of per-shower feature arrays."""
    Z, P, R = shape_zpr
    n = data.shape[0]
    d = np.asarray(data, dtype=np.float64).reshape(n, Z, P, R)
    tot = d.sum((1, 2, 3))
    layer = d.sum((2, 3))                                                     # utils/plots.py:568-576
    safe = np.maximum(tot, 1e-30)
    f = {}
    f["total_E"] = tot                                                        # :885-898
    f["Edep_over_Einc"] = tot / np.asarray(e_inc, dtype=np.float64).reshape(-1)   # :406-425
    # :913-916 counts voxels above 1e-3 GeV; with random weights the scale is arbitrary, so the threshold is relative to the
    # shower's own largest voxel (a per-shower quantity: comparable between independently sampled sets)
    f["n_hits"] = (d > 1e-3 * d.reshape(n, -1).max(1).reshape(n, 1, 1, 1)).sum((1, 2, 3)).astype(np.float64)
    f["depth_mean"] = (layer * np.arange(Z)).sum(1) / safe                    # energy-weighted layer index
    f["depth_width"] = np.sqrt(np.maximum((layer * np.arange(Z) ** 2).sum(1) / safe - f["depth_mean"] ** 2, 0))
    rprof = d.sum((1, 2))                                                     # :619-625
    f["r_centre"] = (rprof * np.arange(R)).sum(1) / safe                      # :463-519 (GetCenter, bin-index units)
    f["r_width"] = np.sqrt(np.maximum((rprof * np.arange(R) ** 2).sum(1) / safe - f["r_centre"] ** 2, 0))
    pprof = d.sum((1, 3))                                                     # :652-657
    ang = 2 * np.pi * np.arange(P) / P
    cx, cy = (pprof * np.cos(ang)).sum(1) / safe, (pprof * np.sin(ang)).sum(1) / safe
    f["phi_spread"] = np.sqrt(np.maximum(-2.0 * np.log(np.maximum(np.hypot(cx, cy), 1e-12)), 0))   # :17-43 circular spread
    f["max_voxel_frac"] = d.reshape(n, -1).max(1) / safe                      # :979-984
    f["layer_sparsity"] = (d > 1e-6 * tot.reshape(-1, 1, 1, 1)).mean((1, 2, 3))   # :685-692
    for z in (0, Z // 2, Z - 1):
        f[f"E_layer_{z}"] = layer[:, z]
    return f


def separation_power(a, b, bins=20):
    """0.5 * sum (h1-h2)^2/(h1+h2) on unit-normalised histograms over the joint range (utils/plots.py:374-382)."""
    lo, hi = min(a.min(), b.min()), max(a.max(), b.max())
    if hi <= lo:
        return 0.0
    h1, _ = np.histogram(a, bins=bins, range=(lo, hi))
    h2, _ = np.histogram(b, bins=bins, range=(lo, hi))
    h1, h2 = h1 / h1.sum(), h2 / h2.sum()
    den = h1 + h2
    m = den > 0
    return float(0.5 * ((h1[m] - h2[m]) ** 2 / den[m]).sum())


def ks_pvalue(a, b):
    from scipy.stats import ks_2samp
    return float(ks_2samp(a, b).pvalue)
