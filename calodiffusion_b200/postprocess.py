"""Device-side mirrors of the reference's post-processing of sampled showers (models/diffusion.py:171-191) - SURVEY 8(f) row 1:

  ReverseNormCaloChall   utils/utils.py:446-573       datasets 2 / 3 (network shape) and datasets 0 / 1 in their original
                                                       flat shape (SHOWER_EMBED 'orig-*': ragged layers from the binning file)
  ReverseNormHGCal       utils/HGCal_utils.py:167-292  HGCal, with the GLaM decoding (averaging or sampled) in the middle

Same names, argument meaning and return values `(data numpy, energy numpy)` as the reference; the arithmetic runs in the
CUDA library (`cd_reverse_norm_ex`, `cd_geom_apply` / `cd_geom_decode_sparse`) on the showers while they are still on
the device, instead of numpy on the host.  Variants the kernels do not cover raise ValueError: the quantile transform
(needs the reference's joblib files) and datasets 0 / 1 in embedded shape (GeomConverter.unconvert).
"""
import ctypes as C
import math

import numpy as np
import torch

from . import _lib

# utils/consts.py:1-209 (dataset_params); 10 / 11 = datasets 0 / 1 in original shape
DATASET_PARAMS = {
    0: dict(logit_mean=-12.4783, logit_std=2.21267, logit_min=-13.81551, logit_max=0.9448, log_mean=0.0, log_std=1.0, log_min=0.0,
            log_max=2.0, totalE_mean=0.0, totalE_std=1.0, layers_mean=0.0, layers_std=1.0),
    10: dict(logit_mean=-11.7610, logit_std=2.84317, logit_min=-13.81551, logit_max=0.2554, log_mean=0.0, log_std=1.0, log_min=0.0,
             log_max=2.0, totalE_mean=0.2359, totalE_std=0.08255, layers_mean=-4.9742, layers_std=4.89629),
    1: dict(logit_mean=-12.1444, logit_std=2.45056, logit_min=-13.81551, logit_max=-1.6615, log_mean=0.0, log_std=1.0, log_min=0.0,
            log_max=2.0, totalE_mean=0.0, totalE_std=1.0, layers_mean=0.0, layers_std=1.0),
    11: dict(logit_mean=-9.9807, logit_std=3.14168, logit_min=-13.81551, logit_max=0.2554, log_mean=0.0, log_std=1.0, log_min=0.0,
             log_max=2.0, totalE_mean=0.3123, totalE_std=0.02872, layers_mean=-4.9023, layers_std=5.17364),
    2: dict(logit_mean=-12.8564, logit_std=1.9123, logit_min=-13.8155, logit_max=0.1153, log_mean=-17.5451, log_std=4.4086,
            log_min=-20.0, log_max=-0.6372, totalE_mean=0.3926, totalE_std=0.05546, layers_mean=-6.35551, layers_std=3.90699),
    3: dict(logit_mean=-13.4753, logit_std=1.1070, logit_min=-13.81551, logit_max=0.2909, log_mean=-1.1245, log_std=3.3451,
            log_min=-18.6905, log_max=0.0, totalE_mean=0.0, totalE_std=1.0, layers_mean=0.0, layers_std=1.0),
    100: dict(logit_mean=-13.7371, logit_std=0.68639, logit_min=-13.81551, logit_max=0.2909, totalE_mean=0.0055, totalE_std=0.00018,
              layers_mean=-4.4450, layers_std=2.37667),
    101: dict(logit_mean=-18.3170, logit_std=1.03153, logit_min=-13.81551, logit_max=0.2909, embed_mean=0.0835, embed_std=3.1083,
              totalE_mean=0.5538, totalE_std=0.01767, layers_mean=-4.5836, layers_std=2.98382),
    111: dict(logit_mean=-17.3442, logit_std=3.26085, logit_min=-13.81551, logit_max=0.2909, embed_mean=0.0, embed_std=1.0,
              totalE_mean=1.1076, totalE_std=0.03535, layers_mean=-4.5836, layers_std=2.98382),
}
_PARAMS = DATASET_PARAMS   # (name kept for callers of the first version)


def _map_constants(c, showerMap, max_deposit):
    """(map_kind, a_scale, a_shift, log_cap) of the inverse voxel map (utils.py:492-517 / HGCal_utils.py:211-237)."""
    if "quantile" in showerMap:
        raise ValueError("the quantile transform (joblib file of the reference) is not supported by the B200 backend")
    kind = "logit" if "logit" in showerMap else ("log" if "log" in showerMap else None)
    if kind is None:
        raise ValueError(f"unknown shower map {showerMap!r}")
    if "norm" in showerMap:
        a, b = c[kind + "_std"], c[kind + "_mean"]
    elif "scaled" in showerMap:
        a = 0.5 * (c[kind + "_max"] - c[kind + "_min"])
        b = a + c[kind + "_min"]
    else:
        a, b = 1.0, 0.0
    return (0 if kind == "logit" else 1), float(a), float(b), float(math.log(max_deposit))


def _run(desc_kw, vox, e, layerE, seg=None, stream=None):
    """One `cd_reverse_norm_ex` call: vox (N, total) on the device -> (data (N, total), energy (N, 1))."""
    lib = _lib.load()
    n, total = vox.shape
    dev = vox.device
    out = torch.empty((n, total), device=dev, dtype=torch.float32)
    energy = torch.empty((n, 1), device=dev, dtype=torch.float32)
    d = _lib.ReverseNormExDesc()
    base = dict(map_kind=0, a_scale=1.0, a_shift=0.0, alpha=1e-6, log_cap=0.0, post_scale=1.0, post_shift=0.0, totalE_mean=0.0,
                totalE_std=1.0, layers_mean=0.0, layers_std=1.0, guard_eps=1e-6, energy_mode=0, emin=1.0, emax=1.0, max_deposit=1.0,
                ecut=0.0, e_stride=1, final_scale=1, total=total)
    base.update(desc_kw)
    for k, v in base.items():
        setattr(d, k, v)
    st = torch.cuda.current_stream(dev).cuda_stream if stream is None else stream
    for lo in range(0, n, 32768):     # grid.y limit
        hi = min(n, lo + 32768)
        _lib.check(lib.cd_reverse_norm_ex(C.byref(d), seg.data_ptr() if seg is not None else None, vox[lo:hi].data_ptr(),
                                          e[lo:hi].data_ptr(), layerE[lo:hi].data_ptr() if layerE is not None else None,
                                          out[lo:hi].data_ptr(), energy[lo:hi].data_ptr(), hi - lo, st), "reverse_norm")
    return out, energy


def _prep(voxels, e, layerE, n_layers, need_layers):
    if not voxels.is_cuda:
        raise RuntimeError("reverse_norm_device needs CUDA tensors; there is no CPU fallback")
    dev = voxels.device
    n = voxels.shape[0]
    v = voxels.to(torch.float32).reshape(n, -1).contiguous()
    ed = e.to(device=dev, dtype=torch.float32).reshape(n, -1).contiguous()
    le = None
    if need_layers:
        if layerE is None:
            raise AssertionError("layerE is required for 'layer' shower maps")   # utils.py:520
        le = layerE.to(device=dev, dtype=torch.float32).reshape(n, -1).contiguous()
        if le.shape[1] != n_layers + 1:
            raise ValueError(f"layerE must be (N, {n_layers + 1}) = [totalE, layers]")
    return v, ed, le


def reverse_norm_device(voxels, e, emax=9999.0, emin=0.0001, max_deposit=2, logE=True, layerE=None, showerMap="log",
                        dataset_num=2, ecut=0.0, stream=None, layer_boundaries=None):
    """ReverseNormCaloChall, tensors in, tensors out (all on the voxels' CUDA device): returns (data (N, voxels), energy (N,1)).
    voxels (N, 1, Z, P, R) for datasets 2 / 3; (N, n_flat) with `layer_boundaries` (np.unique(bins.GetBinEdges())) for the
    original-shape datasets 10 / 11."""
    if dataset_num not in (2, 3, 10, 11):
        raise ValueError("device ReverseNormCaloChall supports dataset_num 2, 3 and the original-shape datasets 0 / 1 (10 / 11)")
    c = DATASET_PARAMS[dataset_num]
    kind, a, b, cap = _map_constants(c, showerMap, max_deposit)
    layer = "layer" in showerMap
    seg = None
    if dataset_num >= 10:
        if layer_boundaries is None:
            raise ValueError("original-shape datasets need the layer boundaries of the binning file")
        bounds = np.asarray(layer_boundaries, np.int32)
        nseg = len(bounds) - 1
        if voxels.dim() != 2 or int(bounds[-1]) != voxels.shape[1] or int(bounds[0]) != 0:
            raise ValueError("voxels must be (N, n_flat) matching the layer boundaries")
        seg = torch.from_numpy(bounds).to(voxels.device)
    else:
        if voxels.dim() != 5:
            raise ValueError("voxels must be (N, 1, Z, P, R)")
        nseg = int(voxels.shape[2])
    v, ed, le = _prep(voxels, e, layerE, nseg, layer)
    if ed.shape[1] != 1:
        raise ValueError("e must hold one value per shower")
    kw = dict(nseg=nseg, map_kind=kind, a_scale=a, a_shift=b, alpha=1e-6, log_cap=cap, guard_eps=1e-6,
              energy_mode=0 if logE else 1, emin=float(emin), emax=float(emax), max_deposit=float(max_deposit), ecut=float(ecut),
              totalE_mean=c["totalE_mean"], totalE_std=c["totalE_std"], layers_mean=c["layers_mean"], layers_std=c["layers_std"])
    return _run(kw, v, ed, le, seg, stream)


def ReverseNormCaloChall(voxels, e, emax=9999.0, emin=0.0001, config=None, shape=None, binning_file="", max_deposit=2,
                         logE=True, layerE=None, showerMap="log", dataset_num=2, orig_shape=False, ecut=0.0, **kwargs):
    """Reference signature (utils/utils.py:446-462).  `voxels`, `e`, `layerE` may be numpy arrays (as in the reference)
    or tensors already on the device; returns numpy `(data, energy)` like the reference."""
    if dataset_num > 3 or dataset_num < 0:
        raise ValueError("Invalid dataset %i!" % dataset_num)         # the reference prints this and exits (utils.py:466-468)
    bounds = None
    if dataset_num <= 1:
        if not orig_shape:
            raise ValueError("datasets 0 / 1 in embedded shape (GeomConverter.unconvert, utils.py:562-565) are not on the B200 path")
        from .geometry import BinningXML
        bins = BinningXML("photon" if dataset_num == 1 else "pion", binning_file)
        bounds = np.unique(bins.GetBinEdges())
        dataset_num += 10                                             # utils.py:475-476
    dev = torch.device("cuda", torch.cuda.current_device())
    as_t = lambda a: a.to(dev) if torch.is_tensor(a) else torch.as_tensor(np.asarray(a), device=dev)
    data, energy = reverse_norm_device(as_t(voxels), as_t(e), emax=emax, emin=emin, max_deposit=max_deposit, logE=logE,
                                       layerE=None if layerE is None else as_t(layerE), showerMap=showerMap,
                                       dataset_num=dataset_num, ecut=ecut, layer_boundaries=bounds)
    return data.cpu().numpy(), energy.cpu().numpy()


def reverse_norm_hgcal_device(voxels, e, emax=9999.0, emin=0.0001, max_deposit=2, layerE=None, showerMap="log", dataset_num=111,
                              embed=False, NN_embed=None, sparse_decoding=False, sparse_per_batch=False, seed=None, u_bank=None,
                              stream=None):
    """ReverseNormHGCal on the device: voxels (N, 1, L, A, R) with `embed` (decoded through NN_embed) or (N, 1, L, cells);
    e (N, k) generator info (column 0 = incident energy).  Returns (data (N, L, cells) tensor, gen_out (N, k) tensor)."""
    c = DATASET_PARAMS[dataset_num]
    kind, a, b, cap = _map_constants(c, showerMap, max_deposit)
    layer = "layer" in showerMap
    n = voxels.shape[0]
    nseg = int(voxels.shape[2])
    v, ed, le = _prep(voxels, e, layerE, nseg, layer)
    # HGCal_utils.py:196-200: emin / emax may be per-column arrays of the generator info; column 0 is the incident energy
    lo_t = torch.as_tensor(np.asarray(emin, np.float32), device=ed.device)
    hi_t = torch.as_tensor(np.asarray(emax, np.float32), device=ed.device)
    gen_out = lo_t + (hi_t - lo_t) * ed                                # (tiny (N, k) tensor)
    emin, emax = float(np.asarray(emin, np.float32).reshape(-1)[0]), float(np.asarray(emax, np.float32).reshape(-1)[0])
    common = dict(nseg=nseg, alpha=1e-8, guard_eps=1e-8, energy_mode=1, emin=float(emin), emax=float(emax),
                  max_deposit=float(max_deposit), e_stride=int(ed.shape[1]), totalE_mean=c["totalE_mean"], totalE_std=c["totalE_std"],
                  layers_mean=c["layers_mean"], layers_std=c["layers_std"])
    if not embed:
        data, _ = _run(dict(common, map_kind=kind, a_scale=a, a_shift=b, log_cap=cap), v, ed, le, None, stream)
        return data.reshape(n, nseg, -1), gen_out
    if NN_embed is None:
        raise ValueError("embed=True needs the HGCalConverter (NN_embed)")
    # stage 1: inverse map on the grid (+ the converter's normalisation, HGCal_utils.py:660-661); no layers, no energy scale
    post = (float(NN_embed.embed_std), float(NN_embed.embed_mean)) if getattr(NN_embed, "norm", False) else (1.0, 0.0)
    grid, _ = _run(dict(common, map_kind=kind, a_scale=a, a_shift=b, log_cap=cap, post_scale=post[0], post_shift=post[1],
                        final_scale=0), v, ed, None, None, stream)
    # stage 2: GLaM decoding to the cells (the normalisation was applied above: call the un-normalised decoder)
    norm_saved = getattr(NN_embed, "norm", False)
    NN_embed.norm = False
    try:
        cells = NN_embed.dec(grid.reshape(voxels.shape), sparse_decoding=sparse_decoding, sparse_per_batch=sparse_per_batch,
                             **({"seed": seed, "u_bank": u_bank} if sparse_decoding else {}))
    finally:
        NN_embed.norm = norm_saved
    # stage 3: per-layer renormalisation and energy scale on the cells
    data, _ = _run(dict(common, map_kind=2), cells.reshape(n, -1).contiguous(), ed, le, None, stream)
    return data.reshape(n, nseg, -1), gen_out


def ReverseNormHGCal(voxels, e, shape=None, emax=9999.0, emin=0.0001, max_deposit=2, logE=True, layerE=None, showerMap="log",
                     dataset_num=2, orig_shape=False, ecut=0.0, embed=False, NN_embed=None, binning_file="", config=None,
                     sparse_decoding=False, sparse_per_batch=False):
    """Reference signature (utils/HGCal_utils.py:167-187); numpy or tensors in, numpy `(data, gen_out)` out."""
    dev = torch.device("cuda", torch.cuda.current_device())
    as_t = lambda a_: a_.to(dev) if torch.is_tensor(a_) else torch.as_tensor(np.asarray(a_), device=dev)
    if embed and NN_embed is None:                                    # HGCal_utils.py:239-247
        from .geometry import HGCalConverter
        NN_embed = HGCalConverter(bins=config["SHAPE_FINAL"], geom_file=binning_file, trainable=config.get("TRAINABLE_EMBED", False),
                                  device=dev)
        NN_embed.init(norm=True, dataset_num=dataset_num)
    data, gen_out = reverse_norm_hgcal_device(as_t(voxels), as_t(e), emax=emax, emin=emin, max_deposit=max_deposit,
                                              layerE=None if layerE is None else as_t(layerE), showerMap=showerMap,
                                              dataset_num=dataset_num, embed=embed, NN_embed=NN_embed,
                                              sparse_decoding=sparse_decoding, sparse_per_batch=sparse_per_batch)
    return data.cpu().numpy(), gen_out.cpu().numpy()
