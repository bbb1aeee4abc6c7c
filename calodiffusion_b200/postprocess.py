"""Device-side mirror of `ReverseNormCaloChall` (utils/utils.py:446-573), the inverse of the training-set
transformations applied to sampled showers (models/diffusion.py:171-191) — SURVEY 8(f) row 1.

Same name, argument meaning and return value `(data (N, voxels) numpy, energy (N,1) numpy)` as the reference for
the CaloChallenge datasets 2 and 3 with the shipped '*logit-norm' shower maps; the arithmetic runs in one CUDA
kernel (`cd_reverse_norm`) on the showers while they are still on the device, instead of numpy on the host.
Unsupported variants (datasets 0/1 with GeomConverter, quantile / log maps, logE=False) raise ValueError.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib

# utils/consts.py:82-117
_PARAMS = {
    2: dict(logit_mean=-12.8564, logit_std=1.9123, totalE_mean=0.3926, totalE_std=0.05546, layers_mean=-6.35551,
            layers_std=3.90699),
    3: dict(logit_mean=-13.4753, logit_std=1.1070, totalE_mean=0.0, totalE_std=1.0, layers_mean=0.0, layers_std=1.0),
}


def reverse_norm_device(voxels, e, emax=9999.0, emin=0.0001, max_deposit=2, logE=True, layerE=None, showerMap="log",
                        dataset_num=2, ecut=0.0, stream=None):
    """Tensors in, tensors out (all on the voxels' CUDA device): returns (data (N, voxels), energy (N,1))."""
    if dataset_num not in _PARAMS:
        raise ValueError("device ReverseNorm supports dataset_num 2 and 3 only")
    if "logit" not in showerMap or "norm" not in showerMap or "quantile" in showerMap:
        raise ValueError(f"device ReverseNorm supports the '*logit-norm' shower maps only (got {showerMap!r})")
    if not logE:
        raise ValueError("device ReverseNorm supports logE=True only")
    if "layer" in showerMap and layerE is None:
        raise AssertionError("layerE is required for 'layer' shower maps")   # utils.py:520
    if not voxels.is_cuda:
        raise RuntimeError("reverse_norm_device needs CUDA tensors; there is no CPU fallback")
    lib = _lib.load()
    c = _PARAMS[dataset_num]
    n = voxels.shape[0]
    if voxels.dim() != 5:
        raise ValueError("voxels must be (N, 1, Z, P, R)")
    Z, PR = int(voxels.shape[2]), int(voxels.shape[3] * voxels.shape[4])
    dev = voxels.device
    v = voxels.to(torch.float32).contiguous()
    ed = e.to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
    if ed.numel() != n:
        raise ValueError("e must hold one value per shower")
    le = None
    if "layer" in showerMap:
        le = layerE.to(device=dev, dtype=torch.float32).reshape(n, -1).contiguous()
        if le.shape[1] != Z + 1:
            raise ValueError(f"layerE must be (N, {Z + 1}) = [totalE, layers]")
    out = torch.empty((n, Z * PR), device=dev, dtype=torch.float32)
    energy = torch.empty((n, 1), device=dev, dtype=torch.float32)
    d = _lib.ReverseNormDesc(Z, PR, c["logit_mean"], c["logit_std"], c["totalE_mean"], c["totalE_std"], c["layers_mean"],
                             c["layers_std"], float(emin), float(emax), float(max_deposit), float(ecut))
    st = torch.cuda.current_stream(dev).cuda_stream if stream is None else stream
    for lo in range(0, n, 32768):     # grid.y limit
        hi = min(n, lo + 32768)
        _lib.check(lib.cd_reverse_norm(C.byref(d), v[lo:hi].data_ptr(), ed[lo:hi].data_ptr(),
                                       le[lo:hi].data_ptr() if le is not None else None, out[lo:hi].data_ptr(),
                                       energy[lo:hi].data_ptr(), hi - lo, st), "reverse_norm")
    return out, energy


def ReverseNormCaloChall(voxels, e, emax=9999.0, emin=0.0001, config=None, shape=None, binning_file="", max_deposit=2,
                         logE=True, layerE=None, showerMap="log", dataset_num=2, orig_shape=False, ecut=0.0, **kwargs):
    """Reference signature (utils/utils.py:446-462).  `voxels`, `e`, `layerE` may be numpy arrays (as in the reference)
    or tensors already on the device; returns numpy `(data, energy)` like the reference."""
    if orig_shape or dataset_num not in _PARAMS:
        raise ValueError("the B200 backend covers dataset_num 2 / 3 in network shape only")
    dev = torch.device("cuda", torch.cuda.current_device())
    as_t = lambda a: a.to(dev) if torch.is_tensor(a) else torch.as_tensor(np.asarray(a), device=dev)
    data, energy = reverse_norm_device(as_t(voxels), as_t(e), emax=emax, emin=emin, max_deposit=max_deposit, logE=logE,
                                       layerE=None if layerE is None else as_t(layerE), showerMap=showerMap,
                                       dataset_num=dataset_num, ecut=ecut)
    return data.cpu().numpy(), energy.cpu().numpy()
