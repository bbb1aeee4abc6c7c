"""calodiffusion_b200 — B200-native (sm_100a) backend for CaloDiffusion's denoiser / sampling-loop hot path.

Host side mirrors the reference interface for that path (reference paths relative to calodiffusion/):
  calodiffusion_b200.diffusion.Diffusion          <- models/diffusion.py:18-104 (contract + sample)
  calodiffusion_b200.calodiffusion.CaloDiffusion  <- models/calodiffusion.py:9-173
  calodiffusion_b200.sample.{DDim,DDPM,Euler,Heun}<- models/sample.py:29-121, 577-726, 772-821
  calodiffusion_b200.loss.hybrid_weight           <- models/loss.py:9-41, 78-95, 163-179 (scalings/tables)
  calodiffusion_b200.utils.{LoadJson,load_attr}   <- utils/utils.py:439-443, 1047-1061
All arithmetic runs in hand-written CUDA inside libcalodiffusion_b200.so (C ABI: include/calodiffusion_b200.h).
There is no CPU fallback: constructing a model without the library or without an sm_100 GPU raises.
"""
__version__ = "0.1.0"
