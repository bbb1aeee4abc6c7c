"""CPU oracle of one training step's loss and gradients (TEST INFRASTRUCTURE ONLY; nothing under calodiffusion_b200/ imports it).

Restates, with plain torch CPU ops + autograd through the functional denoiser of oracle/unet_oracle.py:
  Loss.__call__            models/loss.py:118-142   sigma = exp(rnd_normal * P_std + P_mean)   ('log' noise schedule: continuous)
  hybrid_weight            models/loss.py:168-179   x_noisy = data + sigma*noise; x0 = denoise(x_noisy); weight = 1 + 1/sigma^2
  l2_loss                  models/loss.py:103-104   sum(weight*(pred-target)^2) / (mean(weight) * numel)
Pinned to the live reference's `compute_loss(...).backward()` by tests/golden/train_ds2.npz (oracle/make_golden_train.py)."""
import numpy as np
import torch

from .unet_oracle import DenoiserOracle


def sigma_from_rnd(cfg, rnd_normal):
    """models/loss.py:16-25, 139-140."""
    if "log" in cfg.get("NOISE_SCHED", "linear"):
        p_mean, p_std = -1.2, 1.2
    else:
        p_mean, p_std = -1.0, 1.0
    return (rnd_normal * p_std + p_mean).exp()


def sigma_from_time(n_steps, time):
    """models/loss.py:124-137 (discrete noise schedules, e.g. 'cosine'): sigma_t = sqrt(1-abar_t)/sqrt(abar_t) of the cosine table
    (utils/sampling.py:16-24, loss.py:78-95)."""
    import math
    x = torch.linspace(0, n_steps, n_steps + 1)
    ac = torch.cos(((x / n_steps) + 0.008) / 1.008 * math.pi * 0.5) ** 2
    ac = ac / ac[0]
    betas = torch.clip(1 - (ac[1:] / ac[:-1]), 0.0001, 0.9999)
    abar = torch.cumprod(1.0 - betas, dim=0)
    return (torch.sqrt(1.0 - abar) / torch.sqrt(abar)).gather(-1, time.long())


def sigma_for(cfg, rnd_normal=None, time=None):
    """The noise level Loss.__call__ would use: continuous ('log' in NOISE_SCHED) from rnd_normal, else discrete from time."""
    if "log" in cfg.get("NOISE_SCHED", "linear"):
        return sigma_from_rnd(cfg, rnd_normal)
    return sigma_from_time(cfg.get("NSTEPS", 400), time)


def loss_and_grads(cfg, sd, data, E, layers, noise, sigma):
    """Returns (loss float, {name: grad tensor}) for the parameters in `sd` (CPU fp32)."""
    leaf = {k: v.detach().clone().to(torch.float32).requires_grad_(True) for k, v in sd.items()}
    o = DenoiserOracle(cfg, sd)
    o.sd = leaf
    sig = sigma.reshape([-1] + [1] * (data.dim() - 1))
    x_noisy = data + sig * noise
    x0 = o.denoise(x_noisy, E, sigma, layers)
    weight = (1.0 + 1.0 / sig ** 2)
    loss = (weight * (x0 - data) ** 2).sum() / (torch.mean(weight) * float(np.prod(data.shape)))
    loss.backward()
    return float(loss), {k: v.grad for k, v in leaf.items()}
