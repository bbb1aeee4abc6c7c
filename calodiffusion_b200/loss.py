"""Loss registry mirror (models/loss.py) — only what the sampling path reads: EDM scalings and the cosine tables.

The training losses themselves are out of scope for this backend (SURVEY §8 row T is "next")."""
import math

import torch


def cosine_beta_schedule(nsteps, s=0.008):
    """utils/sampling.py:16-24 (fp32, same op order)."""
    x = torch.linspace(0, nsteps, nsteps + 1)
    alphas_cumprod = torch.cos(((x / nsteps) + s) / (1 + s) * math.pi * 0.5) ** 2
    alphas_cumprod = alphas_cumprod / alphas_cumprod[0]
    betas = 1 - (alphas_cumprod[1:] / alphas_cumprod[:-1])
    return torch.clip(betas, 0.0001, 0.9999)


class Loss:
    def __init__(self, config, n_steps, loss_type="l1"):
        self.config = config
        self.update_step(n_steps)
        self.discrete_time = True
        self.P_mean, self.P_std, self.sigma_data = -1, 1, 0.5
        if "log" in config.get("NOISE_SCHED", "linear"):  # models/loss.py:16-25
            self.discrete_time = False
            self.P_mean, self.P_std, self.sigma_data = -1.2, 1.2, 1.0
        self.loss_type = loss_type

    def get_scaling(self, sigma):
        """models/loss.py:29-41."""
        return {
            "c_skip": self.sigma_data ** 2 / (sigma ** 2 + self.sigma_data ** 2),
            "c_out": sigma * self.sigma_data / (sigma ** 2 + self.sigma_data ** 2) ** 0.5,
            "c_in": 1 / (sigma ** 2 + self.sigma_data ** 2) ** 0.5,
        }

    def update_step(self, steps):
        """models/loss.py:78-95."""
        self.n_steps = steps
        betas = cosine_beta_schedule(self.n_steps)
        alphas_cumprod = torch.cumprod(1.0 - betas, dim=0)
        alphas_cumprod_prev = torch.nn.functional.pad(alphas_cumprod[:-1], (1, 0), value=1.0)
        self.sqrt_alphas_cumprod = torch.sqrt(alphas_cumprod)
        self.sqrt_one_minus_alphas_cumprod = torch.sqrt(1.0 - alphas_cumprod)
        self.posterior_variance = betas * (1.0 - alphas_cumprod_prev) / (1.0 - alphas_cumprod)

    def __call__(self, *a, **k):
        raise NotImplementedError("training losses are not part of the B200 sampling backend")


class hybrid_weight(Loss):
    """models/loss.py:163-179; the only TRAINING_OBJ the shipped configs use, and the only one whose
    denoise branch (c_skip*x + c_out*pred, models/calodiffusion.py:166-167) the CUDA epilogue implements."""
