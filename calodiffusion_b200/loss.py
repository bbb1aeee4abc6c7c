"""Loss registry mirror (models/loss.py): EDM scalings, the cosine tables, and the training objective the shipped configs use
(`hybrid_weight`, LOSS_TYPE l2).  The arithmetic of the loss - noising, the denoiser forward, the weighted l2 and the whole
backward pass - runs in the CUDA library (`cd_train_step`); this module only draws / maps the noise levels as the reference does."""
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

    def __call__(self, model, data, E, noise=None, time=None, layers=None, rnd_normal=None):
        """models/loss.py:118-142: the per-shower noise level, then the objective's loss_function."""
        const_shape = (data.shape[0], *((1,) * (len(data.shape) - 1)))
        if noise is None:
            noise = torch.randn_like(data)
        if self.discrete_time:
            if time is None:
                time = torch.randint(0, self.n_steps, (data.size()[0],), device=data.device).long()
            t = time.cpu()
            sigma = (self.sqrt_one_minus_alphas_cumprod.gather(-1, t) / self.sqrt_alphas_cumprod.gather(-1, t)).reshape(const_shape)
            sigma = sigma.to(data.device)
        else:
            if rnd_normal is None:
                rnd_normal = torch.randn((data.size()[0],), device=data.device)
            sigma = (rnd_normal * self.P_std + self.P_mean).exp().reshape(const_shape)
        return self.loss_function(model, data, E, sigma=sigma, noise=noise, layers=layers)

    def loss_function(self, model, data, E, sigma=None, noise=None, layers=None):
        raise NotImplementedError


class hybrid_weight(Loss):
    """models/loss.py:163-179; the only TRAINING_OBJ the shipped configs use, and the only one whose
    denoise branch (c_skip*x + c_out*pred, models/calodiffusion.py:166-167) the CUDA epilogue implements."""

    def loss_function(self, model, data, E, sigma=None, noise=None, layers=None):
        """x_noisy = data + sigma*noise; x0 = denoise(x_noisy); weight = 1 + 1/sigma^2; l2 (models/loss.py:168-179, 103-104)."""
        if self.loss_type != "l2":
            raise NotImplementedError(f"the B200 training step implements LOSS_TYPE 'l2' (got {self.loss_type!r})")
        return model.training_loss(data, E, sigma, noise, layers, read_loss=getattr(model, "_read_loss", True))
