"""`Diffusion` contract mirror (models/diffusion.py:18-104)."""
import copy
from abc import ABC, abstractmethod

import torch

from . import utils


class Diffusion(torch.nn.Module, ABC):
    def __init__(self, config, n_steps=400, loss_type="l2"):
        super().__init__()
        self.config = config if isinstance(config, dict) else utils.LoadJson(config)
        if not torch.cuda.is_available():
            raise RuntimeError("calodiffusion_b200 needs a CUDA (sm_100a) device; there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.nsteps = n_steps
        self.loss_type = loss_type
        self.hgcal = self.pre_embed = False
        loss_algo = self.config.get("TRAINING_OBJ", "noise_pred")
        self.loss_function = utils.load_attr("loss", loss_algo)(self.config, self.nsteps, self.loss_type)
        sampler_algo = self.config.get("SAMPLER", "DDim")
        self.sampler_algorithm = utils.load_attr("sampler", sampler_algo)(self.config)
        self.NN_embed = None
        if "orig" not in self.config.get("SHOWER_EMBED", ""):
            self._data_shape = self.config["SHAPE_PAD"][1:]
        else:
            self._data_shape = self.config["SHAPE_ORIG"][1:]

    @abstractmethod
    def init_model(self):
        raise NotImplementedError

    @abstractmethod
    def noise_generation(self, shape):
        return torch.randn(shape, device=self.device, dtype=torch.float32)

    @abstractmethod
    def forward(self, x, E, time, layers, controls=None):
        raise NotImplementedError

    @abstractmethod
    def __call__(self, x, **kwargs):
        raise NotImplementedError

    def sample(self, energy, layers, num_steps=400, debug=False, sample_offset=0):
        """models/diffusion.py:77-104: start noise -> sampler -> numpy on host."""
        generated_shape = list(copy.copy(self._data_shape))
        generated_shape.insert(0, energy.shape[0])
        start = self.noise_generation(generated_shape)
        x, xs, x0s = self.sampler_algorithm(self, start, energy, layers, num_steps, sample_offset, debug)
        if debug:
            return x.detach().cpu().numpy(), xs, x0s
        return x.detach().cpu().numpy()

    def compute_loss(self, data, energy, noise, layers, time=None, rnd_normal=None):
        """models/diffusion.py:106-110: loss of one training step.  Like the reference, `time` is accepted but not forwarded
        (a discrete noise schedule draws its own steps inside Loss.__call__); call `self.loss_function(...)` to pass it."""
        return self.loss_function(self, data, energy, noise=noise, layers=layers, rnd_normal=rnd_normal)
