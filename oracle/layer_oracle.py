"""CPU restatement of the layer-energy denoiser of LayerDiffusion (TEST INFRASTRUCTURE ONLY — never imported by the product).

  resnet_forward        <- models/models.py:372-457 (ResDense, ResNet.forward)
  LayerDenoiserOracle   <- models/layerdiffusion.py:111-114 (layer_forward) under CaloDiffusion.denoise
                           (models/calodiffusion.py:154-169, models/loss.py:29-41)
Pinned to the live reference by tests/golden/layer.npz (oracle/make_golden_layer.py).
"""
import torch
import torch.nn.functional as F


def _lin(sd, name, x):
    return F.linear(x, sd[name + ".weight"], sd[name + ".bias"])


def resnet_forward(sd, x, cond, time, pre=""):
    """x (B, dim_in), cond (B, cond_size), time (B,) -> (B, dim_in)."""
    c = _lin(sd, pre + "cond_mlp.4", F.gelu(_lin(sd, pre + "cond_mlp.2", F.gelu(_lin(sd, pre + "cond_mlp.0", cond)))))
    t = time.reshape(-1, 1)
    t = _lin(sd, pre + "time_mlp.5", F.gelu(_lin(sd, pre + "time_mlp.3", F.gelu(_lin(sd, pre + "time_mlp.1", t)))))
    cvec = torch.cat([c, t], dim=-1)
    h = _lin(sd, pre + "in_lay", x)
    i = 0
    while f"{pre}hidden_layers.{i}.dense1.0.weight" in sd:
        L = f"{pre}hidden_layers.{i}."
        a = F.gelu(_lin(sd, L + "dense1.0", h)) + _lin(sd, L + "embeder.1", F.gelu(cvec))
        h = F.gelu(_lin(sd, L + "dense2.0", a)) + h
        i += 1
    return _lin(sd, pre + "out_lay", h)


class LayerDenoiserOracle:
    """`denoise(x, E, sigma, layers)` of LayerDiffusion in its layer state; same call shape as DenoiserOracle so the
    oracle samplers (oracle.unet_oracle.sample_ddim / sample_edm) drive it unchanged."""

    def __init__(self, config, state_dict, dtype=torch.float32):
        self.config, self.dtype = config, dtype
        self.sd = {k: v.detach().to(dtype) for k, v in state_dict.items()}
        self.time_embed = config.get("TIME_EMBED", "sin")
        self.sigma_data = 1.0 if "log" in config.get("NOISE_SCHED", "linear") else 0.5

    def time_map(self, sigma):
        s = sigma.reshape(-1).to(torch.float32)
        if self.time_embed == "sigma":
            return (s / (1 + s ** 2).sqrt()).to(self.dtype)
        if self.time_embed == "log":
            return (0.5 * torch.log(s)).to(self.dtype)
        raise KeyError(self.time_embed)

    def forward(self, x, E, t_emb):
        return resnet_forward(self.sd, x.to(self.dtype), E.to(self.dtype), t_emb)

    def denoise(self, x, E, sigma, layers=None):
        x = x.to(self.dtype)
        sigma = sigma.to(self.dtype).reshape(-1, 1)
        sd2 = self.sigma_data ** 2
        c_skip = sd2 / (sigma ** 2 + sd2)
        c_out = sigma * self.sigma_data / (sigma ** 2 + sd2) ** 0.5
        c_in = 1 / (sigma ** 2 + sd2) ** 0.5
        return c_skip * x + c_out * self.forward(x * c_in, E, self.time_map(sigma))

    __call__ = denoise
