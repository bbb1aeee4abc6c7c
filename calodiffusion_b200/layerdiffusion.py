"""`LayerDiffusion(CaloDiffusion)` on the B200 CUDA library (mirror of models/layerdiffusion.py:12-163).

Two denoisers behind one object: the layer-energy model (a dense ResNet over (B, Z+1) vectors: total deposited energy +
one value per calorimeter layer, models/models.py:390-457) and the base U-Net.  At sampling time the layer model runs
first (`sample_layers`), its output becomes the `layers` conditioning of the base model.  Both run in the library: the
dense ResNet is `layer_mlp_kernel` (one launch per evaluation, update fused, same CUDA-graph loop as the U-Net) behind
`cd_layer_model_create`; `set_layer_state` swaps which engine handle the `Diffusion` entry points drive.
"""
import copy
import ctypes as C
import os

import torch
import torch.nn as nn

from . import _lib, utils
from .calodiffusion import CaloDiffusion, CondUnetB200


def resnet_param_shapes(dim_in, num_layers=3, hidden_dim=256, cond_emb_dim=128, cond_size=1):
    """State-dict names and shapes of the reference's dense ResNet (models/models.py:390-441), in registration order."""
    half = cond_emb_dim // 2
    s = {}

    def lin(name, cout, cin):
        s[name + ".weight"] = (cout, cin)
        s[name + ".bias"] = (cout,)

    lin("time_mlp.1", half // 2, 1)
    lin("time_mlp.3", half, half // 2)
    lin("time_mlp.5", half, half)
    lin("cond_mlp.0", half // 2, cond_size)
    lin("cond_mlp.2", half, half // 2)
    lin("cond_mlp.4", half, half)
    lin("in_lay", hidden_dim, dim_in)
    for i in range(num_layers - 1):
        lin(f"hidden_layers.{i}.embeder.1", hidden_dim, cond_emb_dim)
        lin(f"hidden_layers.{i}.dense1.0", hidden_dim, hidden_dim)
        lin(f"hidden_layers.{i}.dense2.0", hidden_dim, hidden_dim)
    lin("out_lay", dim_in, hidden_dim)
    return s


class LayerDiffusion(CaloDiffusion):
    """Creates a model which has two steps, a layer model and a base model (models/layerdiffusion.py:12-19)."""

    LAYER_NUM_LAYERS, LAYER_HIDDEN, LAYER_COND_DIM = 5, 256, 128     # layerdiffusion.py:36 + ResNet defaults

    def __init__(self, config, n_steps=400, loss_type="l2", max_batch=None):
        super().__init__(config, n_steps, loss_type, max_batch)
        self.layer_loss = False
        sampler_algo = self.config.get("LAYER_SAMPLER", "DDim")
        self.layer_sampler = utils.load_attr("sampler", sampler_algo)(self.config)
        self.layer_steps = self.config.get("LAYER_STEPS", n_steps)
        self.shape_pad = self.config.get("SHAPE_PAD")
        if self.shape_pad is None:
            self.shape_pad = self.config["SHAPE_FINAL"]
        # per-state engine bookkeeping (handle, parameter versions, io buffers, tensor widths), swapped by set_layer_state
        self._layer_dim = int(self.config["SHAPE_FINAL"][2]) + 1
        self._states = {
            False: None,
            True: dict(_handle=None, _handle_versions=None, _io={}, _voxels=self._layer_dim, _cond_size=3 if self.hgcal else 1,
                       layer_cond=False, do_embed=False, _max_batch=self._max_batch),
        }

    def init_model(self):
        cond_size = 3 if self.hgcal else 1
        dim_in = int(self.config["SHAPE_FINAL"][2]) + 1
        shapes = resnet_param_shapes(dim_in, self.LAYER_NUM_LAYERS, self.LAYER_HIDDEN, self.LAYER_COND_DIM, cond_size)
        self.layer_model = CondUnetB200(shapes).to(self.device)
        model = super().init_model()
        self.base_model = model
        return model

    # ------------------------------------------------------------------ state switching
    # `_max_batch` is per handle: each engine is built for its own largest batch (a larger later batch rebuilds only the
    # active state's handle, so the other state's handle must keep the value it was built with)
    _SWAPPED = ("_handle", "_handle_versions", "_io", "_voxels", "_cond_size", "layer_cond", "do_embed", "_max_batch")

    def set_layer_state(self, is_layer=False):
        """layerdiffusion.py:40-48: which of the two denoisers `forward` / `denoise` / the samplers drive."""
        is_layer = bool(is_layer)
        if is_layer == self.layer_loss:
            return
        self._states[self.layer_loss] = {k: getattr(self, k) for k in self._SWAPPED}
        for k, v in self._states[is_layer].items():
            setattr(self, k, v)
        self.layer_loss = is_layer
        self.model = self.layer_model if is_layer else self.base_model

    def _create_handle(self, lib, max_batch):
        if not self.layer_loss:
            return super()._create_handle(lib, max_batch)
        d = _lib.LayerDesc()
        d.dim_in, d.hidden, d.num_layers = self._layer_dim, self.LAYER_HIDDEN, self.LAYER_NUM_LAYERS
        d.cond_size, d.cond_dim = self._cond_size, self.LAYER_COND_DIM
        d.time_embed = 0 if self.time_embed == "sigma" else 1
        d.sigma_data = float(self.loss_function.sigma_data)
        d.max_batch = int(max_batch)
        h = lib.cd_layer_model_create(C.byref(d))
        if not h:
            raise RuntimeError("cd_layer_model_create failed: " + lib.cd_last_error().decode())
        return h

    def _destroy_handle(self):
        super()._destroy_handle()
        states = getattr(self, "_states", None)
        if states:
            other = states.get(not self.layer_loss)
            if other and other.get("_handle") is not None:
                self._lib.cd_unet_destroy(other["_handle"])
                other["_handle"] = None

    def _engine(self, batch):
        # a rebuilt handle of one state must not free the other state's: only the active handle is replaced here
        saved = self._states
        self._states = None
        try:
            return super()._engine(batch)
        finally:
            self._states = saved

    def layer_forward(self, x, E, time, **kwargs):
        """layerdiffusion.py:111-114 (add_RZPhi is the identity on 2-D input)."""
        if not self.layer_loss:
            raise RuntimeError("layer_forward needs set_layer_state(True)")
        return CaloDiffusion.forward(self, x, E, time, None)

    def forward(self, x, E, time, layers=None, controls=None, **kwargs):
        return CaloDiffusion.forward(self, x, E, time, None if self.layer_loss else layers, controls)

    # ------------------------------------------------------------------ checkpoints
    def load_layer_model_state(self, strict=True):
        """layerdiffusion.py:57-82: config['layer_model'] (or <checkpoint>/checkpoint.pth) -> layer_model."""
        try:
            path = self.config["layer_model"]
        except KeyError:
            path = os.path.join(self.config.get("checkpoint", ""), "checkpoint.pth")
            if not os.path.exists(path):
                raise RuntimeError("Could not load layer model from either config or checkpoint path")
        sd = torch.load(path, map_location="cpu", weights_only=False)
        sd = sd if "model_state_dict" not in sd else sd["model_state_dict"]
        if "layer_model" in set(k.split(".")[0] for k in sd.keys()):
            sd = {k.removeprefix("layer_model."): v for k, v in sd.items()}
        self.layer_model.load_state_dict({k: v for k, v in sd.items() if k in self.layer_model.state_dict()}, strict=strict)
        self._states[True]["_handle_versions"] = None

    def load_state_dict(self, state_dict, strict=True):
        """layerdiffusion.py:84-103: layer model from its own file, base model from `state_dict`."""
        self.load_layer_model_state(strict)
        prefixes = set(k.split(".")[0] for k in state_dict.keys())
        if "base_model" in prefixes:
            state_dict = {k.removeprefix("base_model."): v for k, v in state_dict.items() if k.startswith("base_model.")}
        elif "model" in prefixes:
            state_dict = {k.removeprefix("model."): v for k, v in state_dict.items() if k.startswith("model.")}
        out = self.base_model.load_state_dict(state_dict, strict)
        self._handle_versions = None
        return out

    def state_dict(self, *a, **k):
        sd = nn.Module.state_dict(self, *a, **k)
        sd["layer_model"] = self.layer_model.state_dict()
        return sd

    # ------------------------------------------------------------------ sampling
    def sample_layers(self, energy, layers, debug=False, sample_offset=None):
        """layerdiffusion.py:116-131: diffusion over the (B, Z+1) layer-energy vector, conditioned on E only."""
        self.set_layer_state(is_layer=True)
        try:
            start = self.noise_generation((energy.shape[0], self.shape_pad[2] + 1)).to(torch.float32)
            x, _, _ = self.layer_sampler(self, start, energy, layers, self.layer_steps, sample_offset or 0, debug)
        finally:
            self.set_layer_state(is_layer=False)
        return x

    def sample(self, energy, layers, num_steps=400, debug=False, sample_offset=None, return_layers=False):
        """layerdiffusion.py:133-163: layer energies first, then the showers conditioned on them; dict result."""
        generated_shape = list(copy.copy(self._data_shape))
        generated_shape.insert(0, energy.shape[0])
        start = self.noise_generation(generated_shape).to(torch.float32)
        layers = self.sample_layers(energy, layers=None, debug=debug, sample_offset=sample_offset)
        x, xs, x0s = self.sampler_algorithm(self, start, energy, layers, num_steps, sample_offset or 0, debug)
        out = {"x": x.detach().cpu().numpy()}
        if debug:
            out["xs"] = xs
            out["x0s"] = x0s
        if return_layers:
            out["layers"] = layers
        return out
