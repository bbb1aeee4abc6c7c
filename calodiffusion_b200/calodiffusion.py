"""`CaloDiffusion(Diffusion)` on the B200 CUDA library (mirror of models/calodiffusion.py:9-173).

`init_model()` returns an `nn.Module` whose parameters carry the reference's state-dict names and shapes
(`model.init_conv.conv.weight`, `model.downs.0.0.block1.proj.conv.weight`, ... — SURVEY §5), so reference
`.pt/.pth` checkpoints load unchanged; the arithmetic is not in that module but in libcalodiffusion_b200.so,
to which the weights are handed (and re-handed whenever a parameter's version counter changes).
"""
import ctypes as C
import math
from collections import OrderedDict

import numpy as np
import torch
import torch.nn as nn

from . import _lib, geometry, utils
from .diffusion import Diffusion


def level_shapes(shape_zpr, n_levels):
    """models/models.py:619-635: Z -> ceil(Z/2) (compress_Z), P -> P//2, R -> R//2 per down-sampling."""
    out = [tuple(shape_zpr)]
    for _ in range(n_levels - 1):
        z, p, r = out[-1]
        out.append((math.ceil(z / 2.0), p // 2, r // 2))
    return out


def unet_param_shapes(layer_sizes, in_channels, cond_size, cond_dim, shape_zpr):
    """Name -> shape of every CondUnet parameter for the shipped topology (cylindrical, compress_Z,
    block_attn, mid_attn, identity time/cond embedding), in the reference's registration order
    (models/models.py:556-699)."""
    sp = OrderedDict()
    half = cond_dim // 2
    hid = max(cond_size, half // 2)
    ls = list(layer_sizes)
    nl = len(ls) - 1
    shapes = level_shapes(shape_zpr, nl)

    def conv(name, cout, cin, k, bias=True):
        sp[name + ".conv.weight"] = (cout, cin) + tuple(k)
        if bias:
            sp[name + ".conv.bias"] = (cout,)

    def lin(name, cout, cin):
        sp[name + ".weight"] = (cout, cin)
        sp[name + ".bias"] = (cout,)

    def resblock(pre, cin, cout, cond=True):
        if cond:
            lin(pre + ".mlp.1", cout, cond_dim)
        for blk, ci in (("block1", cin), ("block2", cout)):
            conv(f"{pre}.{blk}.proj", cout, ci, (3, 3, 3))
            sp[f"{pre}.{blk}.norm.weight"] = (cout,)
            sp[f"{pre}.{blk}.norm.bias"] = (cout,)
        if cin != cout:
            conv(pre + ".res_conv", cout, cin, (1, 1, 1))

    def attn(pre, c):
        conv(pre + ".fn.fn.to_qkv", 96, c, (1, 1, 1), bias=False)
        conv(pre + ".fn.fn.to_out.0", c, 32, (1, 1, 1))
        sp[pre + ".fn.fn.to_out.1.weight"] = (c,)
        sp[pre + ".fn.fn.to_out.1.bias"] = (c,)
        sp[pre + ".fn.norm.weight"] = (c,)
        sp[pre + ".fn.norm.bias"] = (c,)

    conv("init_conv", ls[0], in_channels, (3, 3, 3))
    lin("time_mlp.1", half // 2, 1)
    lin("time_mlp.3", half, half // 2)
    lin("time_mlp.5", half, half)
    lin("cond_mlp.0", hid, cond_size)
    lin("cond_mlp.2", half, hid)
    lin("cond_mlp.4", half, half)
    for i in range(nl):
        resblock(f"downs.{i}.0", ls[i], ls[i + 1])
        resblock(f"downs.{i}.1", ls[i + 1], ls[i + 1])
        if i < nl - 1:
            conv(f"downs.{i}.2", ls[i + 1], ls[i + 1], (3, 4, 4))
    for i in range(nl):
        lev = nl - 1 - i
        dim_in, dim_out = ls[lev], ls[lev + 1]
        resblock(f"ups.{i}.0", dim_out * 2, dim_in)
        resblock(f"ups.{i}.1", dim_in, dim_in)
        if i < nl - 1:
            # Upsample from level `lev` to `lev-1`: Z kernel 4 iff the finer level's Z is even (models.py:624-628,337)
            zk = 4 if shapes[lev - 1][0] % 2 == 0 else 3
            sp[f"ups.{i}.2.convTrans.weight"] = (dim_in, dim_in, zk, 4, 4)
            sp[f"ups.{i}.2.convTrans.bias"] = (dim_in,)
    for i in range(nl):
        attn(f"downs_attn.{i}", ls[i + 1])
    for i in range(nl):
        attn(f"ups_attn.{i}", ls[nl - 1 - i])
    resblock("mid_block1", ls[nl], ls[nl])
    attn("mid_attn", ls[nl])
    resblock("mid_block2", ls[nl], ls[nl])
    resblock("final_conv.0", ls[1], ls[0], cond=False)
    conv("final_conv.1", 1, ls[0], (1, 1, 1))
    return sp


class CondUnetB200(nn.Module):
    """Parameter container with the reference's names; default init follows torch's Conv/Linear/GroupNorm
    defaults (kaiming-uniform(a=sqrt(5)) == U(+-1/sqrt(fan_in)))."""

    def __init__(self, shapes):
        super().__init__()
        self._shapes = shapes
        for name, shape in shapes.items():
            parts = name.split(".")
            mod = self
            for p in parts[:-1]:
                if p not in mod._modules:
                    mod.add_module(p, nn.Module())
                mod = mod._modules[p]
            if ".norm." in name or ".to_out.1." in name:
                t = torch.ones(shape) if parts[-1] == "weight" else torch.zeros(shape)
            else:
                if len(shape) > 1:
                    fan_in = int(np.prod(shape[1:]))
                    if "convTrans" in name:
                        fan_in = int(shape[0] * np.prod(shape[2:]))
                else:
                    wshape = shapes[name[: -len("bias")] + "weight"]
                    fan_in = int(np.prod(wshape[1:]))
                    if "convTrans" in name:
                        fan_in = int(wshape[0] * np.prod(wshape[2:]))
                bound = 1.0 / math.sqrt(max(fan_in, 1))
                t = (torch.rand(shape) * 2 - 1) * bound
            mod.register_parameter(parts[-1], nn.Parameter(t))

    def forward(self, *a, **k):
        raise RuntimeError("CondUnetB200 holds parameters only; call the owning CaloDiffusion (CUDA library) instead")


class _LossWithGrads(torch.autograd.Function):
    """Connects the loss value computed by `cd_train_step` to the torch parameters: backward() returns the gradients the
    library already left in its flat buffer (divided by the loss scale), so `loss.backward(); optimizer.step()` works unchanged."""

    @staticmethod
    def forward(ctx, loss, model, *params):
        ctx.model = model
        return loss.clone()

    @staticmethod
    def backward(ctx, grad_out):
        tr = ctx.model._train
        inv = grad_out / (tr["loss_scale"] * tr["grad_div"])
        grads = [(tr["G"][o:o + n] * inv).reshape(p.shape) for p, (o, n) in tr["slices"]]
        return (None, None, *grads)


class CaloDiffusion(Diffusion):
    def __init__(self, config, n_steps=400, loss_type="l2", max_batch=None):
        super().__init__(config, n_steps, loss_type)
        cfg = self.config
        self.pre_embed = "pre-embed" in cfg.get("SHOWER_EMBED", "")
        self.hgcal = cfg.get("HGCAL", False)
        if "FCN" in cfg.get("SHOWER_EMBED", ""):
            raise ValueError("SHOWER_EMBED 'FCN' (dense ResNet) is outside the B200 hot path")
        self.time_embed = cfg.get("TIME_EMBED", "sin")
        if self.time_embed not in ("sigma", "log"):
            raise KeyError(self.time_embed)  # models/calodiffusion.py:148-152
        self.dataset_num = cfg.get("DATASET_NUM", 2)
        self.training_objective = cfg.get("TRAINING_OBJ", "noise_pred")
        if "hybrid" not in type(self.loss_function).__name__:
            raise ValueError("??? Training obj %s" % type(self.loss_function).__name__)
        self.layer_cond = "layer" in cfg.get("SHOWERMAP", "")
        for key, want in (("CYLINDRICAL", True), ("COMPRESS_Z", True), ("BLOCK_ATTN", True), ("MID_ATTN", True)):
            if bool(cfg.get(key, False)) != want:
                raise ValueError(f"B200 backend supports {key}={want} only (all shipped configs)")
        if cfg.get("COND_EMBED", "sin") == "sin":
            raise ValueError("sinusoidal COND_EMBED is not used by any shipped config and is not supported")
        shape = cfg["SHAPE_FINAL"][1:]
        self._zpr = (int(shape[1]), int(shape[2]), int(shape[3]))
        self._voxels = int(np.prod(self._zpr))
        self._max_batch = int(max_batch or cfg.get("BATCH", 128))
        self._handle = None
        self._handle_versions = None
        self._train = None
        self._io = {}
        self._lib = _lib.load()
        _lib.check(self._lib.cd_device_check(), "device check")
        self.model = self.init_model()
        self.NN_embed = self.init_embedding_model()
        self.do_embed = self.NN_embed is not None and (not self.pre_embed)
        if self.do_embed:   # the caller-facing tensors are the flat voxels; the maps run inside every evaluation on the device
            if int(np.prod(self.NN_embed.flat_shape())) != int(np.prod(self._data_shape)):
                raise ValueError(f"geometry embedding maps {tuple(self.NN_embed.flat_shape())} voxels but the config's data "
                                 f"shape is {tuple(self._data_shape)}")
            self._voxels = int(np.prod(self._data_shape))
        self.use_graph = True
        self.conv_backend = 0
        self.check_finite = True

    # ------------------------------------------------------------------ contract
    def init_model(self):
        cfg = self.config
        in_channels = 1 + (2 if cfg.get("R_Z_INPUT", False) else 0) + (1 if cfg.get("PHI_INPUT", False) else 0)
        cond_size = 2 + cfg["SHAPE_FINAL"][2] if self.layer_cond else 1
        if self.hgcal:
            cond_size += 2
        self._in_channels, self._cond_size = in_channels, cond_size
        shapes = unet_param_shapes(cfg["LAYER_SIZE_UNET"], in_channels, cond_size, cfg["COND_SIZE_UNET"], self._zpr)
        return CondUnetB200(shapes).to(self.device)

    def init_embedding_model(self, embed=None):
        """models/calodiffusion.py:100-119.  Dataset 1: NNConverter from the binning XML (`BIN_FILE`).  HGCal without
        pre-embedding: HGCalConverter built from the geometry pickle `BIN_FILE` (geometry.HGCalConverter), or, when
        config['GEOM_EMBED'] = dict(enc_mat, dec_mat[, enc_mask, dec_mask]) is given, those matrices (geometry.MatrixConverter)."""
        cfg = self.config
        shower_embed = cfg.get("SHOWER_EMBED", "")
        if "NN" in shower_embed and not self.hgcal:
            particle = "photon" if self.dataset_num == 1 else "pion"
            bins = geometry.BinningXML(particle, cfg["BIN_FILE"])
            return geometry.NNConverter(bins=bins).to(device=self.device)
        if self.hgcal and not self.pre_embed:
            mats = cfg.get("GEOM_EMBED")
            trainable = cfg.get("TRAINABLE_EMBED", False)
            if mats is not None:      # matrices handed in directly (a checkpoint's, or a test's)
                return geometry.MatrixConverter(cfg["SHAPE_FINAL"], mats["enc_mat"], mats["dec_mat"], mats.get("enc_mask"),
                                                mats.get("dec_mask"), trainable=trainable).to(self.device)
            # models/calodiffusion.py:113-117: the GLaM maps built from the geometry pickle named by BIN_FILE
            conv = geometry.HGCalConverter(bins=cfg["SHAPE_FINAL"], geom_file=cfg["BIN_FILE"], device=self.device, trainable=trainable)
            if not trainable:
                conv.init(norm=self.pre_embed, dataset_num=self.dataset_num)
            return conv.to(self.device)
        return None

    def noise_generation(self, shape):
        return super().noise_generation(shape)

    def philox_normal(self, n_showers, seed, shower_offset=0, slot=0):
        """N(0,1) of shape (n_showers, *data_shape) from the library's Philox stream, keyed by global shower index
        (`shower_offset` + row): a sharded batch regenerates exactly the rows of the unsharded one."""
        out = torch.empty([int(n_showers)] + list(self._data_shape), device=self.device, dtype=torch.float32)
        _lib.check(self._lib.cd_fill_normal(out.data_ptr(), int(n_showers), self._voxels, int(seed) & (2 ** 64 - 1),
                                            int(shower_offset), int(slot), self._stream()), "fill_normal")
        return out

    def load_state_dict(self, state_dict, strict=True):
        """models/calodiffusion.py:31-37: strip a foreign top-level prefix (peeks at the 11th key)."""
        base = list(state_dict.keys())[10].split(".")[0]
        if base != "model":
            state_dict = {k.removeprefix(f"{base}."): v for k, v in state_dict.items() if k.split(".")[0] == base}
        out = super().load_state_dict(state_dict, strict)
        self._handle_versions = None
        return out

    # ------------------------------------------------------------------ engine plumbing
    def _desc(self, max_batch):
        cfg = self.config
        d = _lib.UnetDesc()
        d.Z, d.P, d.R = self._zpr
        ls = list(cfg["LAYER_SIZE_UNET"])
        d.n_sizes = len(ls)
        for i, v in enumerate(ls):
            d.layer_sizes[i] = int(v)
        d.in_channels = self._in_channels
        d.rz_input = int(bool(cfg.get("R_Z_INPUT", False)))
        d.phi_input = int(bool(cfg.get("PHI_INPUT", False)))
        d.cond_size = self._cond_size
        d.cond_dim = int(cfg["COND_SIZE_UNET"])
        d.groups = int(cfg.get("BLOCK_GROUPS", 8))
        d.dataset_num = int(self.dataset_num)
        d.time_embed = 0 if self.time_embed == "sigma" else 1
        d.sigma_data = float(self.loss_function.sigma_data)
        d.max_batch = int(max_batch)
        rc = utils.scaled_r_centres(self.dataset_num, self._zpr[2])
        for i, v in enumerate(rc):
            d.r_centres[i] = float(v)
        return d

    def _engine(self, batch):
        self.sync_params_from_engine()
        versions = tuple(p._version for p in self.model.parameters())
        if self.do_embed:
            versions += self.NN_embed._weights_version()
        if self._handle is not None and versions == self._handle_versions and batch <= self._max_batch:
            return self._handle
        if batch > self._max_batch:
            self._max_batch = int(batch)
        self._destroy_handle()
        lib = self._lib
        h = self._create_handle(lib, self._max_batch)
        try:
            for name, p in self.model.named_parameters():
                t = p.detach().to("cpu", torch.float32).contiguous()
                shp = (C.c_int64 * t.dim())(*t.shape)
                _lib.check(lib.cd_unet_set_weight(h, ("model." + name).encode(), t.data_ptr(), shp, t.dim()), "set_weight")
            if self.do_embed:
                for which, m in enumerate(self.NN_embed.csr_maps()):
                    _lib.check(lib.cd_unet_set_geom_map(h, which, m["n_rows"], m["n_cols"], m["rowptr"].ctypes.data,
                                                        m["col"].ctypes.data, m["val"].ctypes.data), "set_geom_map")
            _lib.check(lib.cd_unet_set_option(h, b"conv_backend", int(self.conv_backend)), "set_option")
            _lib.check(lib.cd_unet_set_option(h, b"check_finite", int(self.check_finite)), "set_option")
            _lib.check(lib.cd_unet_finalize(h), "finalize")
            if int(lib.cd_unet_voxels(h)) != self._voxels:
                raise RuntimeError(f"engine works on {int(lib.cd_unet_voxels(h))} voxels per shower, host expects {self._voxels}")
        except Exception:
            lib.cd_unet_destroy(h)
            raise
        self._handle, self._handle_versions = h, versions
        return h

    def _create_handle(self, lib, max_batch):
        desc = self._desc(max_batch)
        h = lib.cd_unet_create(C.byref(desc))
        if not h:
            raise RuntimeError("cd_unet_create failed: " + lib.cd_last_error().decode())
        return h

    def set_conv_backend(self, backend):
        """0: tcgen05 implicit GEMM where supported (default); 1: generic SIMT kernel everywhere."""
        self.conv_backend = int(backend)
        if self._handle is not None:
            _lib.check(self._lib.cd_unet_set_option(self._handle, b"conv_backend", self.conv_backend), "set_option")

    def set_check_finite(self, on=True):
        """True (default): `run_program` raises RuntimeError when a denoiser evaluation produced non-finite output (activations
        left the fp16 range, |c_in*x| >= 65504); False: the non-finite showers are returned, as the reference would."""
        self.check_finite = bool(on)
        if self._handle is not None:
            _lib.check(self._lib.cd_unet_set_option(self._handle, b"check_finite", int(self.check_finite)), "set_option")

    def launch_count(self):
        return int(self._lib.cd_unet_launch_count(self._handle)) if self._handle else 0

    def set_profile(self, mode):
        """1: time every op with CUDA events (eager, diagnostic); 2: same and reset the accumulators; 0: off."""
        if self._handle is None:
            raise RuntimeError("no engine yet: run one evaluation first")
        _lib.check(self._lib.cd_unet_set_option(self._handle, b"profile", int(mode)), "set_option")

    def profile_report(self):
        """{op name: dict(count, ms, flops, bytes)} accumulated while profiling was on."""
        buf = C.create_string_buffer(1 << 16)
        _lib.check(self._lib.cd_unet_profile_report(self._handle, buf, len(buf)), "profile_report")
        out = {}
        for line in buf.value.decode().splitlines():
            name, cnt, ms, fl, by = line.split()
            out[name] = dict(count=int(cnt), ms=float(ms), flops=float(fl), bytes=float(by))
        return out

    def set_train_profile(self, mode):
        """`set_profile` for the training handle: per-op timing of forward and backward (diagnostic)."""
        _lib.check(self._lib.cd_unet_set_option(self._train["h"], b"profile", int(mode)), "set_option")

    def train_profile_report(self):
        h, self._handle = self._handle, self._train["h"]
        try:
            return self.profile_report()
        finally:
            self._handle = h

    def workspace_bytes(self):
        return int(self._lib.cd_unet_workspace_bytes(self._handle)) if self._handle else 0

    def _destroy_handle(self):
        if getattr(self, "_handle", None) is not None:
            self._lib.cd_unet_destroy(self._handle)
            self._handle = None
        if getattr(self, "_train", None) is not None:
            self._lib.cd_unet_destroy(self._train["h"])
            self._train = None

    # ------------------------------------------------------------------ training step
    class _DevBuf:
        """A raw device buffer of the library exposed through __cuda_array_interface__ (zero-copy torch view)."""

        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f4", "data": (int(ptr), False), "version": 2}

    def _train_engine(self, batch):
        """The training handle (option train: no buffer reuse, tape for the backward pass) plus torch views of its flat
        parameter / gradient / Adam buffers and the name -> (offset, numel) table."""
        tr = getattr(self, "_train", None)
        if tr is not None and batch <= tr["max_batch"]:
            return tr
        if tr is not None:
            self.sync_params_from_engine()
            self._lib.cd_unet_destroy(tr["h"])
            self._train = None
        lib = self._lib
        if self.do_embed:
            raise ValueError("training with an in-loop geometry embedding is not supported by the B200 backend")
        h = self._create_handle(lib, int(batch))
        try:
            _lib.check(lib.cd_unet_set_option(h, b"train", 1), "set_option")
            for name, p in self.model.named_parameters():
                t = p.detach().to("cpu", torch.float32).contiguous()
                shp = (C.c_int64 * t.dim())(*t.shape)
                _lib.check(lib.cd_unet_set_weight(h, ("model." + name).encode(), t.data_ptr(), shp, t.dim()), "set_weight")
            _lib.check(lib.cd_unet_finalize(h), "finalize")
        except Exception:
            lib.cd_unet_destroy(h)
            raise
        n = int(lib.cd_train_flat_size(h))
        view = lambda which: torch.as_tensor(self._DevBuf(lib.cd_train_buffer(h, which), n), device=self.device)
        table = {}
        buf = C.create_string_buffer(256)
        off, num = C.c_int64(), C.c_int64()
        for i in range(int(lib.cd_train_num_params(h))):
            _lib.check(lib.cd_train_param_info(h, i, buf, len(buf), C.byref(off), C.byref(num)), "param_info")
            table[buf.value.decode()] = (int(off.value), int(num.value))
        params = dict(self.model.named_parameters())
        slices = [(params[k[len("model."):]], table[k]) for k in table]
        self._train = dict(h=h, max_batch=int(batch), P=view(0), G=view(1), M=view(2), V=view(3), table=table, slices=slices,
                           versions=tuple(p._version for p in self.model.parameters()), stale=False, loss_scale=None, grad_div=1.0)
        return self._train

    def _sync_params_to_engine(self, tr):
        """torch parameters changed (an external optimizer stepped, a checkpoint was loaded): copy them into the flat master
        buffer on the device and re-derive the kernels' weight layouts."""
        versions = tuple(p._version for p in self.model.parameters())
        if versions == tr["versions"]:
            return
        with torch.no_grad():
            torch._foreach_copy_([tr["P"][o:o + n] for _, (o, n) in tr["slices"]], [p.detach().reshape(-1) for p, _ in tr["slices"]])
        _lib.check(self._lib.cd_train_refresh(tr["h"], self._stream()), "train_refresh")
        tr["versions"] = versions

    def sync_params_from_engine(self):
        """After `fused_adam_step` the master copy on the device is ahead of the torch parameters: copy it back (called before
        sampling, state_dict() and whenever a caller wants `model.parameters()` to be current)."""
        tr = getattr(self, "_train", None)
        if tr is None or not tr["stale"]:
            return
        with torch.no_grad():
            torch._foreach_copy_([p.detach().reshape(-1) for p, _ in tr["slices"]], [tr["P"][o:o + n] for _, (o, n) in tr["slices"]])
        tr["versions"] = tuple(p._version for p in self.model.parameters())
        tr["stale"] = False
        self._handle_versions = None

    def state_dict(self, *a, **k):
        self.sync_params_from_engine()
        return super().state_dict(*a, **k)

    def training_loss(self, data, E, sigma, noise, layers, loss_scale=None, read_loss=True):
        """hybrid_weight.loss_function on the device (models/loss.py:168-179): forward, loss and the whole backward pass in one
        library call.  Returns the loss as a 0-d tensor that is connected to the parameters: `loss.backward()` hands every
        parameter its gradient (drop-in for train/train_diffusion.py:58-62 with any torch optimizer); `fused_adam_step()` is
        the all-device alternative.  `self.grad_sync` (set by distributed.enable_data_parallel_training) is called on the flat
        gradient buffer right after the backward pass is enqueued: the ONE collective of a data-parallel step.
        read_loss=False skips the host read of the loss value (no synchronisation; returns None)."""
        B = data.shape[0]
        tr = self._train_engine(B)
        if not tr["stale"]:
            self._sync_params_to_engine(tr)
        d = self._dev(data, (B, self._voxels))
        nz = self._dev(noise, (B, self._voxels))
        sg = self._dev(sigma).reshape(-1)
        if sg.numel() == 1:
            sg = sg.expand(B).contiguous()
        cond = self._cond(E, layers)
        scale = float(loss_scale) if loss_scale is not None else float(B * self._voxels)
        _lib.check(self._lib.cd_train_step(tr["h"], d.data_ptr(), nz.data_ptr(), sg.data_ptr(), cond.data_ptr(), scale, B,
                                           self._stream()), "train_step")
        tr["loss_scale"] = scale
        sync = getattr(self, "grad_sync", None)
        tr["grad_div"] = float(sync(tr["G"])) if sync is not None else 1.0     # enqueued behind the backward kernels
        if not read_loss:
            return None
        out = C.c_double()
        _lib.check(self._lib.cd_train_loss(tr["h"], C.byref(out), self._stream()), "train_loss")
        loss = torch.tensor(out.value, device=self.device, dtype=torch.float32)
        return _LossWithGrads.apply(loss, self, *[p for p, _ in tr["slices"]])

    def grad_buffer(self):
        """Flat fp32 gradient of the last `training_loss` (times its loss scale): the tensor a data-parallel trainer all-reduces."""
        return self._train["G"]

    def fused_adam_step(self, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        """torch.optim.Adam's update (train/train.py:144 defaults) on the device-resident master parameters, followed by the
        re-derivation of the convolution weight layouts.  The torch parameters are refreshed lazily (sync_params_from_engine)."""
        tr = self._train
        _lib.check(self._lib.cd_train_adam(tr["h"], float(lr), float(betas[0]), float(betas[1]), float(eps),
                                           float(tr["loss_scale"] * tr["grad_div"]), self._stream()), "train_adam")
        tr["stale"] = True

    def __del__(self):
        try:
            self._destroy_handle()
        except Exception:
            pass

    def _dev(self, t, shape=None):
        t = t.to(device=self.device, dtype=torch.float32)
        if shape is not None:
            t = t.reshape(shape)
        return t.contiguous()

    def _cond(self, E, layers):
        """models/calodiffusion.py:89-90: cond = E || layers when the shower map is layer-conditioned."""
        E = self._dev(E).reshape(E.shape[0], -1)
        if self.layer_cond and layers is not None:
            E = torch.cat([E, self._dev(layers).reshape(E.shape[0], -1)], dim=1).contiguous()
        if E.shape[1] != self._cond_size:
            raise ValueError(f"conditioning width {E.shape[1]} != {self._cond_size} expected by the model")
        return E

    @staticmethod
    def _stream():
        return torch.cuda.current_stream().cuda_stream

    # ------------------------------------------------------------------ compute entry points
    def forward(self, x, E, time, layers, controls=None):
        """models/calodiffusion.py:86-98 (add_RZPhi + CondUnet.forward); `time` is the embedded time value."""
        if controls is not None:
            raise ValueError("ControlNet hidden states are not supported by the B200 backend")
        B = x.shape[0]
        xd = self._dev(x, (B, self._voxels))
        cond = self._cond(E, layers)
        t = self._dev(time, (B,))
        out = torch.empty_like(xd)
        h = self._engine(B)
        _lib.check(self._lib.cd_unet_forward(h, xd.data_ptr(), t.data_ptr(), cond.data_ptr(), out.data_ptr(), B,
                                             self._stream()), "forward")
        return out.reshape(x.shape)

    def denoise(self, x, E=None, sigma=None, layers=None, controls=None):
        """models/calodiffusion.py:154-169: c_skip*x + c_out*F(c_in*x; f(sigma), cond)."""
        if controls is not None:
            raise ValueError("ControlNet hidden states are not supported by the B200 backend")
        B = x.shape[0]
        xd = self._dev(x, (B, self._voxels))
        cond = self._cond(E, layers)
        s = self._dev(sigma).reshape(-1)
        if s.numel() == 1:
            s = s.expand(B).contiguous()
        out = torch.empty_like(xd)
        h = self._engine(B)
        _lib.check(self._lib.cd_denoise(h, xd.data_ptr(), s.data_ptr(), cond.data_ptr(), out.data_ptr(), B,
                                        self._stream()), "denoise")
        return out.reshape(x.shape)

    def __call__(self, x, **kwargs):
        return self.denoise(x, **kwargs)

    def run_program(self, start, E, layers, evals, start_scale, final_buf, noise_bank=None, seed=0, shower_offset=0,
                    n_rec_x=0, n_rec_x0=0):
        """Run a table of evaluations (see include/calodiffusion_b200.h: cd_eval) as one fused loop."""
        B = start.shape[0]
        # persistent, pointer-stable device buffers per (B, record sizes, bank size): the captured CUDA graph
        # bakes these addresses, so repeated calls replay it instead of re-capturing
        nbank = 0 if noise_bank is None else int(noise_bank.numel() // (B * self._voxels))
        key = (B, n_rec_x, n_rec_x0, nbank)
        io = self._io.get(key)
        if io is None:
            mk = lambda *shape: torch.empty(shape, device=self.device, dtype=torch.float32)
            io = dict(start=mk(B, self._voxels), cond=mk(B, self._cond_size), out=mk(B, self._voxels),
                      xs=mk(n_rec_x, B, self._voxels) if n_rec_x else None,
                      x0s=mk(n_rec_x0, B, self._voxels) if n_rec_x0 else None,
                      bank=mk(nbank, B, self._voxels) if nbank else None)
            self._io = {key: io}  # keep one set alive
        sd, cond, out, xs, x0s, bank = io["start"], io["cond"], io["out"], io["xs"], io["x0s"], io["bank"]
        sd.copy_(start.reshape(B, self._voxels), non_blocking=True)
        cond.copy_(self._cond(E, layers), non_blocking=True)
        if bank is not None:
            bank.copy_(noise_bank.reshape(nbank, B, self._voxels), non_blocking=True)
        arr = (_lib.Eval * len(evals))(*evals)
        h = self._engine(B)
        _lib.check(self._lib.cd_sample_run(
            h, sd.data_ptr(), float(start_scale), cond.data_ptr(), arr, len(evals),
            bank.data_ptr() if bank is not None else None, int(seed) & (2 ** 64 - 1), int(shower_offset), int(final_buf),
            out.data_ptr(), xs.data_ptr() if xs is not None else None, x0s.data_ptr() if x0s is not None else None,
            B, self._stream(), 1 if self.use_graph else 0), "sample_run")
        shp = tuple(start.shape)
        xs_l = [xs[i].reshape(shp).clone() for i in range(n_rec_x)] if xs is not None else None
        x0s_l = [x0s[i].reshape(shp).clone() for i in range(n_rec_x0)] if x0s is not None else None
        return out.reshape(shp).clone(), xs_l, x0s_l
