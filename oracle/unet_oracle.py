"""CPU oracle for the CaloDiffusion denoiser hot path (TEST INFRASTRUCTURE ONLY).

A functional, state-dict-driven restatement (plain torch CPU ops, fp32 or fp64) of the reference's
U-Net denoiser and the DDIM / DDPM / EDM-Euler / EDM-Heun sampling loops.  It exists to check the
CUDA path; nothing under `calodiffusion_b200/` imports it.  Only `tests/`, `__graft_entry__.smoke()`
and `bench.py`'s CPU-baseline legs may import this module.

Parity pin: `tests/golden/*.npz` hold outputs of the LIVE reference (`/root/reference`, imported in
the build container through `oracle/ref_shim.py` by `oracle/make_golden.py`) for seeded weights and
inputs; `tests/test_oracle_golden.py` checks this restatement against them, so the oracle is pinned
to the reference itself (the reference's own tests hold no numerical vectors except the
cylindrical-conv `[9,18,15]` known answer of `calodiffusion/tests/test_cyl_conv.py:7-25`, which is
also checked).

Each function cites the reference lines it restates (paths relative to /root/reference/calodiffusion).
"""
import math

import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------------
# geometry constants  (utils/utils.py:34-40 create_phi_image, :43-150 create_R_Z_image)
# --------------------------------------------------------------------------------------------
_R_BINS = {
    0: [0.0, 1.0, 4.0, 5.0, 7.0, 10.0, 15.0, 20.0, 30.0, 50.0, 80.0, 90.0, 100.0, 130.0, 150.0, 160.0,
        200.0, 250.0, 300.0, 350.0, 400.0, 600.0, 1000.0, 2000.0],
    1: [0.0, 2.0, 4.0, 5.0, 6.0, 8.0, 10.0, 12.0, 15.0, 20.0, 25.0, 30.0, 40.0, 50.0, 60.0, 70.0, 80.0,
        90.0, 100.0, 120.0, 130.0, 150.0, 160.0, 200.0, 250.0, 300.0, 350.0, 400.0, 600.0, 1000.0, 2000.0],
    2: [0, 4.65, 9.3, 13.95, 18.6, 23.25, 27.9, 32.55, 37.2, 41.85],
    3: [0, 2.325, 4.65, 6.975, 9.3, 11.625, 13.95, 16.275, 18.6, 20.925, 23.25, 25.575, 27.9, 30.225,
        32.55, 34.875, 37.2, 39.525, 41.85],
}


def r_bin_centres(dataset_num, n_r):
    """utils/utils.py:44-133: bin edges per dataset; >=100 (HGCal) uses arange(0, R+1)."""
    if dataset_num >= 100:
        edges = [float(i) for i in range(n_r + 1)]
    else:
        edges = _R_BINS[dataset_num]
    centres = [(edges[i] + edges[i + 1]) / 2.0 for i in range(len(edges) - 1)]
    if len(centres) != n_r:
        raise ValueError(f"Mismatch for dataset num {dataset_num}: expecting final dim {len(centres)}, got {n_r}")
    return centres


def geometry_channels(dataset_num, Z, P, R, dtype=torch.float32):
    """Returns (R_image, Z_image, phi_image), each (1,1,Z,P,R).  utils/utils.py:134-150 and :34-40."""
    centres = torch.tensor(r_bin_centres(dataset_num, R), dtype=torch.float32)
    r_img = (centres / centres[-1]).to(dtype).view(1, 1, 1, 1, R).expand(1, 1, Z, P, R).clone()
    z_img = (torch.arange(Z, dtype=torch.float32) / float(Z)).to(dtype).view(1, 1, Z, 1, 1).expand(1, 1, Z, P, R).clone()
    phi = torch.linspace(0.0, 1.0, P, dtype=torch.float32).to(dtype)
    phi_img = phi.view(1, 1, 1, P, 1).expand(1, 1, Z, P, R).clone()
    return r_img, z_img, phi_img


# --------------------------------------------------------------------------------------------
# layers  (models/models.py)
# --------------------------------------------------------------------------------------------
def _circ_pad_phi(x, n):
    """models/models.py:93-94: circular padding on dim -2 (phi) only."""
    if n == 0:
        return x
    return torch.cat([x[..., -n:, :], x, x[..., :n, :]], dim=-2)


def cylindrical_conv(x, w, b=None, stride=(1, 1, 1), padding=(0, 0, 0)):
    """models/models.py:65-96.  `padding` is padding_orig: phi part is circular, Z and R zero."""
    x = _circ_pad_phi(x, padding[1])
    return F.conv3d(x, w, b, stride=stride, padding=(padding[0], 0, padding[2]))


def cylindrical_conv_transpose(x, w, b, stride, output_padding):
    """models/models.py:25-62: circular pad 1 on phi, ConvTranspose3d(padding=(1, kP-1, 1))."""
    x = _circ_pad_phi(x, 1)
    return F.conv_transpose3d(x, w, b, stride=stride, padding=(1, w.shape[3] - 1, 1),
                              output_padding=output_padding)


def _block(sd, pre, x, groups):
    """models/models.py:160-169 Block.forward (scale_shift branch is never taken)."""
    x = cylindrical_conv(x, sd[pre + "proj.conv.weight"], sd[pre + "proj.conv.bias"], padding=(1, 1, 1))
    x = F.group_norm(x, groups, sd[pre + "norm.weight"], sd[pre + "norm.bias"], eps=1e-5)
    return F.silu(x)


def _resnet_block(sd, pre, x, cond, groups):
    """models/models.py:191-200 ResnetBlock.forward: additive conditioning between the blocks."""
    h = _block(sd, pre + "block1.", x, groups)
    if cond is not None and (pre + "mlp.1.weight") in sd:
        e = F.linear(F.silu(cond), sd[pre + "mlp.1.weight"], sd[pre + "mlp.1.bias"])
        h = h + e[:, :, None, None, None]
    h = _block(sd, pre + "block2.", h, groups)
    if (pre + "res_conv.conv.weight") in sd:
        x = cylindrical_conv(x, sd[pre + "res_conv.conv.weight"], sd[pre + "res_conv.conv.bias"])
    return h + x


def _linear_attention_residual(sd, pre, x):
    """models/models.py:111-117 Residual, :321-329 PreNorm, :301-318 LinearAttention (heads=1, d=32)."""
    b, c, Z, P, R = x.shape
    y = F.group_norm(x, 1, sd[pre + "fn.norm.weight"], sd[pre + "fn.norm.bias"], eps=1e-5)
    qkv = cylindrical_conv(y, sd[pre + "fn.fn.to_qkv.conv.weight"], None)
    d = qkv.shape[1] // 3
    q, k, v = [t.reshape(b, d, Z * P * R) for t in qkv.chunk(3, dim=1)]
    q = q.softmax(dim=-2) * (d ** -0.5)
    k = k.softmax(dim=-1)
    context = torch.einsum("bdn,ben->bde", k, v)
    out = torch.einsum("bde,bdn->ben", context, q).reshape(b, d, Z, P, R)
    out = cylindrical_conv(out, sd[pre + "fn.fn.to_out.0.conv.weight"], sd[pre + "fn.fn.to_out.0.conv.bias"])
    out = F.group_norm(out, 1, sd[pre + "fn.fn.to_out.1.weight"], sd[pre + "fn.fn.to_out.1.bias"], eps=1e-5)
    return out + x


def _mlp3(sd, pre, idx, x):
    """models/models.py:583-608: Linear → GELU(erf) → Linear → GELU → Linear."""
    h = F.gelu(F.linear(x, sd[f"{pre}{idx[0]}.weight"], sd[f"{pre}{idx[0]}.bias"]))
    h = F.gelu(F.linear(h, sd[f"{pre}{idx[1]}.weight"], sd[f"{pre}{idx[1]}.bias"]))
    return F.linear(h, sd[f"{pre}{idx[2]}.weight"], sd[f"{pre}{idx[2]}.bias"])


def cond_unet_forward(sd, x, cond, time, groups=8, pre="model."):
    """models/models.py:701-748 CondUnet.forward for the shipped topology
    (cylindrical, compress_Z, block_attn, mid_attn, time_embed=False, cond_embed=False)."""
    n_levels = 0
    while f"{pre}downs.{n_levels}.0.block1.proj.conv.weight" in sd:
        n_levels += 1
    x = cylindrical_conv(x, sd[pre + "init_conv.conv.weight"], sd[pre + "init_conv.conv.bias"], padding=(1, 1, 1))
    c = _mlp3(sd, pre + "cond_mlp.", (0, 2, 4), cond)
    t = _mlp3(sd, pre + "time_mlp.", (1, 3, 5), time.reshape(-1, 1))
    conditions = torch.cat([t, c], dim=-1)

    skips = []
    for i in range(n_levels):
        x = _resnet_block(sd, f"{pre}downs.{i}.0.", x, conditions, groups)
        x = _resnet_block(sd, f"{pre}downs.{i}.1.", x, conditions, groups)
        x = _linear_attention_residual(sd, f"{pre}downs_attn.{i}.", x)
        skips.append(x)
        if f"{pre}downs.{i}.2.conv.weight" in sd:
            x = cylindrical_conv(x, sd[f"{pre}downs.{i}.2.conv.weight"], sd[f"{pre}downs.{i}.2.conv.bias"],
                                 stride=(2, 2, 2), padding=(1, 1, 1))
    x = _resnet_block(sd, pre + "mid_block1.", x, conditions, groups)
    x = _linear_attention_residual(sd, pre + "mid_attn.", x)
    x = _resnet_block(sd, pre + "mid_block2.", x, conditions, groups)
    for i in range(n_levels):
        x = torch.cat((x, skips.pop()), dim=1)
        x = _resnet_block(sd, f"{pre}ups.{i}.0.", x, conditions, groups)
        x = _resnet_block(sd, f"{pre}ups.{i}.1.", x, conditions, groups)
        x = _linear_attention_residual(sd, f"{pre}ups_attn.{i}.", x)
        if f"{pre}ups.{i}.2.convTrans.weight" in sd:
            w = sd[f"{pre}ups.{i}.2.convTrans.weight"]
            tgt = skips[-1].shape[2:]
            # output_padding so the result matches the skip's shape (models/models.py:624-635,669-678)
            op = (tgt[0] - ((x.shape[2] - 1) * 2 + w.shape[2] - 2),
                  tgt[1] - (2 * x.shape[3]),
                  tgt[2] - ((x.shape[4] - 1) * 2 + w.shape[4] - 2))
            x = cylindrical_conv_transpose(x, w, sd[f"{pre}ups.{i}.2.convTrans.bias"], (2, 2, 2), op)
    x = _resnet_block(sd, pre + "final_conv.0.", x, None, groups)
    return cylindrical_conv(x, sd[pre + "final_conv.1.conv.weight"], sd[pre + "final_conv.1.conv.bias"])


# --------------------------------------------------------------------------------------------
# CaloDiffusion wrapper  (models/calodiffusion.py, models/loss.py)
# --------------------------------------------------------------------------------------------
class DenoiserOracle:
    """State-dict + config → `denoise(x, E, sigma, layers)`; mirrors CaloDiffusion.denoise."""

    def __init__(self, config, state_dict, dtype=torch.float32, device="cpu", geom=None):
        self.config = config
        self.geom = geom   # oracle.geom_embed.GeomEmbedOracle: in-loop enc/dec (calodiffusion.py:87-88, 95-96), else None
        self.dtype = dtype
        self.device = torch.device(device)   # "cuda": the same torch ops on cuDNN/cuBLAS (scripts/torch_gpu_baseline.py only)
        self.sd = {k: v.detach().to(self.device, dtype) for k, v in state_dict.items()}
        shape = config["SHAPE_FINAL"][1:]
        self.Z, self.P, self.R = shape[1], shape[2], shape[3]
        self.r_img, self.z_img, self.phi_img = (t.to(self.device) for t in
                                                geometry_channels(config.get("DATASET_NUM", 2), self.Z, self.P, self.R, dtype))
        self.layer_cond = "layer" in config.get("SHOWERMAP", "")
        self.time_embed = config.get("TIME_EMBED", "sin")
        # loss.py:16-25: sigma_data 1.0 for 'log' noise schedules, else 0.5
        self.sigma_data = 1.0 if "log" in config.get("NOISE_SCHED", "linear") else 0.5
        self.groups = config.get("BLOCK_GROUPS", 8)

    def time_map(self, sigma):
        """calodiffusion.py:144-155 (float64 detour, then .float() at :93)."""
        s = sigma.reshape(-1).to(torch.float32)
        if self.time_embed == "sigma":
            t = s / (1 + s ** 2).sqrt()
        elif self.time_embed == "log":
            t = 0.5 * torch.log(s)
        else:
            raise KeyError(self.time_embed)
        return t.to(torch.float64).to(torch.float32).to(self.dtype)

    def unet(self, x, E, t_emb, layers):
        """calodiffusion.py:86-98 forward + :121-142 add_RZPhi."""
        cond = E
        if self.layer_cond and layers is not None:
            cond = torch.cat([E, layers], dim=1)
        if self.geom is not None:
            x = self.geom.enc(x.to(torch.float32)).to(self.dtype)
        B = x.shape[0]
        cats = [x]
        if self.config.get("R_Z_INPUT", False):
            cats += [self.r_img.expand(B, -1, -1, -1, -1), self.z_img.expand(B, -1, -1, -1, -1)]
        if self.config.get("PHI_INPUT", False):
            cats += [self.phi_img.expand(B, -1, -1, -1, -1)]
        inp = torch.cat(cats, dim=1)
        out = cond_unet_forward(self.sd, inp, cond.to(self.dtype), t_emb, self.groups)
        return out if self.geom is None else self.geom.dec(out)

    def denoise(self, x, E, sigma, layers=None):
        """calodiffusion.py:154-169 with loss.py:29-41 scalings ('hybrid' branch)."""
        x = x.to(self.dtype)
        sigma = sigma.to(self.dtype).reshape([-1] + [1] * (x.dim() - 1))
        sd2 = self.sigma_data ** 2
        c_skip = sd2 / (sigma ** 2 + sd2)
        c_out = sigma * self.sigma_data / (sigma ** 2 + sd2) ** 0.5
        c_in = 1 / (sigma ** 2 + sd2) ** 0.5
        pred = self.unet(x * c_in, E.to(self.dtype), self.time_map(sigma), None if layers is None else layers.to(self.dtype))
        return c_skip * x + c_out * pred

    __call__ = denoise


# --------------------------------------------------------------------------------------------
# samplers  (models/sample.py, utils/sampling.py)
# --------------------------------------------------------------------------------------------
def cosine_alphas_cumprod(nsteps, s=0.008):
    """utils/sampling.py:16-24 + sample.py:45-47 (fp32, as the reference computes it)."""
    x = torch.linspace(0, nsteps, nsteps + 1)
    ac = torch.cos(((x / nsteps) + s) / (1 + s) * np.pi * 0.5) ** 2
    ac = ac / ac[0]
    betas = torch.clip(1 - (ac[1:] / ac[:-1]), 0.0001, 0.9999)
    return torch.cumprod(1.0 - betas, dim=0)


def ddpm_tables(num_steps, eta):
    """Per-step scalars of sample.py:41-104 in fp32: returns dict of (num_steps,) tensors indexed by t."""
    ac = cosine_alphas_cumprod(num_steps)
    ac_prev = F.pad(ac[:-1], (1, 0), value=1.0)
    sqrt_ac = torch.sqrt(ac)
    sqrt_1m = torch.sqrt(1.0 - ac)
    t = torch.arange(num_steps)
    denom = sqrt_ac[torch.clamp(t - 1, min=0)]
    sigma = sqrt_1m / sqrt_ac
    ddim_sigma = eta * (((1 - ac_prev) / (1 - ac)) * (1 - ac / ac_prev)) ** 0.5
    sigma_prev = (1.0 - ac_prev - ddim_sigma ** 2).sqrt() / denom
    return {"sigma": sigma, "sigma_prev": sigma_prev, "noise_scale": ddim_sigma / denom,
            "mask": (t > 0).to(torch.float32)}


def sample_ddim(model, start, E, layers, num_steps, eta=0.0, noise_bank=None, sample_offset=0):
    """sample.py:29-121 (DDim eta=0, DDPM eta=1).  `noise_bank[i]` replaces the i-th torch.randn draw.
    Returns (x, xs, x0s) like the reference (xs = post-update x)."""
    tab = ddpm_tables(num_steps, eta)
    steps = list(range(num_steps - 1, -1, -1))[sample_offset:]
    x = start.to(model.dtype) * tab["sigma"][steps[0]].to(model.dtype)
    xs, x0s = [], []
    B = x.shape[0]
    for i, t in enumerate(steps):
        sig = tab["sigma"][t].to(model.dtype)
        x0 = model(x, E, torch.full((B,), float(sig), dtype=model.dtype), layers)
        eps = (x - x0) / sig
        noise = noise_bank[i].to(model.dtype) if noise_bank is not None else torch.randn(x.shape, dtype=model.dtype)
        x = x0 + tab["mask"][t].to(model.dtype) * tab["sigma_prev"][t].to(model.dtype) * eps \
            + tab["noise_scale"][t].to(model.dtype) * noise
        x0s.append(x0)
        xs.append(x)
    return x, xs, x0s


def karras_steps(num_steps, sigma_min=0.002, sigma_max=80.0, rho=7, sample_offset=0):
    """sample.py:667-682: fp32 Karras discretisation with t_N = 0 appended."""
    i = torch.arange(num_steps, dtype=torch.float32)
    t = (sigma_max ** (1 / rho) + i / (num_steps - 1) * (sigma_min ** (1 / rho) - sigma_max ** (1 / rho))) ** rho
    return torch.cat([t, torch.zeros_like(t[:1])])[sample_offset:]


def sample_edm(model, start, E, layers, num_steps, order=1, noise_bank=None, last_step_euler=False,
               sigma_min=0.002, sigma_max=80.0, rho=7, S_churn=0.0, S_min=0.01, S_max=1.0, S_noise=1.003):
    """sample.py:631-662 for_loop with Euler (:784-789) or Heun (:805-821) in-loop update.

    Heun follows the reference literally: d' uses the PREVIOUS state x_cur (sample.py:820) and there is no
    last-step guard, so the final step divides by t_next = 0 (non-finite output).
    `last_step_euler=True` is the guarded variant (utils/sampling.py:232) applied to both sides in parity
    runs of the final showers."""
    t_steps = karras_steps(num_steps, sigma_min, sigma_max, rho)
    dt = model.dtype
    x_next = start.to(dt) * t_steps[0].to(dt)
    xs, x0s = [], []
    B = start.shape[0]
    for i, (t_cur, t_next) in enumerate(zip(t_steps[:-1], t_steps[1:])):
        x_cur = x_next
        gamma = min(S_churn / num_steps, math.sqrt(2) - 1) if S_min <= float(t_cur) <= S_max else 0
        t_hat = (t_cur + gamma * t_cur).to(dt)
        noise = noise_bank[i].to(dt) if noise_bank is not None else torch.randn(x_cur.shape, dtype=dt)
        x_hat = x_cur + (t_hat ** 2 - t_cur.to(dt) ** 2).sqrt() * S_noise * noise
        denoised = model(x_hat, E, torch.full((B,), float(t_hat), dtype=dt), layers)
        d_cur = (x_hat - denoised) / t_hat
        h = t_next.to(dt) - t_hat
        if order == 1 or (last_step_euler and float(t_next) == 0.0):
            x_next = x_hat + h * d_cur
        elif order == "dpm2":
            # sample.py:836-850 DPM2: midpoint in log-sigma; t_next = 0 gives t_mid = 0 and a division by zero
            t_mid = t_hat.log().lerp(t_next.to(dt).log(), 0.5).exp()
            x_2 = x_hat + d_cur * (t_mid - t_hat)
            den2 = model(x_2, E, torch.full((B,), float(t_mid), dtype=dt), layers)
            x_next = x_hat + h * ((x_2 - den2) / t_mid)
        else:
            x_prime = x_hat + h * d_cur
            t_prime = t_hat + h
            den2 = model(x_prime, E, torch.full((B,), float(t_prime), dtype=dt), layers)
            d_prime = (x_cur - den2) / t_next.to(dt)
            x_next = x_hat + h * (0.5 * d_cur + 0.5 * d_prime)
        xs.append(x_cur)
        x0s.append(denoised)
    return x_next, xs, x0s


def lms_coeff(order, t, i, j):
    """utils/sampling.py:77-89 `linear_multistep_coeff`: integral over [t_i, t_{i+1}] of the Lagrange basis polynomial
    of node t_{i-j} among t_i .. t_{i-order+1}.  The reference integrates with scipy quad (epsrel 1e-4); the integrand is
    a polynomial of degree order-1, so an `order`-point Gauss-Legendre rule is exact."""
    if order - 1 > i:
        raise ValueError(f"Order {order} too high for step {i}")
    t = np.asarray(t, dtype=np.float64)
    nodes, weights = np.polynomial.legendre.leggauss(max(order, 1))
    a, b = t[i], t[i + 1]
    tau = 0.5 * (b - a) * nodes + 0.5 * (b + a)
    prod = np.ones_like(tau)
    for k in range(order):
        if k == j:
            continue
        prod = prod * (tau - t[i - k]) / (t[i - j] - t[i - k])
    return float(0.5 * (b - a) * np.dot(weights, prod))


def sample_lms(model, start, E, layers, num_steps, order=4, sigma_min=0.002, sigma_max=80.0, rho=7):
    """sample.py:729-768 `LMS.sampler`: linear multistep over the Karras schedule, one evaluation per step, no noise,
    up to `order` stored derivatives d = (x - D)/t."""
    t_steps = karras_steps(num_steps, sigma_min, sigma_max, rho)
    dt = model.dtype
    B = start.shape[0]
    x_next = start.to(dt) * t_steps[0].to(dt)
    ds = []
    for i, t_cur in enumerate(t_steps[:-1]):
        x_hat = x_next
        denoised = model(x_hat, E, torch.full((B,), float(t_cur), dtype=dt), layers)
        ds.append((x_hat - denoised) / t_cur.to(dt))
        if len(ds) > order:
            ds.pop(0)
        cur_order = min(i + 1, order)
        coeffs = [lms_coeff(cur_order, t_steps.numpy(), i, j) for j in range(cur_order)]
        x_next = x_hat + sum(c * d for c, d in zip(coeffs, reversed(ds)))
    return x_next, [], []


def sample_consistency(model, start, E, layers, num_steps, consis_nsteps=100, noise_bank=None, sigma_min=0.002):
    """sample.py:969-1011 `Consistency.__call__` + utils/sampling.py:1143-1172 `sample_consis`: x0 prediction at up to
    five hard-coded noise levels of the `consis_nsteps` cosine schedule, re-noised to the next level in between.
    Returns (x, xs, last x0) like the reference."""
    ac = cosine_alphas_cumprod(consis_nsteps)
    s1m, sac = torch.sqrt(1.0 - ac), torch.sqrt(ac)
    t_all = [s1m[consis_nsteps - t - 1] / sac[consis_nsteps - t - 1] for t in range(consis_nsteps)]
    idxs = [0] + [int(round(consis_nsteps * f)) for f in (0.5, 0.7, 0.9, 0.95)]
    t_steps = torch.stack([t_all[i] for i in idxs[:num_steps]]) if num_steps > 1 else torch.stack([t_all[0]])
    sigmas = torch.cat([t_steps, torch.zeros_like(t_steps[:1])])
    dt = model.dtype
    B = start.shape[0]
    x = start.to(dt) * sigmas[0].to(dt)
    xs, x0, draws = [], None, 0
    for sigma_cur, sigma_next in zip(sigmas[:-1], sigmas[1:]):
        x0 = model(x, E, torch.full((B,), float(sigma_cur), dtype=dt), layers)
        sigma_next = torch.clip(sigma_next, sigma_min, None)
        if sigma_next > sigma_min:
            noise = noise_bank[draws].to(dt) if noise_bank is not None else torch.randn(x.shape, dtype=dt)
            draws += 1
            x = x0 + noise * torch.sqrt(sigma_next ** 2 - sigma_min ** 2).to(dt)
        else:
            x = x0
        xs.append(x)
    return x, xs, x0


# --------------------------------------------------------------------------------------------
# deterministic synthetic weights (shared recipe; the product never calls this)
# --------------------------------------------------------------------------------------------
def synth_state_dict(shapes, seed=0):
    """Seeded weights for a name→shape table: U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weights/biases,
    GroupNorm affine = 1 + 0.1 N(0,1) / 0.1 N(0,1) so the affine path is exercised."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, shape in shapes.items():
        shape = tuple(shape)
        if ".norm." in name or name.endswith("to_out.1.weight") or name.endswith("to_out.1.bias"):
            base = 1.0 if name.endswith("weight") else 0.0
            sd[name] = base + 0.1 * torch.randn(shape, generator=g)
        else:
            if len(shape) > 1:
                fan_in = int(np.prod(shape[1:]))
                if "convTrans" in name:
                    fan_in = int(shape[0] * np.prod(shape[2:]))
            else:
                fan_in = max(shape[0], 1)
            bound = 1.0 / math.sqrt(fan_in)
            sd[name] = (torch.rand(shape, generator=g) * 2 - 1) * bound
    return sd
