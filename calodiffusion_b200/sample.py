"""Sampler registry mirror (models/sample.py).  Same class names, constructor `(config)`, `SAMPLER_OPTIONS`
keys and `__call__(model, start, energy, layers, num_steps, sample_offset, debug) -> (x, xs, x0s)`.

Each sampler turns its schedule into a table of `cd_eval` records (include/calodiffusion_b200.h) — one per
denoiser evaluation, carrying the coefficients of the state update that the CUDA library fuses into the
final-conv epilogue — and hands the whole N-step loop to `model.run_program`, which replays it as one
CUDA graph.  Schedules are computed with the same fp32 torch expressions as the reference.

Noise: the reference draws `torch.randn` every step.  Here the per-step noise comes from the in-kernel
Philox generator (seeded from torch's global RNG, keyed by global shower index so results do not depend
on the batch sharding) or, when `sampler.noise_bank` is set (parity tests), from that tensor
(slot i = the reference's i-th draw).
"""
import math

import numpy as np
import torch

from . import _lib
from .loss import cosine_beta_schedule


def _eval(**kw):
    e = _lib.Eval()
    e.x_out = e.x0_buf = e.a1 = e.a2 = e.a3 = -1
    e.rec_x0 = e.rec_x = -1
    e.pre_src = e.d_out = -1
    e.pre_a = 1.0
    for k, v in kw.items():
        if k == "k":
            for i, c in enumerate(v):
                e.k[i] = float(c)
        else:
            setattr(e, k, v)
    return e


class Sample:
    def __init__(self, config):
        self.config = config
        self.sample_config = self.config.get("SAMPLER_OPTIONS", {})
        self.noise_bank = None      # optional (n_draws, B, ...) tensor replacing the RNG (parity tests)
        self.shower_offset = 0      # global index of the first shower of this batch (multi-GPU sharding)

    def _seed(self):
        return int(torch.randint(0, 2 ** 62, (1,)).item())

    def __call__(self, model, start, energy, layers, num_steps, sample_offset, debug):
        raise NotImplementedError


class DDim(Sample):
    """models/sample.py:29-109 (eta = 0)."""

    def __init__(self, config):
        super().__init__(config)
        self.ddim_eta = 0.0

    @torch.no_grad()
    def __call__(self, model, start, energy, layers, num_steps, sample_offset, debug):
        betas = cosine_beta_schedule(num_steps)
        alphas_cumprod = torch.cumprod(1.0 - betas, dim=0)
        alphas_cumprod_prev = torch.nn.functional.pad(alphas_cumprod[:-1], (1, 0), value=1.0)
        sqrt_ac = torch.sqrt(alphas_cumprod)
        sqrt_1m = torch.sqrt(1.0 - alphas_cumprod)
        time_steps = torch.flip(torch.arange(num_steps), [0])
        if sample_offset > 0:
            time_steps = time_steps[sample_offset:]
        sigma_start = float(sqrt_1m[time_steps[0]] / sqrt_ac[time_steps[0]])
        evals = []
        for i, t in enumerate(time_steps.tolist()):
            alpha, alpha_prev = alphas_cumprod[t], alphas_cumprod_prev[t]
            denom = sqrt_ac[max(t - 1, 0)]
            sigma = sqrt_1m[t] / sqrt_ac[t]
            ddim_sigma = self.ddim_eta * (((1 - alpha_prev) / (1 - alpha)) * (1 - alpha / alpha_prev)) ** 0.5
            sigma_prev = (1.0 - alpha_prev - ddim_sigma ** 2).sqrt() / denom
            a = float(sigma_prev / sigma) if t > 0 else 0.0          # mask * sigma_prev / sigma
            ns = float(ddim_sigma / denom)
            # x = x0 + a*(x - x0) + ns*noise
            evals.append(_eval(sigma=float(sigma), x_in=0, x_out=0, k=(1.0 - a, a, ns, 0, 0), a1=-2, noise_index=i,
                               rec_x0=i if debug else -1, rec_x=i if debug else -1, rec_x_buf=0))
        n = len(evals)
        x, xs, x0s = model.run_program(start, energy, layers, evals, sigma_start, 0, noise_bank=self.noise_bank,
                                       seed=self._seed(), shower_offset=self.shower_offset,
                                       n_rec_x=n if debug else 0, n_rec_x0=n if debug else 0)
        return x, (xs if debug else []), (x0s if debug else [])


class DDPM(DDim):
    """models/sample.py:112-121 (eta = 1)."""

    def __init__(self, config):
        super().__init__(config)
        self.ddim_eta = 1.0


class EDMAbstract(Sample):
    """models/sample.py:577-726: Karras schedule + churn loop; subclasses pick the in-loop update."""
    order = 1

    def __init__(self, config):
        super().__init__(config)
        noisy = self.config.get("NOISY_SAMPLE", False)
        self.S_churn = 40 if noisy else 0
        self.S_min = self.sample_config.get("S_MIN", 0.01)
        self.S_max = 50 if noisy else 1
        self.S_noise = self.sample_config.get("S_NOISE", 1.003)
        self.sigma_min = self.sample_config.get("SIGMA_MIN", 0.002)
        self.sigma_max = self.sample_config.get("SIGMA_MAX", 80.0)
        self.orig_schedule = self.sample_config.get("ORG_SCHEDULE", False)
        self.rho = self.sample_config.get("RHO", 7)
        # Heun / DPM2: the reference's last step divides by t_next = 0 (non-finite showers, reproduced by default).
        # SAMPLER_OPTIONS {"LAST_STEP_EULER": true} (or the attribute) takes that step as an Euler step instead, the
        # guard of utils/sampling.py:232 - the finite form bench.py measures and says so in its config block.
        self.last_step_euler = bool(self.sample_config.get("LAST_STEP_EULER", False))

    def setup(self, num_steps, sample_offset):
        if self.orig_schedule:
            # models/sample.py:686-699 calls self.alpha_bar(j - 1) with one argument although alpha_bar takes (j, num_steps):
            # the reference raises TypeError here for every sampler, so there is no behaviour to reproduce
            raise ValueError("ORG_SCHEDULE raises inside the reference itself (alpha_bar arity, models/sample.py:691); "
                             "not supported by the B200 backend")
        step_indices = torch.arange(num_steps, dtype=torch.float32)
        t_steps = (self.sigma_max ** (1 / self.rho) + step_indices / (num_steps - 1)
                   * (self.sigma_min ** (1 / self.rho) - self.sigma_max ** (1 / self.rho))) ** self.rho
        t_steps = torch.cat([t_steps, torch.zeros_like(t_steps[:1])])
        return t_steps[sample_offset:]

    @torch.no_grad()
    def __call__(self, model, start, energy, layers, num_steps, sample_offset, debug):
        t_steps = self.setup(num_steps, sample_offset)
        evals = []
        cur, other = 0, 1     # state buffers: x_cur lives in `cur`
        n_steps = len(t_steps) - 1
        for i in range(n_steps):
            t_cur, t_next = t_steps[i], t_steps[i + 1]
            gamma = min(self.S_churn / num_steps, np.sqrt(2) - 1) if self.S_min <= t_cur <= self.S_max else 0
            t_hat = torch.as_tensor(t_cur + gamma * t_cur)
            pre_b = float((t_hat ** 2 - t_cur ** 2).sqrt() * self.S_noise)
            h = t_next - t_hat
            ht = float(h / t_hat)
            common = dict(sigma=float(t_hat), x_in=cur, pre_a=1.0, pre_b=pre_b, pre_noise_index=i,
                          rec_x0=i if debug else -1)
            # with churn (pre_b != 0) x_hat = x_cur + noise replaces x_cur in place; the engine records xs[i] before the noise
            # is added, as the reference does (models/sample.py:660 appends x_cur)
            if self.order == 1 or (self.last_step_euler and float(t_next) == 0.0):
                # x_next = x_hat + h*(x_hat - D)/t_hat
                evals.append(_eval(**common, x_out=other, k=(-ht, 1.0 + ht, 0, 0, 0),
                                   rec_x=i if debug else -1, rec_x_buf=cur))
                cur, other = other, cur
            elif self.order == "dpm2":
                # sample.py:836-850: t_mid = sqrt(t_hat*t_next) (lerp in log sigma); x_2 = x_hat + d_cur*(t_mid - t_hat) -> buf3;
                # x_next = x_hat + h*(x_2 - D_2)/t_mid.  t_next = 0: t_mid = 0, evaluation at sigma 0 and h/0 (reference)
                t_mid = torch.as_tensor(t_hat).log().lerp(torch.as_tensor(t_next).log(), 0.5).exp()
                r = float((t_mid - t_hat) / t_hat)
                # (xs[i] is recorded by the FIRST evaluation: with churn the engine takes it before the noise is added)
                evals.append(_eval(**common, x_out=3, k=(-r, 1.0 + r, 0, 0, 0), rec_x=i if debug else -1, rec_x_buf=cur))
                with np.errstate(divide="ignore"):
                    hm = float(np.float32(h) / np.float32(t_mid))
                sig2 = float(t_mid) if float(t_mid) > 0.0 else 1e-30
                evals.append(_eval(sigma=sig2, x_in=3, x_out=other, k=(-hm, hm, 1.0, 0, 0), a1=cur))
                cur, other = other, cur
            else:
                # first evaluation: D -> buf2, x' -> buf3
                evals.append(_eval(**common, x_out=3, x0_buf=2, k=(-ht, 1.0 + ht, 0, 0, 0)))
                # second: D' at t' = t_hat + h ; x_next = x_hat + h*(0.5*(x_hat-D)/t_hat + 0.5*(x_cur-D')/t_next)
                t_prime = t_hat + h
                with np.errstate(divide="ignore"):
                    hn = float(np.float32(0.5) * np.float32(h) / np.float32(t_next))  # inf at the last step (reference)
                k = (-hn, 0.0, 1.0 + 0.5 * ht + hn, -0.5 * ht, 0)
                sig2 = float(t_prime)
                if not sig2 > 0.0:
                    # reference evaluates the network at sigma = 0 here (log 0 / division by zero): results are
                    # non-finite there as well; evaluate at a tiny sigma and let the inf coefficients propagate
                    sig2 = 1e-30
                if pre_b != 0.0:
                    # churn: d' uses the UN-noised x_cur (models/sample.py:820 reads self.x_next) while everything else uses
                    # x_hat, so x_hat goes to its own buffer (`other`) and x_cur stays in `cur`:
                    #   x_next = x_hat + h*(0.5*(x_hat - D)/t_hat + 0.5*(x_cur - D')/t_next)
                    evals[-1].x_in, evals[-1].pre_src = other, cur
                    k = (-hn, 0.0, hn, -0.5 * ht, 1.0 + 0.5 * ht)
                    evals.append(_eval(sigma=sig2, x_in=3, x_out=other, k=k, a1=cur, a2=2, a3=other,
                                       rec_x=i if debug else -1, rec_x_buf=cur))
                else:
                    evals.append(_eval(sigma=sig2, x_in=3, x_out=other, k=k, a1=cur, a2=2,
                                       rec_x=i if debug else -1, rec_x_buf=cur))
                cur, other = other, cur
        n = n_steps
        x, xs, x0s = model.run_program(start, energy, layers, evals, float(t_steps[0]), cur, noise_bank=self.noise_bank,
                                       seed=self._seed(), shower_offset=self.shower_offset,
                                       n_rec_x=n if debug else 0, n_rec_x0=n if debug else 0)
        return x, (xs if debug else []), (x0s if debug else [])


class Euler(EDMAbstract):
    """models/sample.py:772-789."""
    order = 1


class Heun(EDMAbstract):
    """models/sample.py:792-821 — including its use of x_cur in d' (:820) and the unguarded last step."""
    order = 2


def linear_multistep_coeff(order, t, i, j):
    """utils/sampling.py:77-89: integral over [t_i, t_{i+1}] of the Lagrange basis polynomial of node t_{i-j} among
    t_i .. t_{i-order+1}.  The reference calls scipy quad on this degree-(order-1) polynomial; an `order`-point
    Gauss-Legendre rule integrates it exactly."""
    if order - 1 > i:
        raise ValueError(f"Order {order} too high for step {i}")
    t = np.asarray(t, dtype=np.float64)
    nodes, weights = np.polynomial.legendre.leggauss(max(int(order), 1))
    half, mid = 0.5 * (t[i + 1] - t[i]), 0.5 * (t[i + 1] + t[i])
    tau = half * nodes + mid
    prod = np.ones_like(tau)
    for k in range(order):
        if k != j:
            prod *= (tau - t[i - k]) / (t[i - j] - t[i - k])
    return float(half * np.dot(weights, prod))


class LMS(EDMAbstract):
    """models/sample.py:729-768: linear multistep (SAMPLER_OPTIONS ORDER, default 4) over the Karras schedule; one
    denoiser evaluation per step, no noise, returns empty xs / x0s like the reference.

    Runs as a fused `cd_eval` program like the other samplers (one CUDA graph): every evaluation also stores its ODE
    derivative d_i = (x - D)/t_i in a ring of state buffers (cd_eval.d_out / d_scale) and the update reads the previous
    ORDER-1 of them:  x_next = x + c_0*d_i + c_1*d_{i-1} + ...  with c_0*d_i folded into the (D, x) coefficients."""
    MAX_ORDER = 4   # x + the current derivative + three ring reads (a1..a3) per record

    def __init__(self, config):
        super().__init__(config)
        self.order = self.sample_config.get("ORDER", 4)

    @torch.no_grad()
    def __call__(self, model, start, energy, layers, num_steps, sample_offset, debug):
        if not 1 <= int(self.order) <= self.MAX_ORDER:
            raise ValueError(f"LMS ORDER must be in [1, {self.MAX_ORDER}] on the B200 backend (got {self.order})")
        t_steps = self.setup(num_steps, sample_offset)
        t64 = t_steps.double().numpy()
        ring = (1, 2, 3, 4)                     # state buffers holding d_i, slot i % 4
        evals = []
        for i in range(len(t_steps) - 1):
            t_hat = float(t_steps[i])
            cur_order = min(i + 1, int(self.order))
            c = [linear_multistep_coeff(cur_order, t64, i, j) for j in range(cur_order)] + [0.0] * 4
            prev = [ring[(i - j) % 4] if j < cur_order else -1 for j in (1, 2, 3)]
            evals.append(_eval(sigma=t_hat, x_in=0, x_out=0, k=(-c[0] / t_hat, 1.0 + c[0] / t_hat, c[1], c[2], c[3]),
                               a1=prev[0], a2=prev[1], a3=prev[2], d_out=ring[i % 4], d_scale=1.0 / t_hat))
        x, _, _ = model.run_program(start, energy, layers, evals, float(t_steps[0]), 0, noise_bank=None, seed=0,
                                    shower_offset=self.shower_offset)
        return x, [], []


class Restart(EDMAbstract):
    """models/sample.py:853-955.  The restart branch is guarded by `index + 1 in self.restart_list.keys()` (:877) with an
    int on the left; the default RESTART_LIST (and any list read from config.json) has string keys, so the reference never
    restarts and every step is the Euler update x_hat + h*d_cur (:872-875) - that behaviour is what runs here, as the same
    fused cd_eval program as `Euler`.  A RESTART_LIST whose int keys would fire the restart loop raises `ValueError`.
    (The reference fills `xs` with the builtin `next` (:951); the per-step states are returned here instead.)"""
    order = 1

    def __init__(self, config):
        super().__init__(config)
        default_restart = {"0": [4, 1, 19.35, 40.79], "1": [4, 1, 1.09, 1.92], "2": [4, 4, 0.59, 1.09],
                           "3": [4, 1, 0.30, 0.59], "4": [4, 4, 0.06, 0.30]}
        self.restart_list = self.sample_config.get("RESTART_LIST", default_restart)
        self.restart_gamma = self.sample_config.get("RESTART_GAMMA", 0.05)

    @torch.no_grad()
    def __call__(self, model, start, energy, layers, num_steps, sample_offset, debug):
        n_steps = num_steps + 1 - sample_offset - 1
        if any(isinstance(k, int) and not isinstance(k, bool) and 1 <= k <= n_steps for k in self.restart_list.keys()):
            raise ValueError("Restart with an int-keyed RESTART_LIST (active restart loops) is not supported by the "
                             "B200 backend")
        return super().__call__(model, start, energy, layers, num_steps, sample_offset, debug)


class DPM2(EDMAbstract):
    """models/sample.py:824-850: midpoint (in log sigma) second-order step, two evaluations per step - including the
    unguarded last step (t_mid = 0: the reference's final showers are non-finite; `last_step_euler` is the guarded form)."""
    order = "dpm2"


class Consistency(Sample):
    """models/sample.py:957-1011 + utils/sampling.py:1143-1172 (`sample_consis`): x0 prediction at up to five hard-coded
    levels of the CONSIS_NSTEPS cosine schedule, re-noised to the next level (x = x0 + sqrt(s_next^2 - s_min^2)*noise) in
    between.  Returns `(x, xs, last x0)` like the reference (the third item is a tensor there, not a list)."""
    sigma_min = 0.002

    def __init__(self, config):
        super().__init__(config)
        self.consis_nsteps = self.config.get("CONSIS_NSTEPS", 100)

    @torch.no_grad()
    def __call__(self, model, start, energy, layers, num_steps, sample_offset, debug):
        n = self.consis_nsteps
        alphas_cumprod = torch.cumprod(1.0 - cosine_beta_schedule(n), dim=0)
        sqrt_ac, sqrt_1m = torch.sqrt(alphas_cumprod), torch.sqrt(1.0 - alphas_cumprod)
        t_all = [sqrt_1m[n - t - 1] / sqrt_ac[n - t - 1] for t in range(n)]
        sample_idxs = [0] + [int(round(n * f)) for f in (0.5, 0.7, 0.9, 0.95)]
        t_steps = torch.stack([t_all[i] for i in sample_idxs[:num_steps]]) if num_steps > 1 else torch.stack([t_all[0]])
        sigmas = torch.cat([t_steps, torch.zeros_like(t_steps[:1])])
        evals, draws = [], 0
        last = len(sigmas) - 2
        for i in range(len(sigmas) - 1):
            sigma_next = torch.clip(sigmas[i + 1], self.sigma_min, None)
            rec = dict(rec_x0=0 if i == last else -1, rec_x=i if debug else -1, rec_x_buf=0)
            if sigma_next > self.sigma_min:
                ns = float(torch.sqrt(sigma_next ** 2 - self.sigma_min ** 2))
                evals.append(_eval(sigma=float(sigmas[i]), x_in=0, x_out=0, k=(1.0, 0, ns, 0, 0), a1=-2,
                                   noise_index=draws, **rec))
                draws += 1
            else:
                evals.append(_eval(sigma=float(sigmas[i]), x_in=0, x_out=0, k=(1.0, 0, 0, 0, 0), **rec))
        x, xs, x0s = model.run_program(start, energy, layers, evals, float(sigmas[0]), 0, noise_bank=self.noise_bank,
                                       seed=self._seed(), shower_offset=self.shower_offset,
                                       n_rec_x=len(evals) if debug else 0, n_rec_x0=1)
        return x, (xs if debug else []), x0s[0]
