"""Samplers: DPM-Solver++(2M) (the reference's default, swarm/job_arguments.py:56-58,146-148),
Euler, Euler-ancestral (vid2vid, swarm/video/pix2pix.py:19) and DDIM, always with Karras sigmas
when asked (`use_karras_sigmas=True`, swarm/diffusion/diffusion_func.py:71-74).

B200-first formulation: every one of these update rules is *linear* in (x, x0, previous x0,
noise).  So the host precomputes, in fp64, one row of 8 coefficients per step

    x0    = e0*x + e1*model_out            (epsilon / v-prediction / sample parametrisation)
    x_new = c0*x + c1*x0 + c2*x0_prev + c3*noise
    next model input = x_new * s_in        (Euler's scale_model_input; 1 for DPM-Solver/DDIM)

and the whole scheduler step -- including classifier-free guidance -- is ONE fused kernel
(`sw_sched_step`, csrc/elementwise.cu) that indexes the table with a device-side step counter, so
a captured CUDA graph can be replayed for every step.  `apply_step_torch` evaluates the same row
with torch ops (CPU path and oracle).  Spec: SURVEY.md appendix A.1-A.5.

DPM-Solver++ uses the sigma-space form (alpha = 1/sqrt(sigma^2+1), lambda = -log sigma, final
sigma 0): identical to the table form away from t=0 and free of the duplicate-timestep 0/0 the
0.16-era rounded-Karras table form hits at 50 steps.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Optional

import numpy as np
import torch

SD_SCHEDULER_CONFIG = dict(num_train_timesteps=1000, beta_start=0.00085, beta_end=0.012,
                           beta_schedule="scaled_linear", prediction_type="epsilon", steps_offset=1)

KINDS = ("dpmpp_2m", "euler", "euler_a", "ddim")


def _betas(cfg) -> np.ndarray:
    n, b0, b1 = cfg["num_train_timesteps"], cfg["beta_start"], cfg["beta_end"]
    sched = cfg.get("beta_schedule", "scaled_linear")
    if sched == "scaled_linear":
        return np.linspace(b0 ** 0.5, b1 ** 0.5, n, dtype=np.float64) ** 2
    if sched == "linear":
        return np.linspace(b0, b1, n, dtype=np.float64)
    if sched == "squaredcos_cap_v2":
        f = lambda t: math.cos((t + 0.008) / 1.008 * math.pi / 2) ** 2
        return np.array([min(1 - f((i + 1) / n) / f(i / n), 0.999) for i in range(n)], dtype=np.float64)
    raise ValueError(f"unsupported beta_schedule {sched}")


def karras_sigmas(sigma_min: float, sigma_max: float, n: int, rho: float = 7.0) -> np.ndarray:
    ramp = np.linspace(0, 1, n)
    lo, hi = sigma_min ** (1 / rho), sigma_max ** (1 / rho)
    return (hi + ramp * (lo - hi)) ** rho


def sigma_to_t(sigma: np.ndarray, log_sigmas: np.ndarray) -> np.ndarray:
    """Fractional training timestep of each sigma by log-sigma interpolation."""
    log_sigma = np.log(np.maximum(sigma, 1e-10))
    dists = log_sigma[None, :] - log_sigmas[:, None]
    low = np.cumsum(dists >= 0, axis=0).argmax(axis=0).clip(max=log_sigmas.shape[0] - 2)
    high = low + 1
    w = np.clip((log_sigmas[low] - log_sigma) / (log_sigmas[low] - log_sigmas[high]), 0, 1)
    return (1 - w) * low + w * high


@dataclass
class StepPlan:
    """Everything a denoising loop needs, for one (scheduler, steps) choice."""

    kind: str
    timesteps: np.ndarray            # [n] model timesteps (float)
    sigmas: np.ndarray               # [n+1] k-diffusion sigmas, last = final
    coef: np.ndarray                 # [n, 8] fp32 {e0,e1,c0,c1,c2,c3,s_in,0}
    init_noise_sigma: float
    first_scale: float               # scale_model_input for the very first model call
    needs_noise: bool = False
    prediction_type: str = "epsilon"
    alphas_cumprod: np.ndarray = field(default=None, repr=False)

    @property
    def num_steps(self) -> int:
        return len(self.timesteps)

    def slice_from(self, start: int) -> "StepPlan":
        """img2img: run only steps [start, n). The multistep history restarts at first order."""
        coef = self.coef[start:].copy()
        if self.kind == "dpmpp_2m" and len(coef):
            coef[0] = _dpm_row(self.sigmas[start], self.sigmas[start + 1], None, self.prediction_type,
                               second_order=False)
        first = 1.0 / math.sqrt(self.sigmas[start] ** 2 + 1) if self.kind in ("euler", "euler_a") else 1.0
        return StepPlan(self.kind, self.timesteps[start:], self.sigmas[start:], coef, self.init_noise_sigma, first,
                        self.needs_noise, self.prediction_type, self.alphas_cumprod)

    def add_noise(self, x0: torch.Tensor, noise: torch.Tensor, step: int = 0) -> torch.Tensor:
        """Noised sample at the start of `step` (img2img / inpaint starting point)."""
        s = float(self.sigmas[step])
        if self.kind in ("euler", "euler_a"):
            return x0 + noise * s
        a = 1.0 / math.sqrt(s * s + 1.0)
        return a * x0 + (s * a) * noise


def _x0_coefs(sigma: float, prediction: str, sigma_space: bool):
    """x0 = e0*x + e1*model_out, for a sample x expressed either in sigma space (x = x0 + sigma*eps,
    Euler) or in variance-preserving space (x = alpha*x0 + sigma_t*eps, DPM-Solver / DDIM)."""
    a = 1.0 / math.sqrt(sigma * sigma + 1.0)
    st = sigma * a
    if sigma_space:
        if prediction == "epsilon":
            return 1.0, -sigma
        if prediction == "v_prediction":
            return 1.0 / (sigma * sigma + 1.0), -sigma / math.sqrt(sigma * sigma + 1.0)
        return 0.0, 1.0
    if prediction == "epsilon":
        return 1.0 / a, -st / a
    if prediction == "v_prediction":
        return a, -st
    return 0.0, 1.0


def _dpm_row(s_cur, s_next, s_prev, prediction, second_order):
    e0, e1 = _x0_coefs(s_cur, prediction, sigma_space=False)
    a_s, a_t = 1 / math.sqrt(s_cur ** 2 + 1), 1 / math.sqrt(s_next ** 2 + 1)
    sig_s, sig_t = s_cur * a_s, s_next * a_t
    if s_next == 0.0:  # final step: x_t = x0
        return [e0, e1, 0.0, 1.0, 0.0, 0.0, 1.0, 0.0]
    lam_s, lam_t = -math.log(s_cur), -math.log(s_next)
    h = lam_t - lam_s
    A = sig_t / sig_s
    B = -a_t * math.expm1(-h)
    if second_order and s_prev is not None:
        h0 = lam_s - (-math.log(s_prev))
        r0 = h0 / h
        return [e0, e1, A, B * (1 + 0.5 / r0), -B * 0.5 / r0, 0.0, 1.0, 0.0]
    return [e0, e1, A, B, 0.0, 0.0, 1.0, 0.0]


def make_plan(kind: str, num_steps: int, config: Optional[dict] = None, use_karras_sigmas: bool = True) -> StepPlan:
    cfg = dict(SD_SCHEDULER_CONFIG)
    cfg.update(config or {})
    if kind not in KINDS:
        raise ValueError(f"unknown scheduler kind {kind}")
    if num_steps < 1:
        raise ValueError("num_inference_steps must be >= 1")
    prediction = cfg.get("prediction_type", "epsilon")
    n_train = cfg["num_train_timesteps"]
    acp = np.cumprod(1.0 - _betas(cfg))
    table = np.sqrt((1 - acp) / acp)  # k-diffusion sigma per training timestep (ascending)
    log_table = np.log(table)

    if kind == "ddim":
        ratio = n_train // num_steps
        ts = (np.arange(num_steps) * ratio).round()[::-1].astype(np.int64) + cfg.get("steps_offset", 0)
        ts = ts.clip(max=n_train - 1)
        coef = []
        eta = float(cfg.get("eta", 0.0) or 0.0)  # DDIM's stochasticity knob: 0 = deterministic, 1 = DDPM-like variance
        for t in ts:
            prev = t - ratio
            a_t = acp[t]
            a_p = acp[prev] if prev >= 0 else (1.0 if cfg.get("set_alpha_to_one", False) else acp[0])
            sig = math.sqrt((1 - a_t) / a_t)
            e0, e1 = _x0_coefs(sig, prediction, sigma_space=False)
            # x_prev = sqrt(a_p) x0 + sqrt(1 - a_p - s^2) eps + s z, eps = (x - sqrt(a_t) x0) / sqrt(1 - a_t),
            # s = eta sqrt((1 - a_p) / (1 - a_t)) sqrt(1 - a_t / a_p)
            s_noise = eta * math.sqrt(max((1 - a_p) / (1 - a_t) * (1 - a_t / a_p), 0.0))
            c0 = math.sqrt(max(1 - a_p - s_noise * s_noise, 0.0)) / math.sqrt(1 - a_t)
            coef.append([e0, e1, c0, math.sqrt(a_p) - c0 * math.sqrt(a_t), 0.0, s_noise, 1.0, 0.0])
        sig = np.concatenate([table[ts], [0.0]])
        return StepPlan(kind, ts.astype(np.float64), sig, np.asarray(coef, np.float32), 1.0, 1.0, eta > 0.0, prediction, acp)

    if use_karras_sigmas:
        sig = karras_sigmas(float(table[0]), float(table[-1]), num_steps)
        ts = sigma_to_t(sig, log_table)
    else:
        if kind == "dpmpp_2m":
            ts = np.linspace(0, n_train - 1, num_steps + 1).round()[::-1][:-1].astype(np.float64)
        else:
            ts = np.linspace(0, n_train - 1, num_steps, dtype=np.float64)[::-1].copy()
        sig = np.interp(ts, np.arange(n_train), table)
    if kind == "dpmpp_2m":
        ts = ts.round()
    sig = np.concatenate([sig, [0.0]])

    coef = []
    if kind == "dpmpp_2m":
        lower_order_final = num_steps < 15
        for i in range(num_steps):
            last = i == num_steps - 1
            second = i > 0 and not (last and lower_order_final) and sig[i + 1] > 0
            coef.append(_dpm_row(sig[i], sig[i + 1], sig[i - 1] if i > 0 else None, prediction, second))
        return StepPlan(kind, ts, sig, np.asarray(coef, np.float32), 1.0, 1.0, False, prediction, acp)

    # Euler family (sigma space)
    for i in range(num_steps):
        s, sn = sig[i], sig[i + 1]
        e0, e1 = _x0_coefs(s, prediction, sigma_space=True)
        s_in_next = 1.0 / math.sqrt(sn * sn + 1.0)
        if kind == "euler":
            r = sn / s
            coef.append([e0, e1, r, 1 - r, 0.0, 0.0, s_in_next, 0.0])
        else:
            up = math.sqrt(max(sn * sn * (s * s - sn * sn) / (s * s), 0.0))
            down = math.sqrt(max(sn * sn - up * up, 0.0))
            r = down / s
            coef.append([e0, e1, r, 1 - r, 0.0, up, s_in_next, 0.0])
    first = 1.0 / math.sqrt(float(sig[0]) ** 2 + 1)
    # diffusers' Euler schedulers start from randn * sigma_max, except with "leading" timestep spacing (what the SDXL
    # checkpoints configure), where it is randn * sqrt(sigma_max^2 + 1)
    leading = cfg.get("timestep_spacing", "linspace") == "leading"
    init_sigma = math.sqrt(float(sig[0]) ** 2 + 1) if leading else float(sig[0])
    return StepPlan(kind, ts, sig, np.asarray(coef, np.float32), init_sigma, first, kind == "euler_a", prediction, acp)


def guided(model_out: torch.Tensor, n_cond: int, g_text: float, g_image: float = 0.0) -> torch.Tensor:
    """Classifier-free guidance over a stacked batch (appendix A.5)."""
    if n_cond == 1:
        return model_out
    parts = model_out.chunk(n_cond)
    if n_cond == 2:
        eu, ec = parts
        return eu + g_text * (ec - eu)
    et, em, eu = parts
    return eu + g_text * (et - em) + g_image * (em - eu)


def apply_step_torch(plan: StepPlan, i: int, x: torch.Tensor, model_out: torch.Tensor, x0_prev: Optional[torch.Tensor],
                     noise: Optional[torch.Tensor] = None):
    """Reference evaluation of one table row with torch ops. Returns (x_new, x0, next_model_input_scale)."""
    e0, e1, c0, c1, c2, c3, s_in, _ = (float(v) for v in plan.coef[i])
    xf = x.float()
    x0 = e0 * xf + e1 * model_out.float()
    xn = c0 * xf + c1 * x0
    if c2 != 0.0 and x0_prev is not None:
        xn = xn + c2 * x0_prev
    if c3 != 0.0 and noise is not None:
        xn = xn + c3 * noise.float()
    return xn, x0, s_in


# ---------------------------------------------------------------- fp64 oracle (tests)
def reference_dpmpp_2m(plan: StepPlan, x, eps_fn):
    """Textbook DPM-Solver++(2M) in fp64 written from the formulas, independent of
    the coefficient table. `eps_fn(x, i)` returns the model output at step i."""
    x = np.asarray(x, dtype=np.float64)
    sig = plan.sigmas
    prev_x0 = None
    n = plan.num_steps
    for i in range(n):
        s, t = sig[i], sig[i + 1]
        a_s = 1 / math.sqrt(s * s + 1)
        out = np.asarray(eps_fn(x, i), dtype=np.float64)
        if plan.prediction_type == "epsilon":
            x0 = (x - s * a_s * out) / a_s
        elif plan.prediction_type == "v_prediction":
            x0 = a_s * x - s * a_s * out
        else:
            x0 = out
        if t == 0:
            x, prev_x0 = x0, x0
            continue
        a_t = 1 / math.sqrt(t * t + 1)
        h = -math.log(t) + math.log(s)
        first = i == 0 or (i == n - 1 and n < 15)
        if first:
            x = (t * a_t) / (s * a_s) * x - a_t * math.expm1(-h) * x0
        else:
            h0 = -math.log(s) + math.log(sig[i - 1])
            r0 = h0 / h
            d1 = (x0 - prev_x0) / r0
            x = (t * a_t) / (s * a_s) * x - a_t * math.expm1(-h) * x0 - 0.5 * a_t * math.expm1(-h) * d1
        prev_x0 = x0
    return x
