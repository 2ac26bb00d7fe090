"""Noise schedulers of the slider path (DDIM / DDPM / Euler-ancestral / LMS) as affine latent updates.

The reference builds diffusers==0.20.2 schedulers (conceptmod/textsliders/model_util.py:388-436) and
calls `.set_timesteps / .scale_model_input / .step / .add_noise / .init_noise_sigma / .timesteps`
(train_util.py:103,287,324; train_lora.py:157,209-213).  diffusers is not available offline, so the
published update rules are restated here; DDIM is cross-checked against the
reference's own restatement `prev_step` (trainscripts/imagesliders/train_lora-scale.py:31-39).

Every rule is   x' = c_x * x + c_eps * model_output + c_noise * z   with host-computed fp64
coefficients, so one launch of the fused CFG+step kernel (ops.cfg_step) performs it.
"""
from __future__ import annotations

from types import SimpleNamespace

import numpy as np
import torch

from . import ops


def sd_alphas_cumprod(beta_start=0.00085, beta_end=0.012, n=1000) -> torch.Tensor:
    """`scaled_linear` schedule, computed in float32 like diffusers does."""
    betas = torch.linspace(beta_start ** 0.5, beta_end ** 0.5, n, dtype=torch.float32) ** 2
    return torch.cumprod(1.0 - betas, dim=0)


class _StepOutput(SimpleNamespace):
    def __getitem__(self, i):          # `scheduler.step(..., return_dict=False)[0]` (train_util.py:445)
        return (self.prev_sample,)[i]


class SchedulerBase:
    order = 1

    def __init__(self, beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear",
                 num_train_timesteps=1000, clip_sample=False, prediction_type="epsilon"):
        if beta_schedule != "scaled_linear":
            raise NotImplementedError("only the SD `scaled_linear` schedule is used by the reference")
        if clip_sample:
            raise NotImplementedError("clip_sample=True is never used by the reference (model_util.py:401,411)")
        self.config = SimpleNamespace(num_train_timesteps=num_train_timesteps, beta_start=beta_start,
                                      beta_end=beta_end, beta_schedule=beta_schedule, clip_sample=clip_sample,
                                      prediction_type=prediction_type)
        self.alphas_cumprod = sd_alphas_cumprod(beta_start, beta_end, num_train_timesteps)
        self._ac = self.alphas_cumprod.double().numpy()
        self.final_alpha_cumprod = torch.tensor(1.0)
        self.num_inference_steps = None
        self.timesteps = torch.arange(num_train_timesteps - 1, -1, -1, dtype=torch.int64)
        self.init_noise_sigma = torch.tensor(1.0)

    # -- diffusers-compatible surface ---------------------------------------------------------
    def set_timesteps(self, num_inference_steps: int, device=None):
        self.num_inference_steps = num_inference_steps
        ratio = self.config.num_train_timesteps // num_inference_steps
        ts = (np.arange(0, num_inference_steps) * ratio).round()[::-1].copy().astype(np.int64)
        self.timesteps = torch.from_numpy(ts)
        if device is not None:
            self.timesteps = self.timesteps.to(device)

    def scale_model_input(self, sample, timestep=None):
        return sample

    def input_scale(self, timestep) -> float:
        return 1.0

    def add_noise(self, original_samples, noise, timesteps):
        t = int(torch.as_tensor(timesteps).reshape(-1)[0])
        a = float(self._ac[t])
        return ops.cfg_step(original_samples.contiguous(), None, None, noise.contiguous(),
                            c_x=a ** 0.5, c_noise=(1.0 - a) ** 0.5)

    def step(self, model_output, timestep, sample, generator=None, noise=None, return_dict=True, **kw):
        c_x, c_eps, c_n = self.step_coefficients(timestep)
        z = None
        if c_n != 0.0:
            z = noise if noise is not None else torch.randn(
                model_output.shape, generator=generator, device=model_output.device, dtype=model_output.dtype)
        mo = model_output.contiguous()
        prev = ops.cfg_step(sample.contiguous(), mo, mo, z, guidance=0.0, c_x=c_x, c_eps=c_eps, c_noise=c_n)
        self._advance()
        return _StepOutput(prev_sample=prev)

    def _advance(self):
        pass

    def fused_cfg_step(self, latents, eps_u, eps_c, guidance_scale, timestep, generator=None):
        """CFG combine (train_util.py:296-300) and `scheduler.step` (train_util.py:324) in ONE launch:
        x' = c_x x + c_eps (eps_u + g (eps_c - eps_u)) + c_noise z."""
        c_x, c_eps, c_n = self.step_coefficients(timestep)
        z = None
        if c_n != 0.0:
            z = torch.randn(latents.shape, generator=generator, device=latents.device, dtype=latents.dtype)
        out = ops.cfg_step(latents.contiguous(), eps_u, eps_c, z, guidance=guidance_scale, c_x=c_x, c_eps=c_eps,
                           c_noise=c_n)
        self._advance()
        return out

    # -- the arithmetic -------------------------------------------------------------------------
    def step_coefficients(self, timestep) -> tuple[float, float, float]:
        raise NotImplementedError

    def _t(self, timestep) -> int:
        return int(timestep.item()) if torch.is_tensor(timestep) else int(timestep)

    def _prev_alpha(self, t: int) -> tuple[float, float]:
        prev_t = t - self.config.num_train_timesteps // self.num_inference_steps
        a_t = float(self._ac[t])
        a_p = float(self._ac[prev_t]) if prev_t >= 0 else 1.0
        return a_t, a_p


class DDIMScheduler(SchedulerBase):
    """eta = 0, set_alpha_to_one, 'leading' spacing."""

    def step_coefficients(self, timestep):
        a_t, a_p = self._prev_alpha(self._t(timestep))
        if self.config.prediction_type == "epsilon":
            # x0 = (x - sqrt(1-a_t) e)/sqrt(a_t);  x' = sqrt(a_p) x0 + sqrt(1-a_p) e
            c_x = (a_p / a_t) ** 0.5
            c_e = (1.0 - a_p) ** 0.5 - (a_p * (1.0 - a_t) / a_t) ** 0.5
        elif self.config.prediction_type == "v_prediction":
            # x0 = sqrt(a_t) x - sqrt(1-a_t) v;  e = sqrt(a_t) v + sqrt(1-a_t) x
            c_x = (a_p * a_t) ** 0.5 + ((1.0 - a_p) * (1.0 - a_t)) ** 0.5
            c_e = ((1.0 - a_p) * a_t) ** 0.5 - (a_p * (1.0 - a_t)) ** 0.5
        else:
            raise ValueError(self.config.prediction_type)
        return c_x, c_e, 0.0


class DDPMScheduler(SchedulerBase):
    """variance_type 'fixed_small', no clipping."""

    def step_coefficients(self, timestep):
        t = self._t(timestep)
        a_t, a_p = self._prev_alpha(t)
        b_t, b_p = 1.0 - a_t, 1.0 - a_p
        cur_a = a_t / a_p
        cur_b = 1.0 - cur_a
        k0 = a_p ** 0.5 * cur_b / b_t          # weight of x0 in the posterior mean
        kx = cur_a ** 0.5 * b_p / b_t          # weight of x_t
        if self.config.prediction_type == "epsilon":
            c_x = kx + k0 / a_t ** 0.5
            c_e = -k0 * (b_t / a_t) ** 0.5
        elif self.config.prediction_type == "v_prediction":
            c_x = kx + k0 * a_t ** 0.5
            c_e = -k0 * b_t ** 0.5
        else:
            raise ValueError(self.config.prediction_type)
        var = max(b_p / b_t * cur_b, 1e-20)
        c_n = var ** 0.5 if t > 0 else 0.0
        return c_x, c_e, c_n


class EulerAncestralDiscreteScheduler(SchedulerBase):
    """'linspace' spacing (the 0.20.2 default of this class), epsilon / v prediction."""

    def __init__(self, **kw):
        kw.pop("clip_sample", None)
        super().__init__(**kw)
        self._sig_all = ((1.0 - self._ac) / self._ac) ** 0.5
        self.set_timesteps(self.config.num_train_timesteps)

    def set_timesteps(self, num_inference_steps: int, device=None):
        self.num_inference_steps = num_inference_steps
        n = self.config.num_train_timesteps
        ts = np.linspace(0, n - 1, num_inference_steps, dtype=float)[::-1].copy()
        sig = np.interp(ts, np.arange(0, n), self._sig_all)
        self._sigmas = np.concatenate([sig, [0.0]])
        self.sigmas = torch.from_numpy(self._sigmas.astype(np.float32))
        self.timesteps = torch.from_numpy(ts.astype(np.float32))
        if device is not None:
            self.timesteps = self.timesteps.to(device)
        self.init_noise_sigma = torch.tensor(float(self._sigmas.max()), dtype=torch.float32)
        self._index = 0

    def _idx(self, timestep) -> int:
        t = float(timestep.item()) if torch.is_tensor(timestep) else float(timestep)
        return int(np.argmin(np.abs(self.timesteps.cpu().numpy().astype(np.float64) - t)))

    def input_scale(self, timestep) -> float:
        s = float(self._sigmas[self._idx(timestep)])
        return 1.0 / (s * s + 1.0) ** 0.5

    def scale_model_input(self, sample, timestep=None):
        return ops.cfg_step(sample.contiguous(), None, None, None, c_x=self.input_scale(timestep))

    def step_coefficients(self, timestep):
        i = self._idx(timestep)
        s_from, s_to = float(self._sigmas[i]), float(self._sigmas[i + 1])
        s_up = (s_to ** 2 * (s_from ** 2 - s_to ** 2) / s_from ** 2) ** 0.5
        s_down = (s_to ** 2 - s_up ** 2) ** 0.5
        dt = s_down - s_from
        if self.config.prediction_type == "epsilon":
            # x0 = x - s e ; derivative = e ; x' = x + e*dt + z*s_up
            return 1.0, dt, s_up
        if self.config.prediction_type == "v_prediction":
            # x0 = v*(-s/sqrt(s^2+1)) + x/(s^2+1); derivative = (x - x0)/s
            q = s_from ** 2 + 1.0
            return 1.0 + dt * (1.0 - 1.0 / q) / s_from, dt / q ** 0.5, s_up
        raise ValueError(self.config.prediction_type)

    def add_noise(self, original_samples, noise, timesteps):
        s = float(self._sigmas[self._idx(torch.as_tensor(timesteps).reshape(-1)[0])])
        return ops.cfg_step(original_samples.contiguous(), None, None, noise.contiguous(), c_x=1.0, c_noise=s)


class LMSDiscreteScheduler(EulerAncestralDiscreteScheduler):
    """Linear multistep (order <= 4) scheduler, the SD-1 inference default of the reference
    (eval-scripts/generate_images_sd1.py:51; constructor args conceptmod/textsliders/model_util.py:415-423).
    diffusers==0.20.2 arithmetic restated ('linspace' spacing, same sigma table and `scale_model_input` as Euler-a):
        derivative_t = (x - x0_t) / sigma_t                      (= the CFG-combined eps for epsilon prediction)
        x' = x + sum_{i < order} c_i * derivative_{t-i},   c_i = integral over [sigma_t, sigma_{t+1}] of the
        Lagrange basis polynomial of node sigma_{t-i} among the last `order` sigmas.
    The update needs the derivative history, so it is not a single affine step: one CFG launch produces the
    derivative, then one axpy launch of the same kernel per history term (<= 4)."""

    order = 4

    def set_timesteps(self, num_inference_steps: int, device=None):
        super().set_timesteps(num_inference_steps, device)
        self._sig32 = self._sigmas.astype(np.float32).astype(np.float64)      # diffusers keeps the table in float32
        self.derivatives = []

    def lms_coefficient(self, order: int, t: int, current_order: int) -> float:
        from scipy import integrate
        sig = self._sig32

        def basis(tau):
            prod = 1.0
            for k in range(order):
                if current_order == k:
                    continue
                prod *= (tau - sig[t - k]) / (sig[t - current_order] - sig[t - k])
            return prod

        return float(integrate.quad(basis, sig[t], sig[t + 1], epsrel=1e-4)[0])

    def step_coefficients(self, timestep):
        raise NotImplementedError("LMS is a multistep rule: use fused_cfg_step / step (no single affine coefficients; "
                                  "the CUDA-graph engine supports ddim / ddpm / euler_a)")

    def _derivative(self, latents, eps_u, eps_c, guidance_scale, sigma):
        if self.config.prediction_type == "epsilon":          # x0 = x - sigma e  ->  derivative = e
            return ops.cfg_step(None, eps_u, eps_c, None, guidance=guidance_scale)
        if self.config.prediction_type == "v_prediction":     # x0 = -sigma/sqrt(q) v + x/q  ->  (x - x0)/sigma
            q = sigma * sigma + 1.0
            return ops.cfg_step(latents.contiguous(), eps_u, eps_c, None, guidance=guidance_scale, c_x=sigma / q,
                                c_eps=1.0 / q ** 0.5)
        raise ValueError(self.config.prediction_type)

    def fused_cfg_step(self, latents, eps_u, eps_c, guidance_scale, timestep, generator=None):
        i = self._idx(timestep)
        d = self._derivative(latents, eps_u, eps_c, guidance_scale, float(self._sig32[i]))
        self.derivatives.append(d)
        if len(self.derivatives) > self.order:
            self.derivatives.pop(0)
        order = min(i + 1, self.order, len(self.derivatives))
        prev = latents.contiguous()
        for cur, dv in zip(range(order), reversed(self.derivatives)):
            prev = ops.cfg_step(prev, dv, dv, None, guidance=0.0, c_x=1.0, c_eps=self.lms_coefficient(order, i, cur))
        return prev

    def step(self, model_output, timestep, sample, order: int = 4, generator=None, noise=None, return_dict=True, **kw):
        mo = model_output.contiguous()
        return _StepOutput(prev_sample=self.fused_cfg_step(sample, mo, mo, 0.0, timestep))

    def add_noise(self, original_samples, noise, timesteps):
        s = float(self._sig32[self._idx(torch.as_tensor(timesteps).reshape(-1)[0])])
        return ops.cfg_step(original_samples.contiguous(), None, None, noise.contiguous(), c_x=1.0, c_noise=s)
