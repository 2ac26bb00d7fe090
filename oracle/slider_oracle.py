"""CPU oracle of the concept-slider hot path — TEST INFRASTRUCTURE ONLY.

A plain-PyTorch (CPU) restatement of the reference's arithmetic, each function citing the reference
lines it follows.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product package (sliders_conceptmod_b200/) never does.

Pinning: checked against golden vectors produced by importing the reference's own classes in the
authoring container (tests/golden/make_golden.py -> tests/golden/*.pt, tests/test_oracle_golden.py).
The scheduler rules live in diffusers==0.20.2, which is absent from /root/reference: DDIM is pinned by
the in-repo `prev_step` (trainscripts/imagesliders/train_lora-scale.py:31-39); DDPM / Euler-ancestral
are restated from the published formulas and are "parity unpinned" (see DESIGN.md).
"""
from __future__ import annotations

import math
from types import SimpleNamespace

import torch
import torch.nn.functional as F


# ----------------------------------------------------------------------------- LoRA layer (a1, a2)
def lora_forward(org_forward, lora_down, lora_up, multiplier, scale, x):
    """conceptmod/textsliders/lora.py:134-138 — same op order, so bf16 rounds where the reference rounds."""
    return org_forward(x) + lora_up(lora_down(x)) * multiplier * scale


def dora_forward(x, weight, bias, lora_down_w, lora_up_w, dora_scale, multiplier, scale, stride=1, padding=0):
    """conceptmod/textsliders/dora.py:122-163 (DoRAModule.apply_weight_decompose + forward), same op order:
    W' = (W + B.A) * (dora_scale / ||W + B.A||) with the norm per INPUT channel over outputs and taps (detached),
    y = org(x) + ops(x, W' - W) * scale * multiplier."""
    if weight.dim() == 4:
        lora_weight = torch.mm(lora_up_w.flatten(start_dim=1), lora_down_w.flatten(start_dim=1)).reshape(weight.shape)
    else:
        lora_weight = lora_up_w @ lora_down_w
    w = weight + lora_weight
    norm = (w.transpose(0, 1).reshape(w.shape[1], -1).norm(dim=1, keepdim=True)
            .reshape(w.shape[1], *[1] * (w.dim() - 1)).transpose(0, 1)).detach()
    delta = (w * (dora_scale / norm) - weight).to(x.dtype)
    if weight.dim() == 4:
        return (F.conv2d(x, weight, bias, stride=stride, padding=padding)
                + F.conv2d(x, delta, padding=padding, stride=stride) * scale * multiplier)
    return F.linear(x, weight, bias) + F.linear(x, delta) * scale * multiplier


def attach(network):
    """Re-point every wrapped layer of a (sliders_conceptmod_b200) LoRANetwork at the oracle arithmetic.
    Module selection / naming stays the product's host logic (pinned separately by golden key lists)."""
    for m in network.unet_loras:
        org = m._org
        base_forward = type(org).forward.__get__(org)     # the un-patched Linear/Conv2d forward

        def fwd(x, m=m, base_forward=base_forward):
            return lora_forward(base_forward, m.lora_down, m.lora_up, m.multiplier, m.scale, x)

        org.forward = fwd
    return network


# ----------------------------------------------------------------------------- objective (a5)
def guidance_loss(target, positive, neutral, q, guidance_scale, action):
    """trainscripts/imagesliders/prompt_util.py:108-148 (q = unconditional) and
    conceptmod/textsliders/prompt_util.py:134-174 (q = negative), loss_fn = MSELoss() (train_lora.py:96)."""
    if action == "erase":
        tgt = neutral - guidance_scale * (positive - q)
    elif action == "enhance":
        tgt = neutral + guidance_scale * (positive - q)
    else:
        raise ValueError("action must be erase or enhance")
    return F.mse_loss(target, tgt)


# ----------------------------------------------------------------------------- CFG (a6)
def cfg_combine(noise_pred, guidance_scale):
    """conceptmod/textsliders/train_util.py:297-300."""
    u, c = noise_pred.chunk(2)
    return u + guidance_scale * (c - u)


# ----------------------------------------------------------------------------- schedulers (a7)
def alphas_cumprod():
    """model_util.py:397-404 constructor arguments: scaled_linear 0.00085 -> 0.012 over 1000 steps."""
    betas = torch.linspace(0.00085 ** 0.5, 0.012 ** 0.5, 1000, dtype=torch.float32) ** 2
    return torch.cumprod(1.0 - betas, dim=0)


class OracleScheduler:
    """Minimal scheduler object with the surface the reference loop uses."""

    def __init__(self, kind="ddim", prediction_type="epsilon"):
        self.kind, self.prediction_type = kind, prediction_type
        self.alphas_cumprod = alphas_cumprod()
        self.final_alpha_cumprod = torch.tensor(1.0)
        self.config = SimpleNamespace(num_train_timesteps=1000)
        self.init_noise_sigma = torch.tensor(1.0)
        self.num_inference_steps = None
        if kind in ("euler_a", "lms"):
            self._sig_all = ((1 - self.alphas_cumprod.double()) / self.alphas_cumprod.double()).sqrt().numpy()
            self.set_timesteps(1000)

    def set_timesteps(self, n, device=None):
        import numpy as np
        self.num_inference_steps = n
        self.derivatives = []
        if self.kind in ("euler_a", "lms"):
            ts = np.linspace(0, 999, n, dtype=float)[::-1].copy()
            sig = np.interp(ts, np.arange(0, 1000), self._sig_all)
            self.sigmas = torch.from_numpy(np.concatenate([sig, [0.0]]).astype(np.float32))
            self.timesteps = torch.from_numpy(ts.astype(np.float32))
            self.init_noise_sigma = self.sigmas.max()
        else:
            ratio = 1000 // n
            self.timesteps = torch.from_numpy((np.arange(0, n) * ratio).round()[::-1].copy().astype(np.int64))

    def _sigma_index(self, t):
        return int((self.timesteps - float(t)).abs().argmin())

    def scale_model_input(self, sample, t):
        if self.kind in ("euler_a", "lms"):
            s = self.sigmas[self._sigma_index(t)]
            return sample / ((s ** 2 + 1) ** 0.5)
        return sample

    def add_noise(self, x0, noise, timesteps):
        a = self.alphas_cumprod[int(torch.as_tensor(timesteps).reshape(-1)[0])]
        return a ** 0.5 * x0 + (1 - a) ** 0.5 * noise

    def step(self, model_output, timestep, sample, noise=None):
        if self.kind == "ddim":
            prev = ddim_prev_step(model_output, int(timestep), self, sample)
        elif self.kind == "ddpm":
            prev = ddpm_step(model_output, int(timestep), self, sample, noise)
        elif self.kind == "lms":
            prev = lms_step(model_output, timestep, self, sample)
        else:
            prev = euler_a_step(model_output, timestep, self, sample, noise)
        return SimpleNamespace(prev_sample=prev)


def ddim_prev_step(model_output, timestep, scheduler, sample):
    """trainscripts/imagesliders/train_lora-scale.py:31-39 (eta = 0 DDIM), statement by statement."""
    prev_timestep = timestep - scheduler.config.num_train_timesteps // scheduler.num_inference_steps
    alpha_prod_t = scheduler.alphas_cumprod[timestep]
    alpha_prod_t_prev = scheduler.alphas_cumprod[prev_timestep] if prev_timestep >= 0 else scheduler.final_alpha_cumprod
    beta_prod_t = 1 - alpha_prod_t
    if scheduler.prediction_type == "v_prediction":
        pred_original_sample = alpha_prod_t ** 0.5 * sample - beta_prod_t ** 0.5 * model_output
        model_output = alpha_prod_t ** 0.5 * model_output + beta_prod_t ** 0.5 * sample
    else:
        pred_original_sample = (sample - beta_prod_t ** 0.5 * model_output) / alpha_prod_t ** 0.5
    pred_sample_direction = (1 - alpha_prod_t_prev) ** 0.5 * model_output
    return alpha_prod_t_prev ** 0.5 * pred_original_sample + pred_sample_direction


def ddpm_step(model_output, t, scheduler, sample, noise):
    """DDPM ancestral step, variance 'fixed_small' (diffusers 0.20.2 DDPMScheduler.step)."""
    prev_t = t - 1000 // scheduler.num_inference_steps
    a_t = scheduler.alphas_cumprod[t]
    a_p = scheduler.alphas_cumprod[prev_t] if prev_t >= 0 else torch.tensor(1.0)
    b_t, b_p = 1 - a_t, 1 - a_p
    cur_a = a_t / a_p
    cur_b = 1 - cur_a
    if scheduler.prediction_type == "v_prediction":
        x0 = a_t ** 0.5 * sample - b_t ** 0.5 * model_output
    else:
        x0 = (sample - b_t ** 0.5 * model_output) / a_t ** 0.5
    mean = (a_p ** 0.5 * cur_b / b_t) * x0 + (cur_a ** 0.5 * b_p / b_t) * sample
    if t > 0:
        var = torch.clamp(b_p / b_t * cur_b, min=1e-20)
        mean = mean + var ** 0.5 * noise
    return mean


def euler_a_step(model_output, timestep, scheduler, sample, noise):
    """Euler-ancestral step (diffusers 0.20.2 EulerAncestralDiscreteScheduler.step)."""
    i = scheduler._sigma_index(timestep)
    sigma = scheduler.sigmas[i]
    if scheduler.prediction_type == "v_prediction":
        x0 = model_output * (-sigma / (sigma ** 2 + 1) ** 0.5) + (sample / (sigma ** 2 + 1))
    else:
        x0 = sample - sigma * model_output
    s_from, s_to = scheduler.sigmas[i], scheduler.sigmas[i + 1]
    s_up = (s_to ** 2 * (s_from ** 2 - s_to ** 2) / s_from ** 2) ** 0.5
    s_down = (s_to ** 2 - s_up ** 2) ** 0.5
    derivative = (sample - x0) / sigma
    return sample + derivative * (s_down - sigma) + noise * s_up


def lms_coefficient(scheduler, order, t, current_order):
    """diffusers 0.20.2 LMSDiscreteScheduler.get_lms_coefficient: integral of the Lagrange basis polynomial of node
    sigma[t - current_order] over [sigma[t], sigma[t + 1]] (scipy quad, epsrel 1e-4)."""
    from scipy import integrate
    sig = scheduler.sigmas.double().numpy()

    def lms_derivative(tau):
        prod = 1.0
        for k in range(order):
            if current_order == k:
                continue
            prod *= (tau - sig[t - k]) / (sig[t - current_order] - sig[t - k])
        return prod

    return integrate.quad(lms_derivative, sig[t], sig[t + 1], epsrel=1e-4)[0]


def lms_step(model_output, timestep, scheduler, sample, order=4):
    """Linear multistep step (diffusers 0.20.2 LMSDiscreteScheduler.step, the scheduler the reference builds at
    conceptmod/textsliders/model_util.py:415-423 and eval-scripts/generate_images_sd1.py:51): parity unpinned
    (third-party arithmetic restated from the published rule)."""
    i = scheduler._sigma_index(timestep)
    sigma = scheduler.sigmas[i]
    if scheduler.prediction_type == "v_prediction":
        x0 = model_output * (-sigma / (sigma ** 2 + 1) ** 0.5) + (sample / (sigma ** 2 + 1))
    else:
        x0 = sample - sigma * model_output
    derivative = (sample - x0) / sigma
    scheduler.derivatives.append(derivative)
    if len(scheduler.derivatives) > order:
        scheduler.derivatives.pop(0)
    order = min(i + 1, order, len(scheduler.derivatives))
    coeffs = [lms_coefficient(scheduler, order, i, cur) for cur in range(order)]
    return sample + sum(c * d for c, d in zip(coeffs, reversed(scheduler.derivatives)))


# ----------------------------------------------------------------------------- sampler helpers (a6, a7)
def predict_noise(unet, scheduler, timestep, latents, text_embeddings, guidance_scale=7.5, **unet_kw):
    """conceptmod/textsliders/train_util.py:276-302 (and :449-489 with added_cond_kwargs)."""
    x = torch.cat([latents] * 2)
    x = scheduler.scale_model_input(x, timestep)
    noise_pred = unet(x, timestep, encoder_hidden_states=text_embeddings, **unet_kw).sample
    return cfg_combine(noise_pred, guidance_scale)


@torch.no_grad()
def diffusion(unet, scheduler, latents, text_embeddings, total_timesteps=1000, start_timesteps=0,
              guidance_scale=7.5, noises=None, **unet_kw):
    """conceptmod/textsliders/train_util.py:306-327."""
    for i, timestep in enumerate(scheduler.timesteps[start_timesteps:total_timesteps]):
        noise_pred = predict_noise(unet, scheduler, timestep, latents, text_embeddings, guidance_scale, **unet_kw)
        z = None if noises is None else noises[i]
        latents = scheduler.step(noise_pred, timestep, latents, noise=z).prev_sample
    return latents


def train_iteration(unet, network, scheduler, emb, latents, timesteps_to, guidance_scale, action,
                    max_denoising_steps=50, denoise_guidance=3.0, predict_guidance=1.0, noises=None, pooled=None,
                    time_ids=None):
    """One iteration of conceptmod/textsliders/train_lora.py:155-298 up to loss.backward().  With
    `predict_guidance = denoise_guidance = config.train.cfg`, an `emb["negative"]` entry and `pooled` / `time_ids`
    (SD-XL added_cond_kwargs, rebuilt per call from the same prompt as the text embedding) it is
    conceptmod/textsliders/train_lora_xl.py:165-346.  emb / pooled: dict(target, positive, neutral, unconditional
    [, negative]) of [b,77,D] / [b,1280].  `noises`: the ancestral draws of the partial denoise, in order (euler_a /
    ddpm).  Returns the loss tensor; LoRA gradients are left in .grad."""
    cat = lambda u, c: torch.cat([u, c])

    def pair(key):
        kw = {}
        if pooled is not None:
            kw["added_cond_kwargs"] = {"text_embeds": cat(pooled["unconditional"], pooled[key]),
                                       "time_ids": cat(time_ids, time_ids)}
        return cat(emb["unconditional"], emb[key]), kw

    g = predict_guidance
    with torch.no_grad():
        scheduler.set_timesteps(max_denoising_steps)
        with network:
            e, kw = pair("target")
            denoised = diffusion(unet, scheduler, latents, e, start_timesteps=0, total_timesteps=timesteps_to,
                                 guidance_scale=denoise_guidance, noises=noises, **kw)
        scheduler.set_timesteps(1000)
        t = scheduler.timesteps[int(timesteps_to * 1000 / max_denoising_steps)]
        e, kw = pair("positive")
        pos = predict_noise(unet, scheduler, t, denoised, e, g, **kw)
        e, kw = pair("neutral")
        neu = predict_noise(unet, scheduler, t, denoised, e, g, **kw)
        e, kw = pair("negative" if "negative" in emb else "unconditional")
        unc = predict_noise(unet, scheduler, t, denoised, e, g, **kw)
    with network:
        e, kw = pair("target")
        tgt = predict_noise(unet, scheduler, t, denoised, e, g, **kw)
    loss = guidance_loss(tgt, pos, neu, unc, guidance_scale, action)
    loss.backward()
    return loss, dict(denoised=denoised, positive=pos, neutral=neu, unconditional=unc, target=tgt.detach())


def image_slider_iteration(unet, network, scheduler, emb, lat_low, lat_high, noise, timesteps_to, scale,
                           max_denoising_steps=50):
    """One iteration of trainscripts/imagesliders/train_lora-scale.py:196-334 from the encoded latents on: both
    latents get the SAME noise at timesteps[timesteps_to] (train_util.py:231-233), the adapter at +scale predicts
    the noise of the "high" latents under the positive prompt, at -scale that of the "low" latents under the
    neutral prompt; the two MSE losses back-propagate into the same .grad (no zero_grad between, :327-334)."""
    cat = lambda u, c: torch.cat([u, c])
    with torch.no_grad():
        scheduler.set_timesteps(max_denoising_steps)
        t_add = scheduler.timesteps[timesteps_to:timesteps_to + 1]
        noisy_low = scheduler.add_noise(lat_low, noise, t_add)
        noisy_high = scheduler.add_noise(lat_high, noise, t_add)
        scheduler.set_timesteps(1000)
        t = scheduler.timesteps[int(timesteps_to * 1000 / max_denoising_steps)]
    network.set_lora_slider(scale)
    with network:
        tgt_high = predict_noise(unet, scheduler, t, noisy_high, cat(emb["unconditional"], emb["positive"]), 1)
    loss_high = ((tgt_high.float() - noise.float()) ** 2).mean()
    loss_high.backward()
    network.set_lora_slider(-scale)
    with network:
        tgt_low = predict_noise(unet, scheduler, t, noisy_low, cat(emb["unconditional"], emb["neutral"]), 1)
    loss_low = ((tgt_low.float() - noise.float()) ** 2).mean()
    loss_low.backward()
    return loss_high, loss_low


@torch.no_grad()
def inference_sweep(unet, network, scheduler, text_embeddings, latents, scale, start_noise, guidance_scale,
                    num_inference_steps, noises=None, **unet_kw):
    """Denoising loop of eval-scripts/generate_images_sd1.py:150-176 / generate_images_xl.py:325-358: the adapter
    is off (scale 0) while t > start_noise, then on at `scale`; CFG; scheduler.step."""
    scheduler.set_timesteps(num_inference_steps)
    for i, t in enumerate(scheduler.timesteps):
        network.set_lora_slider(0 if t > start_noise else scale)
        x = scheduler.scale_model_input(torch.cat([latents] * 2), t)
        with network:
            eps = unet(x, t, encoder_hidden_states=text_embeddings, **unet_kw).sample
        eps = cfg_combine(eps, guidance_scale)
        z = None if noises is None else noises[i]
        latents = scheduler.step(eps, t, latents, noise=z).prev_sample
    return latents
