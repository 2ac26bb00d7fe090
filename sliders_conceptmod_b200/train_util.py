"""Sampler helpers of the slider path (reference conceptmod/textsliders/train_util.py).

`predict_noise` / `diffusion` (:276-327), `predict_noise_xl` / `diffusion_xl` (:449-489, :677-708),
`concat_embeddings` (:267-272), `get_initial_latents` (:91-105), `get_add_time_ids` (:976-1011),
`get_optimizer` / `get_lr_scheduler` (:1014-1082).  Same signatures; the classifier-free-guidance
combine and the scheduler update run as ONE fused CUDA kernel per denoising step (ops.cfg_step).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops

UNET_IN_CHANNELS = 4
VAE_SCALE_FACTOR = 8
UNET_ATTENTION_TIME_EMBED_DIM = 256
TEXT_ENCODER_2_PROJECTION_DIM = 1280
UNET_PROJECTION_CLASS_EMBEDDING_INPUT_DIM = 2816


def get_random_noise(batch_size: int, height: int, width: int, generator: torch.Generator = None) -> torch.Tensor:
    return torch.randn((batch_size, UNET_IN_CHANNELS, height // VAE_SCALE_FACTOR, width // VAE_SCALE_FACTOR),
                       generator=generator, device="cpu")


def get_initial_latents(scheduler, n_imgs: int, height: int, width: int, n_prompts: int, generator=None):
    noise = get_random_noise(n_imgs, height, width, generator=generator).repeat(n_prompts, 1, 1, 1)
    return noise * scheduler.init_noise_sigma.to(noise.device)


def concat_embeddings(unconditional: torch.Tensor, conditional: torch.Tensor, n_imgs: int):
    return torch.cat([unconditional, conditional]).repeat_interleave(n_imgs, dim=0)


class _CFGCombineFn(torch.autograd.Function):
    """guided = eps_u + g*(eps_c - eps_u) over the two halves of the UNet output (train_util.py:297-300)."""

    @staticmethod
    def forward(ctx, noise_pred, guidance_scale):
        noise_pred = noise_pred.contiguous()
        half = noise_pred.shape[0] // 2
        ctx.g = float(guidance_scale)
        return ops.cfg_step(None, noise_pred[:half], noise_pred[half:], guidance=ctx.g)

    @staticmethod
    def backward(ctx, gy):
        gy = gy.contiguous()
        half = gy.shape[0]
        gin = torch.empty((2 * half,) + tuple(gy.shape[1:]), device=gy.device, dtype=gy.dtype)
        ops.cfg_step(gy, None, None, None, c_x=1.0 - ctx.g, out=gin[:half])
        ops.cfg_step(gy, None, None, None, c_x=ctx.g, out=gin[half:])
        return gin, None


def _scaled_cfg_input(scheduler, latents, timestep):
    x = torch.cat([latents] * 2)
    return scheduler.scale_model_input(x, timestep)


def predict_noise(unet, scheduler, timestep, latents, text_embeddings, guidance_scale=7.5):
    latent_model_input = _scaled_cfg_input(scheduler, latents, timestep)
    noise_pred = unet(latent_model_input, timestep, encoder_hidden_states=text_embeddings).sample
    return _CFGCombineFn.apply(noise_pred, guidance_scale)


def _fused_cfg_and_step(scheduler, noise_pred, timestep, latents, guidance_scale):
    """a6 + a7 fused: CFG combine and the scheduler's latent update (one launch for the affine rules
    x' = c_x x + c_eps eps + c_noise z; LMS adds one axpy launch per history term)."""
    noise_pred = noise_pred.contiguous()
    half = noise_pred.shape[0] // 2
    return scheduler.fused_cfg_step(latents, noise_pred[:half], noise_pred[half:], guidance_scale, timestep)


@torch.no_grad()
def diffusion(unet, scheduler, latents, text_embeddings, total_timesteps: int = 1000, start_timesteps=0, **kwargs):
    guidance_scale = kwargs.get("guidance_scale", 7.5)
    fused = hasattr(scheduler, "fused_cfg_step")
    for timestep in scheduler.timesteps[start_timesteps:total_timesteps]:
        if fused:
            x = _scaled_cfg_input(scheduler, latents, timestep)
            noise_pred = unet(x, timestep, encoder_hidden_states=text_embeddings).sample
            latents = _fused_cfg_and_step(scheduler, noise_pred, timestep, latents, guidance_scale)
        else:
            noise_pred = predict_noise(unet, scheduler, timestep, latents, text_embeddings, **kwargs)
            latents = scheduler.step(noise_pred, timestep, latents).prev_sample
    return latents


def predict_noise_xl(unet, scheduler, timestep, latents, text_embeddings, add_text_embeddings, add_time_ids,
                     guidance_scale=7.5, guidance_rescale=0.7):
    latent_model_input = _scaled_cfg_input(scheduler, latents, timestep)
    added_cond_kwargs = {"text_embeds": add_text_embeddings, "time_ids": add_time_ids}
    noise_pred = unet(latent_model_input, timestep, encoder_hidden_states=text_embeddings,
                      added_cond_kwargs=added_cond_kwargs).sample
    # the reference computes rescale_noise_cfg(...) and discards it (train_util.py:479-489): plain CFG
    return _CFGCombineFn.apply(noise_pred, guidance_scale)


@torch.no_grad()
def diffusion_xl(unet, scheduler, latents, text_embeddings, add_text_embeddings, add_time_ids,
                 guidance_scale: float = 1.0, total_timesteps: int = 1000, start_timesteps=0):
    fused = hasattr(scheduler, "fused_cfg_step")
    added = {"text_embeds": add_text_embeddings, "time_ids": add_time_ids}
    for timestep in scheduler.timesteps[start_timesteps:total_timesteps]:
        if fused:
            x = _scaled_cfg_input(scheduler, latents, timestep)
            noise_pred = unet(x, timestep, encoder_hidden_states=text_embeddings, added_cond_kwargs=added).sample
            latents = _fused_cfg_and_step(scheduler, noise_pred, timestep, latents, guidance_scale)
        else:
            noise_pred = predict_noise_xl(unet, scheduler, timestep, latents, text_embeddings, add_text_embeddings,
                                          add_time_ids, guidance_scale=guidance_scale, guidance_rescale=0.7)
            latents = scheduler.step(noise_pred, timestep, latents).prev_sample
    return latents


def get_add_time_ids(height: int, width: int, dynamic_crops: bool = False, dtype: torch.dtype = torch.float32):
    if dynamic_crops:
        random_scale = torch.rand(1).item() * 2 + 1
        original_size = (int(height * random_scale), int(width * random_scale))
        crops_coords_top_left = (torch.randint(0, original_size[0] - height, (1,)).item(),
                                 torch.randint(0, original_size[1] - width, (1,)).item())
        target_size = (height, width)
    else:
        original_size, crops_coords_top_left, target_size = (height, width), (0, 0), (height, width)
    add_time_ids = list(original_size + crops_coords_top_left + target_size)
    passed = UNET_ATTENTION_TIME_EMBED_DIM * len(add_time_ids) + TEXT_ENCODER_2_PROJECTION_DIM
    if passed != UNET_PROJECTION_CLASS_EMBEDDING_INPUT_DIM:
        raise ValueError(f"Model expects an added time embedding vector of length "
                         f"{UNET_PROJECTION_CLASS_EMBEDDING_INPUT_DIM}, but a vector of {passed} was created.")
    return torch.tensor([add_time_ids], dtype=dtype)


def get_optimizer(name: str):
    name = name.lower()
    if name.startswith("dadapt"):
        import dadaptation
        if name == "dadaptadam":
            return dadaptation.DAdaptAdam
        if name == "dadaptlion":
            return dadaptation.DAdaptLion
        raise ValueError("DAdapt optimizer must be dadaptadam or dadaptlion")
    if name.endswith("8bit"):
        import bitsandbytes as bnb
        if name == "adam8bit":
            return bnb.optim.Adam8bit
        if name == "lion8bit":
            return bnb.optim.Lion8bit
        raise ValueError("8bit optimizer must be adam8bit or lion8bit")
    if name == "adam":
        return torch.optim.Adam
    if name == "adamw":
        return torch.optim.AdamW
    if name == "lion":
        from lion_pytorch import Lion
        return Lion
    if name == "prodigy":
        import prodigyopt
        return prodigyopt.Prodigy
    raise ValueError("Optimizer must be adam, adamw, lion or Prodigy")


def get_lr_scheduler(name: Optional[str], optimizer, max_iterations: Optional[int], lr_min: Optional[float], **kwargs):
    sched = torch.optim.lr_scheduler
    if name == "cosine":
        return sched.CosineAnnealingLR(optimizer, T_max=max_iterations, eta_min=lr_min, **kwargs)
    if name == "cosine_with_restarts":
        return sched.CosineAnnealingWarmRestarts(optimizer, T_0=max_iterations // 10, T_mult=2, eta_min=lr_min, **kwargs)
    if name == "step":
        return sched.StepLR(optimizer, step_size=max_iterations // 100, gamma=0.999, **kwargs)
    if name == "constant":
        return sched.ConstantLR(optimizer, factor=1, **kwargs)
    if name == "linear":
        return sched.LinearLR(optimizer, factor=0.5, total_iters=max_iterations // 100, **kwargs)
    raise ValueError("Scheduler must be cosine, cosine_with_restarts, step, linear or constant")


def get_random_resolution_in_bucket(bucket_resolution: int = 512) -> tuple[int, int]:
    max_step, min_step = bucket_resolution // 64, (bucket_resolution // 2) // 64
    height = torch.randint(min_step, max_step, (1,)).item() * 64
    width = torch.randint(min_step, max_step, (1,)).item() * 64
    return height, width


# prompt encoding on a CLIP text model (reference train_util.py:108-154, 224-247) — see text_encoder.py
from .text_encoder import (encode_prompts, encode_prompts_xl, text_encode, text_encode_xl,  # noqa: E402,F401
                           text_tokenize)
