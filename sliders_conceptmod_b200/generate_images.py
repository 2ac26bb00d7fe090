"""Slider inference sweep — the denoising loop of eval-scripts/generate_images_sd1.py:129-192 and
eval-scripts/generate_images_xl.py:325-358 (BASELINE config 5), up to the final latents.

Per denoising step the reference does: `set_lora_slider(0 if t > start_noise else scale)`, CFG-batched UNet call
inside `with network:`, `eps_u + g (eps_c - eps_u)`, `scheduler.step`.  Here the CFG combine and the scheduler
update are ONE fused kernel (`slider_cfg_step`) and the adapter switch is the fused GEMM's `s` argument (s == 0
takes the plain-GEMM path).  VAE decode, PNG / matplotlib output and the prompt CSVs are outside the hot path.
"""
from __future__ import annotations

import argparse
from typing import Optional, Sequence

import torch

from . import model_util, ops
from .lora import LoRANetwork
from .workload import target_modules


@torch.no_grad()
def denoise_with_slider(unet, network: LoRANetwork, scheduler, text_embeddings: torch.Tensor, latents: torch.Tensor,
                        scale: float, start_noise: int, guidance_scale: float, num_inference_steps: int,
                        added_cond_kwargs: Optional[dict] = None, noise_generator: Optional[torch.Generator] = None):
    """One (prompt batch, scale) trajectory.  text_embeddings = cat([uncond, cond]) [2b, 77, D]; latents [b,4,h,w]
    already multiplied by init_noise_sigma.  Returns the final latents."""
    scheduler.set_timesteps(num_inference_steps)
    kw = {} if added_cond_kwargs is None else {"added_cond_kwargs": added_cond_kwargs}
    b = latents.shape[0]
    for t in scheduler.timesteps:
        network.set_lora_slider(scale=0 if t > start_noise else scale)
        x = scheduler.scale_model_input(torch.cat([latents] * 2), t)
        with network:
            eps = unet(x, t, encoder_hidden_states=text_embeddings, **kw).sample.contiguous()
        latents = scheduler.fused_cfg_step(latents, eps[:b], eps[b:], guidance_scale, t, generator=noise_generator)
    return latents


@torch.no_grad()
def sweep(unet, network: LoRANetwork, scheduler, prompt_embeds: torch.Tensor, uncond_embeds: torch.Tensor,
          scales: Sequence[float], seed: int, height: int = 512, width: int = 512, start_noise: int = 800,
          guidance_scale: float = 7.5, num_inference_steps: int = 50, added_cond: Optional[dict] = None,
          rank: int = 0, world: int = 1):
    """`for scale in scales:` of the reference scripts (sd1 :138, xl :495-506): identical seed per scale.
    With world > 1 the (scale) tuples are dealt round-robin to ranks — no collective.
    Returns {scale: final latents}."""
    dev, dt = prompt_embeds.device, prompt_embeds.dtype
    b = prompt_embeds.shape[0]
    emb = torch.cat([uncond_embeds, prompt_embeds])
    out = {}
    for k, scale in enumerate(scales):
        if k % world != rank:
            continue
        gen = torch.manual_seed(seed)
        lat = torch.randn((b, 4, height // 8, width // 8), generator=gen).to(dev)
        scheduler.set_timesteps(num_inference_steps)
        lat = (lat * float(scheduler.init_noise_sigma)).to(dt)
        out[scale] = denoise_with_slider(unet, network, scheduler, emb, lat, scale, start_noise, guidance_scale,
                                         num_inference_steps, added_cond_kwargs=added_cond)
    return out


def main(args):
    dev = torch.device(f"cuda:{args.device}")
    dt = torch.bfloat16
    xl = "xl" in args.model_name.lower()
    if xl:
        _, encs, unet, sched = model_util.load_models_xl(args.model_name, scheduler_name=args.scheduler)
        enc = encs[0]
    else:
        _, enc, unet, sched = model_util.load_models(args.model_name, scheduler_name=args.scheduler)
    unet.to(dev, dt).requires_grad_(False).eval()
    network = LoRANetwork(unet, rank=args.rank, multiplier=1.0, alpha=1.0, target_replace=target_modules(not xl),
                          train_method=args.train_method).to(dev, dt)
    if args.lora is not None:
        sd = torch.load(args.lora) if not args.lora.endswith(".safetensors") else __import__(
            "safetensors.torch", fromlist=["load_file"]).load_file(args.lora)
        network.load_state_dict(sd)
    prompts = [args.prompt] * args.num_samples
    pe = enc.encode(prompts).to(dev, dt)
    ue = enc.encode([""] * args.num_samples).to(dev, dt)
    added = None
    if xl:
        from .train_util import get_add_time_ids
        ids = get_add_time_ids(args.res, args.res).to(dev, dt).repeat(2 * args.num_samples, 1)
        pooled = torch.cat([enc.encode_pooled([""] * args.num_samples), enc.encode_pooled(prompts)]).to(dev, dt)
        added = {"text_embeds": pooled, "time_ids": ids}
    scales = [float(s) for s in args.scales.split(",")]
    res = sweep(unet, network, sched, pe, ue, scales, args.seed, args.res, args.res, args.start_noise,
                args.guidance_scale, args.ddim_steps, added)
    for s, lat in res.items():
        print(f"scale {s:+.1f}: latents {tuple(lat.shape)} mean {lat.float().mean().item():+.4f} std {lat.float().std().item():.4f}")
    if args.save is not None:
        torch.save({s: v.cpu() for s, v in res.items()}, args.save)


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="generate_images", description="slider inference sweep (latents only)")
    p.add_argument("--model_name", default="CompVis/stable-diffusion-v1-4")
    p.add_argument("--lora", default=None, help="slider weights (.pt / .safetensors) written by the trainers")
    p.add_argument("--prompt", default="image of a person")
    p.add_argument("--rank", type=int, default=4)
    p.add_argument("--train_method", default="noxattn")
    p.add_argument("--scheduler", default="ddim")
    p.add_argument("--scales", default="-2,-1,0,1,2")
    p.add_argument("--start_noise", type=int, default=800)
    p.add_argument("--guidance_scale", type=float, default=7.5)
    p.add_argument("--ddim_steps", type=int, default=50)
    p.add_argument("--num_samples", type=int, default=1)
    p.add_argument("--res", type=int, default=512)
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--device", type=int, default=0)
    p.add_argument("--save", default=None)
    return p


if __name__ == "__main__":
    main(build_parser().parse_args())
