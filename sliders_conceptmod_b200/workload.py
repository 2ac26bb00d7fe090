"""Synthetic slider workloads of BASELINE.json (random-init UNet stand-in, synthetic text embeddings).

Shared by bench.py, __graft_entry__.smoke() and the GPU tests so that all three drive the same code.
"""
from __future__ import annotations

import copy
from dataclasses import dataclass

import torch

from . import lora as lora_mod
from . import model_util
from .engine import EngineConfig, SliderEngine
from .lora import LoRANetwork
from .unet import UNet2DConditionModel, UNetConfig


@dataclass
class Workload:
    name: str
    family: str            # "sd14" | "sdxl" | "tiny"
    resolution: int
    batch_size: int = 1
    rank: int = 4
    alpha: float = 1.0
    training_method: str = "noxattn"
    c3lier: bool = True
    scheduler: str = "ddim"
    guidance_scale: float = 4.0
    action: str = "enhance"
    attributes: tuple = ("male", "female")
    max_denoising_steps: int = 50
    lr: float = 2e-4
    max_grad_norm: float | None = None


WORKLOADS = {
    # BASELINE.json configs[1]: the configuration the headline metric is quoted on
    "sd14_text_slider_rank4_512px": Workload("sd14_text_slider_rank4_512px", "sd14", 512),
    # configs[0]: CPU-runnable case (64x64 *latents* = 512 px, 10 iterations) — same shapes, used for parity
    "sd14_cpu_case": Workload("sd14_cpu_case", "sd14", 512, attributes=()),
    # configs[2]: SD-XL text slider, attention-only LoRA, AdamW 1e-4 + clip 0.2
    "sdxl_text_slider_rank4_1024px": Workload("sdxl_text_slider_rank4_1024px", "sdxl", 1024, c3lier=False,
                                              scheduler="euler_a", max_denoising_steps=12, lr=1e-4,
                                              max_grad_norm=0.2, guidance_scale=4.0),
    "tiny": Workload("tiny", "tiny", 128, max_denoising_steps=10),
}


def unet_config(family: str) -> UNetConfig:
    return {"sd14": UNetConfig.sd14, "sdxl": UNetConfig.sdxl, "tiny": UNetConfig.tiny}[family]()


def target_modules(c3lier: bool) -> list[str]:
    mods = list(lora_mod.UNET_TARGET_REPLACE_MODULE_TRANSFORMER)
    if c3lier:   # train_lora.py:44-46
        mods += lora_mod.UNET_TARGET_REPLACE_MODULE_CONV
    return mods


def prompt_pairs(wl: Workload, embed_dim: int, tokens: int = 77) -> list[dict]:
    """The shipped prompts.yaml age slider (2 entries) x attributes (prompt_util.py:183-193), embedded by
    the synthetic text encoder."""
    base = [
        dict(target="male person", positive="male person, very old", unconditional="male person, very young",
             neutral="male person"),
        dict(target="female person", positive="female person, very old", unconditional="female person, very young",
             neutral="female person"),
    ]
    enc = model_util.SyntheticTextEncoder(embed_dim, pooled_dim=1280, tokens=tokens)
    out = []
    atts = wl.attributes or ("",)
    for p in base:
        for att in atts:
            q = {k: (att + " " + v if att else v) for k, v in p.items()}
            e = {k: enc.encode([v]) for k, v in q.items()}
            e["pooled"] = {k: enc.encode_pooled([v]) for k, v in q.items()}
            out.append(e)
    return out


def build(wl: Workload, device, dtype=torch.bfloat16, seed: int = 0, use_graphs: bool = True, lora_b_std: float = 0.0):
    """Returns (engine, unet, network, pairs)."""
    cfg = unet_config(wl.family)
    torch.manual_seed(seed)
    with torch.device(device):
        unet = UNet2DConditionModel(cfg)
    unet.to(dtype=dtype)
    unet.requires_grad_(False)
    unet.eval()
    network = LoRANetwork(unet, rank=wl.rank, multiplier=1.0, alpha=wl.alpha,
                          target_replace=target_modules(wl.c3lier), train_method=wl.training_method)
    network.to(device, dtype=dtype)
    if lora_b_std > 0:
        g = torch.Generator(device="cpu").manual_seed(seed + 1)
        for m in network.unet_loras:
            m.lora_up.weight.data.copy_((torch.randn(m.lora_up.weight.shape, generator=g) * lora_b_std).to(dtype))
    sched = model_util.create_noise_scheduler(wl.scheduler)
    ecfg = EngineConfig(max_denoising_steps=wl.max_denoising_steps, lr=wl.lr, max_grad_norm=wl.max_grad_norm,
                        use_graphs=use_graphs)
    xl_cond = {"text_embeds": 1280, "time_ids": 6} if cfg.addition_embed else None
    engine = SliderEngine(unet, network, sched, ecfg, wl.batch_size, wl.resolution, wl.resolution,
                          (77, cfg.cross_attention_dim), dtype, device, xl_cond=xl_cond)
    pairs = prompt_pairs(wl, cfg.cross_attention_dim)
    return engine, unet, network, pairs


def load_pair(engine: SliderEngine, wl: Workload, pair: dict):
    xl = None
    if engine.xl is not None:
        from .train_util import get_add_time_ids
        ids = get_add_time_ids(wl.resolution, wl.resolution)
        xl = {"text_embeds": pair["pooled"], "time_ids": {k: ids for k in ("target", "positive", "neutral", "unconditional")}}
    engine.load_pair(pair["target"], pair["positive"], pair["neutral"], pair["unconditional"], wl.guidance_scale,
                     wl.action, xl=xl)
