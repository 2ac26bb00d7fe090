"""Model / scheduler factories (reference conceptmod/textsliders/model_util.py:112-137, 358-436).

Pretrained weights cannot be fetched here (no network), so `load_models` builds the random-init
UNet stand-in of the requested family plus a deterministic synthetic text encoder; the scheduler
factory keeps the reference's constructor arguments verbatim.
"""
from __future__ import annotations

import dataclasses
import zlib
from typing import Literal

import torch

from .schedulers import DDIMScheduler, DDPMScheduler, EulerAncestralDiscreteScheduler, LMSDiscreteScheduler
from .unet import UNet2DConditionModel, UNetConfig

AVAILABLE_SCHEDULERS = Literal["ddim", "ddpm", "lms", "euler_a"]


def create_noise_scheduler(scheduler_name: AVAILABLE_SCHEDULERS = "ddpm",
                           prediction_type: Literal["epsilon", "v_prediction"] = "epsilon"):
    name = scheduler_name.lower().replace(" ", "_")
    kw = dict(beta_start=0.00085, beta_end=0.012, beta_schedule="scaled_linear", num_train_timesteps=1000,
              prediction_type=prediction_type)
    if name == "ddim":
        return DDIMScheduler(clip_sample=False, **kw)
    if name == "ddpm":
        return DDPMScheduler(clip_sample=False, **kw)
    if name == "euler_a":
        return EulerAncestralDiscreteScheduler(**kw)
    if name == "lms":
        return LMSDiscreteScheduler(**kw)
    raise ValueError(f"Unknown scheduler name: {name}")


class SyntheticTokenizer:
    model_max_length = 77

    def __call__(self, prompts, **kw):
        return list(prompts)


class SyntheticTextEncoder(torch.nn.Module):
    """Stands in for CLIP: a prompt string maps to a fixed N(0,1) embedding seeded by its CRC32."""

    def __init__(self, dim: int = 768, pooled_dim: int | None = None, tokens: int = 77):
        super().__init__()
        self.dim, self.pooled_dim, self.tokens = dim, pooled_dim, tokens
        self._anchor = torch.nn.Parameter(torch.zeros(1), requires_grad=False)

    @property
    def device(self):
        return self._anchor.device

    def encode(self, prompts: list[str]) -> torch.Tensor:
        outs = []
        for p in prompts:
            g = torch.Generator().manual_seed(zlib.crc32(p.encode("utf-8")))
            outs.append(torch.randn(self.tokens, self.dim, generator=g))
        return torch.stack(outs).to(self._anchor.device, self._anchor.dtype if self._anchor.dtype.is_floating_point else torch.float32)

    def encode_pooled(self, prompts: list[str]) -> torch.Tensor:
        outs = []
        for p in prompts:
            g = torch.Generator().manual_seed(zlib.crc32(("pooled:" + p).encode("utf-8")))
            outs.append(torch.randn(self.pooled_dim, generator=g))
        return torch.stack(outs).to(self._anchor.device, self._anchor.dtype)


def encode_prompts(tokenizer, text_encoder, prompts: list[str]) -> torch.Tensor:
    """train_util.encode_prompts (:123-134): the synthetic encoder maps prompt strings directly; a CLIP stand-in
    (text_encoder.CLIPTextModel, `load_models(..., clip_stand_in=True)`) goes through tokenizer + transformer."""
    if isinstance(text_encoder, SyntheticTextEncoder):
        return text_encoder.encode(list(prompts))
    from .text_encoder import encode_prompts as clip_encode_prompts
    return clip_encode_prompts(tokenizer, text_encoder, list(prompts))


def _family(name_or_path: str) -> str:
    n = name_or_path.lower()
    return "sdxl" if ("xl" in n) else ("tiny" if "tiny" in n else "sd14")


def _seeded_unet(cfg: UNetConfig, seed: int) -> UNet2DConditionModel:
    """The random-init stand-in replaces a *pretrained* checkpoint, so it is a fixed function of `seed` and must not
    disturb the caller's RNG stream: the reference trainers are unseeded (train_lora.py:32-76) and a caller-provided
    seed has to keep governing the LoRA init, the prompt-pair / timestep draws and the noise."""
    with torch.random.fork_rng(devices=[]):
        torch.manual_seed(seed)
        return UNet2DConditionModel(cfg)


def load_models(pretrained_model_name_or_path: str, scheduler_name: AVAILABLE_SCHEDULERS, v2: bool = False,
                v_pred: bool = False, weight_dtype: torch.dtype = torch.float32, seed: int = 0,
                clip_stand_in: bool = False):
    fam = _family(pretrained_model_name_or_path)
    cfg = {"sd14": UNetConfig.sd14, "sdxl": UNetConfig.sdxl, "tiny": UNetConfig.tiny}[fam]()
    unet = _seeded_unet(cfg, seed)
    scheduler = create_noise_scheduler(scheduler_name, prediction_type="v_prediction" if v_pred else "epsilon")
    if clip_stand_in and fam == "sd14":       # CLIP-L architecture, random init: 768-wide states = the UNet's context width
        from .text_encoder import load_text_encoders
        toks, encs = load_text_encoders("SD", seed=seed)
        return toks[0], encs[0], unet, scheduler
    return SyntheticTokenizer(), SyntheticTextEncoder(cfg.cross_attention_dim), unet, scheduler


def load_models_xl(pretrained_model_name_or_path: str, scheduler_name: AVAILABLE_SCHEDULERS,
                   weight_dtype: torch.dtype = torch.float32, seed: int = 0):
    if "tiny" in pretrained_model_name_or_path.lower():     # XL topology features at test size
        cfg = dataclasses.replace(UNetConfig.tiny(64), addition_embed=True, use_linear_projection=True)
    else:
        cfg = UNetConfig.sdxl()
    unet = _seeded_unet(cfg, seed)
    scheduler = create_noise_scheduler(scheduler_name)
    enc = SyntheticTextEncoder(cfg.cross_attention_dim, pooled_dim=1280)
    return [SyntheticTokenizer(), SyntheticTokenizer()], [enc, enc], unet, scheduler
