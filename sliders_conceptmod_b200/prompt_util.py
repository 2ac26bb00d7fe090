"""Prompt schema, embedding cache and the slider guidance objective.

Reference: trainscripts/imagesliders/prompt_util.py:31-174 (upstream dialect,
`unconditional_latents=`) and conceptmod/textsliders/prompt_util.py:50-201 (fork dialect,
`negative_latents=`).  Both are  MSE(target, neutral +/- g*(positive - q))  and are served by ONE
fused CUDA kernel (loss + gradient in a single pass): `ops.guidance_loss`.
"""
from __future__ import annotations

import copy
from typing import Literal, Optional, Union

import torch
import yaml
from pydantic import BaseModel, model_validator

from . import ops

ACTION_TYPES = Literal["erase", "enhance"]


class PromptEmbedsXL:
    text_embeds: torch.FloatTensor
    pooled_embeds: torch.FloatTensor

    def __init__(self, *args) -> None:
        self.text_embeds = args[0]
        self.pooled_embeds = args[1]


PROMPT_EMBEDDING = Union[torch.FloatTensor, PromptEmbedsXL]


class PromptEmbedsCache:
    def __init__(self):
        self.prompts: dict[str, PROMPT_EMBEDDING] = {}

    def __setitem__(self, name: str, value: PROMPT_EMBEDDING) -> None:
        self.prompts[name] = value

    def __getitem__(self, name: str) -> Optional[PROMPT_EMBEDDING]:
        return self.prompts.get(name)


class PromptSettings(BaseModel):
    target: str
    positive: Optional[str] = None      # if None, target will be used
    negative: Optional[str] = None      # fork only (conceptmod/textsliders/prompt_util.py:66)
    unconditional: str = ""
    neutral: Optional[str] = None       # if None, unconditional will be used
    action: ACTION_TYPES = "erase"
    guidance_scale: float = 1.0
    resolution: int = 512
    dynamic_resolution: bool = False
    batch_size: int = 1
    dynamic_crops: bool = False

    @model_validator(mode="before")
    @classmethod
    def fill_prompts(cls, values):
        values = dict(values)
        if "target" not in values:
            raise ValueError("target must be specified")
        if "positive" not in values:
            values["positive"] = values["target"]
        if "unconditional" not in values:
            values["unconditional"] = ""
        if "neutral" not in values:
            values["neutral"] = values["unconditional"]
        return values

    def json(self, **kw):   # pydantic-v1 spelling used by the reference (train_lora.py:39)
        return self.model_dump_json(**kw)


class _GuidanceLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, target, positive, neutral, q, guidance, sign):
        loss, grad = ops.guidance_loss(target, positive, neutral, q, guidance, sign,
                                       want_grad=ctx.needs_input_grad[0])
        ctx.grad = grad
        return loss

    @staticmethod
    def backward(ctx, gout):
        g = ctx.grad
        if g is None:
            raise RuntimeError("guidance loss: backward called twice or without a target gradient")
        ctx.grad = None
        ops.scale_by_device_scalar(g, gout.to(torch.float32).reshape(1).contiguous())
        return g, None, None, None, None, None


def guidance_mse(target, positive, neutral, q, guidance_scale: float, sign: int) -> torch.Tensor:
    """mean((target - (neutral + sign*g*(positive - q)))^2) as an fp32 0-dim tensor; differentiable w.r.t.
    `target` only (the other three carry requires_grad=False in the reference, train_lora.py:280-282)."""
    return _GuidanceLossFn.apply(target, positive, neutral, q, float(guidance_scale), int(sign))


def plain_mse(target, reference) -> torch.Tensor:
    """MSELoss()(target, reference) through the same fused kernel (p = q = NULL): the image-slider objective
    (trainscripts/imagesliders/train_lora-scale.py:303,327), differentiable w.r.t. `target`."""
    return _GuidanceLossFn.apply(target, None, reference.to(target.dtype), None, 0.0, 1)


def _check_loss_fn(loss_fn) -> None:
    if not (isinstance(loss_fn, torch.nn.MSELoss) and loss_fn.reduction == "mean"):
        raise NotImplementedError(
            "PromptEmbedsPair: the fused objective implements torch.nn.MSELoss() (reduction='mean'), the only "
            "criterion the reference trainers construct (train_lora.py:96)")


class PromptEmbedsPair:
    """Accepts both constructor forms:
       upstream  (loss_fn, target, positive, unconditional, neutral, settings)
       fork      (loss_fn, target, positive, unconditional, neutral, negative, settings)"""

    def __init__(self, loss_fn, target, positive, unconditional, neutral, *rest) -> None:
        if len(rest) == 1:
            negative, settings = None, rest[0]
        elif len(rest) == 2:
            negative, settings = rest
        else:
            raise TypeError("PromptEmbedsPair(loss_fn, target, positive, unconditional, neutral[, negative], settings)")
        _check_loss_fn(loss_fn)
        self.loss_fn = loss_fn
        self.target, self.positive, self.unconditional, self.neutral = target, positive, unconditional, neutral
        self.negative = negative
        self.guidance_scale = settings.guidance_scale
        self.resolution = settings.resolution
        self.dynamic_resolution = settings.dynamic_resolution
        self.batch_size = settings.batch_size
        self.dynamic_crops = settings.dynamic_crops
        self.action = settings.action

    def embeddings(self):
        return [self.target, self.positive, self.negative, self.unconditional, self.neutral]

    @staticmethod
    def _q(unconditional_latents, negative_latents):
        if (unconditional_latents is None) == (negative_latents is None):
            raise TypeError("pass exactly one of unconditional_latents= (upstream) or negative_latents= (fork)")
        return unconditional_latents if unconditional_latents is not None else negative_latents

    def _erase(self, target_latents, positive_latents, neutral_latents, unconditional_latents=None,
               negative_latents=None):
        """Target latents are going not to have the positive concept."""
        q = self._q(unconditional_latents, negative_latents)
        return guidance_mse(target_latents, positive_latents, neutral_latents, q, self.guidance_scale, -1)

    def _enhance(self, target_latents, positive_latents, neutral_latents, unconditional_latents=None,
                 negative_latents=None):
        """Target latents are going to have the positive concept."""
        q = self._q(unconditional_latents, negative_latents)
        return guidance_mse(target_latents, positive_latents, neutral_latents, q, self.guidance_scale, +1)

    def loss(self, **kwargs):
        if self.action == "erase":
            return self._erase(**kwargs)
        elif self.action == "enhance":
            return self._enhance(**kwargs)
        raise ValueError("action must be erase or enhance")


def load_prompts_from_yaml(path, attributes=[]):
    with open(path, "r") as f:
        prompts = yaml.safe_load(f)
    if prompts is None or len(prompts) == 0:
        raise ValueError("prompts file is empty")
    if len(attributes) != 0:
        newprompts = []
        for i in range(len(prompts)):
            for att in attributes:
                c = copy.deepcopy(prompts[i])
                for key in ("target", "positive", "neutral", "unconditional"):
                    c[key] = att + " " + c[key]
                if c.get("negative") is not None:   # fork raises KeyError here when absent
                    c["negative"] = att + " " + c["negative"]
                newprompts.append(c)
    else:
        newprompts = copy.deepcopy(prompts)
    return [PromptSettings(**p) for p in newprompts]
