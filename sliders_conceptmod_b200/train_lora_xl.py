"""SD-XL text-slider trainer — drop-in for conceptmod/textsliders/train_lora_xl.py
(`train(config, prompts, device, on_step_complete, peft_type, rank, save_file)` :41-394,
`main(args)` :397-419, programmatic `train_lora(...)` :421-458, CLI :460-514).

Keeps the reference's quirks that change results: attention-only targets, alpha left at its
default, AdamW(lr 1e-4, wd 1e-6) + CosineAnnealingLR(T_max 50, eta_min 1e-6) + clip_grad_norm 0.2 regardless of
the config, plain CFG (the rescale is computed and dropped upstream), `.safetensors` output.
"""
from __future__ import annotations

import argparse
from typing import Callable, Optional

import torch

from . import config_util, prompt_util
from .config_util import RootConfig
from .prompt_util import PromptSettings
from .slider_loop import TextSliderTrainer


def train(config: RootConfig, prompts: list[PromptSettings], device, on_step_complete: Optional[Callable] = None,
          peft_type: str = "lora", rank: int = 4, save_file: bool = True, seed=None):
    trainer = TextSliderTrainer(config, prompts, device, xl=True, rank=rank, seed=seed, peft_type=peft_type)
    if config.logging.verbose:
        print(trainer.metadata)
    return trainer.run(on_step_complete=on_step_complete, save_file=save_file, suffix=".safetensors")


def _apply_cli(config, args):
    if args.name is not None:
        config.save.name = args.name
    config.network.alpha = args.alpha
    config.network.rank = args.rank
    config.save.name += f"_alpha{args.alpha}"
    config.save.name += f"_rank{config.network.rank}"
    config.save.name += f"_{config.network.training_method}"
    config.save.path += f"/{config.save.name}"


def main(args):
    config = config_util.load_config_from_yaml(args.config_file)
    attributes = [a.strip() for a in args.attributes.split(",")] if args.attributes is not None else []
    _apply_cli(config, args)
    prompts = prompt_util.load_prompts_from_yaml(config.prompts_file, attributes)
    device = torch.device(f"cuda:{args.device}")
    train(config, prompts, device, None, peft_type=args.peft_type, rank=args.rank)


def train_lora(target, positive, negative, unconditional, alpha=1.0, rank=4, device=0, name=None, attributes=None,
               batch_size=1, config_file="data/config-xl.yaml", resolution=512, steps=None, on_step_complete=None,
               peft_type="lora", save_file=True):
    """Programmatic entry of the fork (train_lora_xl.py:421-458): one prompt quadruple, `enhance`."""
    config = config_util.load_config_from_yaml(config_file)
    if name is not None:
        config.save.name = name
    config.network.alpha = alpha
    config.network.rank = rank
    config.save.name += f"_alpha{alpha}_rank{rank}_{config.network.training_method}"
    config.save.path += f"/{config.save.name}"
    if steps is not None:
        config.train.iterations = steps
    settings = PromptSettings(target=target, positive=positive, negative=negative, unconditional=unconditional,
                              neutral=target, action="enhance", guidance_scale=4, resolution=resolution,
                              batch_size=batch_size)
    prompts = [settings]
    if attributes:
        prompts = []
        for att in [a.strip() for a in attributes.split(",")]:
            prompts.append(PromptSettings(**{**settings.model_dump(),
                                             **{k: att + " " + getattr(settings, k)
                                                for k in ("target", "positive", "neutral", "unconditional", "negative")}}))
    dev = torch.device(f"cuda:{device}") if isinstance(device, int) else device
    return train(config, prompts, dev, on_step_complete, peft_type=peft_type, rank=rank, save_file=save_file)


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser()
    parser.add_argument("--config_file", required=True, help="Config file for training.")
    parser.add_argument("--alpha", type=float, required=True, help="LoRA weight.")
    parser.add_argument("--rank", type=int, required=False, default=4, help="Rank of LoRA.")
    parser.add_argument("--device", type=int, required=False, default=0, help="Device to train on.")
    parser.add_argument("--name", type=str, required=False, default=None, help="Name of the slider.")
    parser.add_argument("--attributes", type=str, required=False, default=None,
                        help="attributes to disentangle (comma separated string)")
    parser.add_argument("--peft_type", type=str, required=False, default="lora", help="dora or lora (default)")
    return parser


if __name__ == "__main__":
    main(build_parser().parse_args())
