"""SD-1.x / 2.x text-slider trainer — drop-in for conceptmod/textsliders/train_lora.py
(`train(config, prompts, device)` :32-340, `main(args)` :343-364, CLI flags :367-417).

    python -m sliders_conceptmod_b200.train_lora --config_file data/config.yaml --alpha 1.0 --rank 4 \
        --name ageslider --attributes 'male, female'

The iteration body lives in slider_loop.TextSliderTrainer; `--engine` runs the same iteration as three
replayed CUDA graphs over static buffers with the flat-buffer AdamW (engine.SliderEngine; AdamW + the
configured LR schedule only) and is the path `torchrun` data-parallel training uses.
"""
from __future__ import annotations

import argparse

import torch

from . import config_util, prompt_util
from .config_util import RootConfig
from .prompt_util import PromptSettings
from .slider_loop import TextSliderTrainer


def train(config: RootConfig, prompts: list[PromptSettings], device, engine: bool = False, seed=None):
    trainer = TextSliderTrainer(config, prompts, device, xl=False, seed=seed)
    if config.logging.verbose:
        print(trainer.metadata)
    if engine:
        from .engine_loop import run_with_engine
        return run_with_engine(trainer, suffix=".pt")
    trainer.run(suffix=".pt")            # train_lora.py:311-328 writes `{name}_{i}steps.pt` / `{name}_last.pt`
    print("Done.")


def main(args):
    config = config_util.load_config_from_yaml(args.config_file)
    if args.name is not None:
        config.save.name = args.name
    attributes = []
    if args.attributes is not None:
        attributes = [a.strip() for a in args.attributes.split(",")]
    config.network.alpha = args.alpha
    config.network.rank = args.rank
    config.save.name += f"_alpha{args.alpha}"
    config.save.name += f"_rank{config.network.rank}"
    config.save.name += f"_{config.network.training_method}"
    config.save.path += f"/{config.save.name}"
    prompts = prompt_util.load_prompts_from_yaml(config.prompts_file, attributes)
    device = torch.device(f"cuda:{args.device}")
    train(config=config, prompts=prompts, device=device, engine=args.engine)


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser()
    parser.add_argument("--config_file", required=False, default="data/config.yaml", help="Config file for training.")
    parser.add_argument("--alpha", type=float, required=True, help="LoRA weight.")
    parser.add_argument("--rank", type=int, required=False, default=4, help="Rank of LoRA.")
    parser.add_argument("--device", type=int, required=False, default=0, help="Device to train on.")
    parser.add_argument("--name", type=str, required=False, default=None, help="Name of the slider.")
    parser.add_argument("--attributes", type=str, required=False, default=None,
                        help="attributes to disentangle (comma separated string)")
    parser.add_argument("--engine", action="store_true",
                        help="B200 extension: CUDA-graph engine + flat AdamW (also the torchrun data-parallel path)")
    return parser


if __name__ == "__main__":
    main(build_parser().parse_args())
