"""`--engine` path of the SD-1.x trainer: the reference iteration (train_lora.py:155-309) issued through
engine.SliderEngine — three replayed CUDA graphs over static buffers, gradients accumulated straight into one
flat fp32 buffer, ONE all-reduce per step under torchrun, one fused AdamW launch.  Random draws (prompt pair,
timesteps_to, initial noise) use the same host generators in the same order as the eager loop; under data
parallel rank k offsets its seed by k and deals prompt pairs round-robin."""
from __future__ import annotations

from pathlib import Path

import torch

from . import parallel, train_util
from .engine import EngineConfig, SliderEngine


def run_with_engine(trainer, suffix: str = ".pt", log_every: int = 10):
    cfg = trainer.config
    if cfg.train.optimizer.lower() != "adamw":
        raise NotImplementedError("--engine implements AdamW (the reference default, data/config.yaml:15)")
    rank, world, _ = parallel.init_from_env()
    pair0 = trainer.prompt_pairs[0]
    emb = pair0.target
    ecfg = EngineConfig(max_denoising_steps=cfg.train.max_denoising_steps, denoise_guidance=float(trainer.denoise_guidance),
                        lr=cfg.train.lr, use_graphs=True)
    for k, v in trainer.optimizer.defaults.items():
        if k == "betas":
            ecfg.betas = tuple(v)
        elif k == "eps":
            ecfg.eps = float(v)
        elif k == "weight_decay":
            ecfg.weight_decay = float(v)
    engine = SliderEngine(trainer.unet, trainer.network, trainer.scheduler, ecfg, pair0.batch_size, pair0.resolution,
                          pair0.resolution, tuple(emb.shape[1:]), trainer.weight_dtype, trainer.device)
    if world > 1:
        parallel.broadcast_(trainer.network.flat_param)
        torch.manual_seed(parallel.rank_seed(torch.initial_seed(), rank))
    save_path = Path(cfg.save.path)
    current = None
    for i in range(cfg.train.iterations):
        idx = torch.randint(0, len(trainer.prompt_pairs), (1,)).item()
        if world > 1:
            idx = parallel.shard_indices(len(trainer.prompt_pairs), rank, world, i)
        pair = trainer.prompt_pairs[idx]
        if (pair.batch_size, pair.resolution) != (pair0.batch_size, pair0.resolution) or pair.dynamic_resolution:
            raise NotImplementedError("--engine needs one static (batch_size, resolution) for all prompt pairs")
        if idx != current:
            engine.load_pair(pair.target, pair.positive, pair.neutral, pair.unconditional, pair.guidance_scale, pair.action)
            current = idx
        timesteps_to = torch.randint(1, cfg.train.max_denoising_steps, (1,)).item()
        trainer.scheduler.set_timesteps(cfg.train.max_denoising_steps)
        latents = train_util.get_initial_latents(trainer.scheduler, pair.batch_size, pair.resolution, pair.resolution, 1)
        engine.lr = trainer.optimizer.param_groups[0]["lr"]
        loss = engine.train_step(latents, timesteps_to)
        trainer.optimizer.step()           # no grads attached: only advances the schedule bookkeeping
        trainer.lr_scheduler.step()
        if rank == 0 and (i % log_every == 0 or i == cfg.train.iterations - 1):
            print(f"iter {i}: Loss*1k {loss.item() * 1000:.4f} lr {engine.lr:.3e}", flush=True)
        if rank == 0 and i % cfg.save.per_steps == 0 and i != 0 and i != cfg.train.iterations - 1:
            save_path.mkdir(parents=True, exist_ok=True)
            trainer.network.save_weights(save_path / f"{cfg.save.name}_{i}steps{suffix}", dtype=trainer.save_weight_dtype)
    if rank == 0:
        save_path.mkdir(parents=True, exist_ok=True)
        trainer.network.save_weights(save_path / f"{cfg.save.name}_last{suffix}", dtype=trainer.save_weight_dtype)
    return engine
