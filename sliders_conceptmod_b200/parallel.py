"""Data-parallel plumbing: one process per GPU, independent (prompt pair, timestep, noise) samples per
rank, ONE all-reduce of the flat LoRA-gradient buffer per optimizer step.  The reference
has no distributed code; equivalence is defined as: N ranks x 1 sample with mean-reduced gradients ==
the reference iterated over the same N samples with gradient accumulation, divided by N."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def init_from_env(backend: str | None = None) -> tuple[int, int, int]:
    """Reads RANK / WORLD_SIZE / LOCAL_RANK / MASTER_* (torchrun). Returns (rank, world, local_rank)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, world, local


def shard_indices(n_items: int, rank: int, world: int, step: int = 0) -> int:
    """Which prompt pair rank `rank` trains on at optimizer step `step`: round-robin over pairs, rotated
    every step so every rank sees every pair."""
    return (rank + step * world) % n_items


def rank_seed(base_seed: int, rank: int) -> int:
    return base_seed + rank


def allreduce_mean_(flat_grad: torch.Tensor) -> torch.Tensor:
    """In-place mean over ranks of the flat gradient buffer (no-op for a single process)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM)
        flat_grad.div_(dist.get_world_size())
    return flat_grad


def broadcast_(flat_param: torch.Tensor, src: int = 0) -> torch.Tensor:
    """Make every rank start from rank `src`'s LoRA factors."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(flat_param, src=src)
    return flat_param
