"""Worker of tests/test_gpu_dp2.py: one process per GPU over NCCL, K SliderEngine steps on per-rank samples."""
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def samples(cfg, world: int, steps: int, dtype):
    """Deterministic per-(step, rank) inputs: prompt-pair embeddings (rank k trains pair k), initial noise, timesteps_to."""
    g = torch.Generator().manual_seed(77)
    pairs = [{k: (4.0 * torch.randn(1, 77, cfg.cross_attention_dim, generator=g)).to(dtype).float()
              for k in ("target", "positive", "neutral", "unconditional")} for _ in range(world)]
    lats = [[torch.randn(1, 4, 16, 16, generator=g).to(dtype).float() for _ in range(world)] for _ in range(steps)]
    t_to = [[int(torch.randint(1, 5, (1,), generator=g)) for _ in range(world)] for _ in range(steps)]
    return pairs, lats, t_to


def build_engine(dtype, device):
    from helpers import build_pair
    from sliders_conceptmod_b200 import model_util
    from sliders_conceptmod_b200.engine import EngineConfig, SliderEngine
    from sliders_conceptmod_b200.unet import UNetConfig
    cfg = UNetConfig.tiny()
    _, _, unet_g, net_g = build_pair(cfg, device, dtype)
    eng = SliderEngine(unet_g, net_g, model_util.create_noise_scheduler("ddim"),
                       EngineConfig(max_denoising_steps=10, use_graphs=True, lr=1e-3), 1, 128, 128,
                       (77, cfg.cross_attention_dim), dtype, device)
    return cfg, eng, net_g


def run(rank: int, world: int, port: int, steps: int, out_path: str):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    dtype = torch.bfloat16
    cfg, eng, net = build_engine(dtype, dev)
    assert eng.world == world
    pairs, lats, t_to = samples(cfg, world, steps, dtype)
    e = pairs[rank]
    eng.load_pair(e["target"], e["positive"], e["neutral"], e["unconditional"], 4.0, "enhance")
    losses = []
    for s in range(steps):
        losses.append(float(eng.train_step(lats[s][rank], t_to[s][rank])))
    torch.cuda.synchronize()
    flat = net.flat_param.float()
    hi, lo = flat.clone(), flat.clone()
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    spread = float((hi - lo).abs().max())       # replicas stay bit-identical: same reduced gradient, same update
    if rank == 0:
        torch.save({"flat_param": flat.cpu(), "losses": losses, "replica_spread": spread}, out_path)
    dist.barrier()
    dist.destroy_process_group()
