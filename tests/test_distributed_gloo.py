"""world_size-2 gloo test of the data-parallel plumbing: N ranks x 1 sample with the mean-reduced flat
gradient == the (oracle) reference iterated over the same N samples with gradient accumulation / N."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import ROOT, build_pair  # noqa: F401


def _flat_grad(net):
    return torch.cat([p.grad.flatten() for m in net.unet_loras for p in (m.lora_down.weight, m.lora_up.weight)])


def _sample(rank, cfg):
    g = torch.Generator().manual_seed(100 + rank)
    emb = {k: torch.randn(1, 77, cfg.cross_attention_dim, generator=g) for k in ("target", "positive", "neutral", "unconditional")}
    return emb, torch.randn(1, 4, 16, 16, generator=g), 1 + rank


def _worker(rank, world, port, out):
    sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from oracle import slider_oracle
    from sliders_conceptmod_b200 import parallel
    from sliders_conceptmod_b200.unet import UNetConfig
    torch.set_num_threads(2)
    r, w, _ = parallel.init_from_env("gloo")
    assert (r, w) == (rank, world)
    cfg = UNetConfig.tiny()
    unet, net, _, _ = build_pair(cfg, "cpu", torch.float32)
    flat = torch.cat([p.data.flatten() for p in net.parameters()])
    if rank == 1:
        flat += 1.0
    parallel.broadcast_(flat)                    # every rank starts from rank 0's factors
    emb, lat, steps = _sample(parallel.shard_indices(2, rank, world), cfg)
    slider_oracle.train_iteration(unet, net, slider_oracle.OracleScheduler("ddim"), emb, lat, steps, 4.0, "enhance", max_denoising_steps=10)
    g = parallel.allreduce_mean_(_flat_grad(net))
    if rank == 0:
        torch.save({"grad": g, "flat0": flat}, out)
    dist.destroy_process_group()


def test_two_ranks_equal_gradient_accumulation(tmp_path):
    out = tmp_path / "g.pt"
    mp.spawn(_worker, args=(2, 29731, str(out)), nprocs=2, join=True)
    got = torch.load(out)
    from oracle import slider_oracle
    from sliders_conceptmod_b200.unet import UNetConfig
    cfg = UNetConfig.tiny()
    unet, net, _, _ = build_pair(cfg, "cpu", torch.float32)
    assert torch.equal(got["flat0"], torch.cat([p.data.flatten() for p in net.parameters()]))
    for rank in range(2):                       # accumulate both samples in one process (no zero_grad between)
        emb, lat, steps = _sample(rank, cfg)
        slider_oracle.train_iteration(unet, net, slider_oracle.OracleScheduler("ddim"), emb, lat, steps, 4.0, "enhance", max_denoising_steps=10)
    ref = _flat_grad(net) / 2
    assert torch.allclose(got["grad"], ref, rtol=1e-4, atol=1e-7)


def test_shard_indices_cover_all_pairs():
    from sliders_conceptmod_b200 import parallel
    seen = {parallel.shard_indices(4, r, 8, s) for r in range(8) for s in range(2)}
    assert seen == {0, 1, 2, 3}
    assert parallel.rank_seed(1234, 3) == 1237
