"""Data-parallel equivalence ON HARDWARE: 2 ranks x 1 sample per step over NCCL, one all-reduce of the flat
LoRA-gradient buffer, == ONE process accumulating the same 2 samples per step (mean) before the same AdamW update.
The reference has no distributed code; this is the definition of equivalence its gradient accumulation implies
(conceptmod/textsliders/train_lora.py:298-300 applied to N samples).  Needs 2 GPUs: `gpurun --gpus 2 -- python -m pytest
tests/test_gpu_dp2.py -m gpu -s` (skips on a 1-GPU box; the CPU/gloo form of the same statement is
tests/test_distributed_gloo.py)."""
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_two_ranks_equal_gradient_accumulation(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    import dp_worker
    from helpers import cosine
    steps, world = 3, 2
    out = str(tmp_path / "dp2.pt")
    mp.spawn(dp_worker.run, args=(world, _free_port(), steps, out), nprocs=world, join=True)
    got = torch.load(out)
    # single process: the same two samples per step, gradients accumulated then averaged, one AdamW launch
    dtype = torch.bfloat16
    cfg, eng, net = dp_worker.build_engine(dtype, "cuda:0")
    p0 = net.flat_param.float().cpu().clone()
    pairs, lats, t_to = dp_worker.samples(cfg, world, steps, dtype)
    losses = []
    for s in range(steps):
        acc = torch.zeros_like(net.flat_grad)
        for r in range(world):
            e = pairs[r]
            eng.load_pair(e["target"], e["positive"], e["neutral"], e["unconditional"], 4.0, "enhance")
            loss = eng.train_step(lats[s][r], t_to[s][r], optimizer=False)
            if r == 0:
                losses.append(float(loss))
            acc += net.flat_grad
        net.flat_grad.copy_(acc / world)
        eng.optimizer_step()
    torch.cuda.synchronize()
    ref = net.flat_param.float().cpu()
    cw = cosine(got["flat_param"], ref)
    cd = cosine(got["flat_param"] - p0, ref - p0)
    mx = (got["flat_param"] - ref).abs().max().item()
    print(f"2 ranks over NCCL vs 1 process accumulating 2 samples, {steps} steps (bf16): weight cosine {cw:.7f}, "
          f"update cosine {cd:.6f}, max-abs {mx:.3e}; rank-0 losses {got['losses']} vs {losses}")
    assert cw >= 0.999 and cd >= 0.99
    assert got["replica_spread"] == 0.0, "ranks diverged"
    # step 0 runs on identical weights: rank 0's loss equals the accumulating process's first-sample loss
    assert abs(got["losses"][0] - losses[0]) <= 1e-3 * abs(losses[0])
