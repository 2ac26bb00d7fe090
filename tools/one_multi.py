"""Runs the grouped time_emb_proj launch (slider_lora_forward_small_multi) of the SD-1.4 workload a few times."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops

dt = torch.bfloat16
K, M, r = 1280, 2, 4
Ns = [320] * 5 + [640] * 5 + [1280] * 12
layers = []
for N in Ns:
    layers.append(((torch.randn(N, 1, K, device="cuda") / K ** 0.5).to(dt), torch.zeros(N, device="cuda", dtype=dt),
                   (torch.randn(r, 1, K, device="cuda") / K ** 0.5).to(dt), (torch.randn(N, r, device="cuda") * 0.02).to(dt)))
x = torch.randn(M, K, device="cuda").to(dt)
plan = ops.SmallMultiPlan(M, K, dt, x.device, layers)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
for _ in range(4):
    flush.zero_()
    ops.lora_forward_small_multi(x, plan, [0.25] * len(Ns))
torch.cuda.synchronize()
