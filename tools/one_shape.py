"""Calls the fused forward on the given shapes a few times (target for `ncu --metrics gpu__time_duration.sum`)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops

shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:] if "," in a]
dt, r, s = torch.bfloat16, 4, 0.25
for (B, H, W, C, N, k) in shapes:
    x = torch.randn(B, H, W, C, device="cuda", dtype=dt)
    w = (torch.randn(N, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
    a = (torch.randn(r, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
    b = (torch.randn(N, r, device="cuda") * 0.02).to(dt)
    bias = torch.zeros(N, device="cuda", dtype=dt)
    xin = x.reshape(W, C) if (k == 1 and B == 1 and H == 1) else x
    for _ in range(4):
        ops.lora_forward(xin, w, bias, a, b, s, ksize=k)
    torch.cuda.synchronize()
