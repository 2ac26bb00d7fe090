"""ncu target: the low-rank backward (slider_lora_grad) at given shapes (B,H,W,Cin,Cout,k)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops

shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:] if "," in a] or [(2, 16, 16, 1280, 1280, 3)]
dt, r = torch.bfloat16, 4
for (B, H, W, C, N, k) in shapes:
    M = B * H * W
    shape = (M, C) if k == 1 and B == 1 and H == 1 else (B, H, W, C)
    x = torch.randn(*shape, device="cuda").to(dt)
    gy = torch.randn(*shape[:-1], N, device="cuda").to(dt)
    h = torch.randn(M, r, device="cuda")
    b = (torch.randn(N, r, device="cuda") * 0.02).to(dt)
    dA = torch.zeros(r, k * k, C, device="cuda")
    dB = torch.zeros(N, r, device="cuda")
    for _ in range(2):
        ops.lora_grad(gy, x, h, b, 0.25, k, dA, dB)
    torch.cuda.synchronize()
print("done")
