"""ncu target: a few launches of the fused base+LoRA GEMM at given shapes (B,H,W,Cin,Cout,k per arg, comma separated).
  python tools/ncu_gemm.py 1,1,8192,320,320,1 2,64,64,320,320,3
"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops

shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:] if "," in a] or [(1, 1, 8192, 320, 320, 1)]
dt, r, s = torch.bfloat16, 4, 0.25
for (B, H, W, C, N, k) in shapes:
    x = torch.randn(B, H, W, C, device="cuda", dtype=dt)
    w = (torch.randn(N, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
    a = (torch.randn(r, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
    b = (torch.randn(N, r, device="cuda") * 0.02).to(dt)
    bias = torch.zeros(N, device="cuda", dtype=dt)
    xin = x.reshape(W, C) if (k == 1 and B == 1 and H == 1) else x
    for _ in range(3):
        y, _ = ops.lora_forward(xin, w, bias, a, b, s, ksize=k)
    torch.cuda.synchronize()
print("done")
