"""Timing probe: is lora_gemm_kernel's main loop bound by L2 -> SM operand traffic?  Runs big convolutions with the
producer's A loads and / or W loads switched off (SLIDER_DBG_SKIP bit 0 / bit 1, results are garbage) and prints
the in-graph time per launch.  Dev tool."""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops

SHAPES = [(2, 64, 64, 640, 640, 3), (2, 32, 32, 1280, 1280, 3), (2, 64, 64, 320, 320, 3), (2, 32, 32, 640, 640, 3),
          (16, 128, 128, 320, 320, 3), (1, 1, 8192, 320, 960, 1)]
dt, r, s = torch.bfloat16, 4, 0.25
for (B, H, W, C, N, k) in SHAPES:
    x = torch.randn(B, H, W, C, device="cuda", dtype=dt)
    w = (torch.randn(N, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
    a = (torch.randn(r, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
    b = (torch.randn(N, r, device="cuda") * 0.02).to(dt)
    xin = x.reshape(-1, C) if k == 1 else x
    reps = 10

    def run():
        for _ in range(reps):
            ops.lora_forward(xin, w, None, a, b, s, ksize=k)
    run(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        run()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (3 * reps)
    fl = 2.0 * B * H * W * N * k * k * C
    print(f"skip={os.environ.get('SLIDER_DBG_SKIP', '0')} shape={(B, H, W, C, N, k)} us={us:.2f} tflops={fl / us / 1e6:.0f}", flush=True)
