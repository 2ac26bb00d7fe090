"""Graph-timed latency of the fused GEMM on tile-starved (split-K) shapes: 16 launches over rotating buffers
captured in one CUDA graph, CUDA-event time per launch."""
import json
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops, _lib
import ctypes
DEPTH = int(sys.argv[1]) if len(sys.argv) > 1 else 6
_lib.lib().slider_debug_splitk_depth.argtypes = [ctypes.c_int]
_lib.lib().slider_debug_splitk_depth(DEPTH)

SHAPES = [(2, 16, 16, 1280, 1280, 3), (2, 8, 8, 1280, 1280, 3), (2, 8, 8, 2560, 1280, 3), (2, 16, 16, 2560, 1280, 3),
          (2, 16, 16, 1920, 1280, 3), (1, 1, 512, 1280, 1280, 1), (1, 1, 512, 5120, 1280, 1), (2, 32, 32, 640, 640, 3),
          (2, 32, 32, 1280, 640, 3), (1, 1, 2048, 640, 640, 1), (2, 64, 64, 320, 320, 3)]
dt, r, s, NB = torch.bfloat16, 4, 0.25, 8
for (B, H, W, C, N, k) in SHAPES:
    xs = [torch.randn(B, H, W, C, device="cuda", dtype=dt) for _ in range(NB)]
    ws = [(torch.randn(N, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt) for _ in range(NB)]
    a = (torch.randn(r, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
    b = (torch.randn(N, r, device="cuda") * 0.02).to(dt)
    bias = torch.zeros(N, device="cuda", dtype=dt)
    fix = lambda x: x.reshape(W, C) if (k == 1 and B == 1 and H == 1) else x
    run = lambda: [ops.lora_forward(fix(xs[i % NB]), ws[i % NB], bias, a, b, s, ksize=k) for i in range(16)]
    run(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        run()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 160
    flops = 2.0 * B * H * W * N * C * k * k
    print(json.dumps({"depth": DEPTH, "shape": [B, H, W, C, N, k], "us": round(us, 2), "tflops": round(flops / us / 1e6, 1)}), flush=True)
