import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops
dt = torch.bfloat16
def bench(M, K, N, nsets, lora=True):
    xs = torch.randn(M, K, device="cuda", dtype=dt)
    Ws = [torch.randn(N, 1, K, device="cuda", dtype=dt) * 0.02 for _ in range(nsets)]
    a = torch.randn(4, 1, K, device="cuda", dtype=dt) * 0.02
    b = torch.randn(N, 4, device="cuda", dtype=dt) * 0.02
    bias = torch.zeros(N, device="cuda", dtype=dt)
    s = 0.25 if lora else 0.0
    for w in Ws[:2]: ops.lora_forward(xs, w, bias, a, b, s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for w in Ws: ops.lora_forward(xs, w, bias, a, b, s)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (5 * nsets)
for (M, K, N) in [(2, 1280, 1280), (2, 1280, 320), (2, 1280, 640), (16, 320, 640)]:
    print((M, K, N), "cold weights (100 sets):", round(bench(M, K, N, 100), 2), "us; warm (2 sets):", round(bench(M, K, N, 2), 2), "us; warm no-lora:", round(bench(M, K, N, 2, False), 2))
