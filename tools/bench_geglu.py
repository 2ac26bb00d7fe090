"""Feed-forward projection + GEGLU gate at the SD-1.4 shapes: the fused tcgen05 launch (slider_linear_geglu) against
cuBLAS + slider_geglu, each timed as the average over a CUDA graph of 16 launches on rotating buffers (inputs and
weights cycle through > 126 MB so that every launch sees cold operands, as in the real forward)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.nn.functional as F
from sliders_conceptmod_b200 import ops

dt = torch.bfloat16
ROT = 16


def timed(fn, reps=5):
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(ROT):
            fn(i)
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for i in range(ROT):
                fn(i)
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * ROT)


for (M, K, D) in [(8192, 320, 1280), (2048, 640, 2560), (512, 1280, 5120), (128, 1280, 5120)]:
    xs = [torch.randn(M, K, device="cuda", dtype=dt) for _ in range(ROT)]
    ws = [(torch.randn(2 * D, K, device="cuda") / K ** 0.5).to(dt) for _ in range(ROT)]
    bs = [torch.randn(2 * D, device="cuda", dtype=dt) * 0.1 for _ in range(ROT)]
    idx = ops.geglu_permutation(D, xs[0].device)
    wp = [w.index_select(0, idx).contiguous() for w in ws]
    bp = [b.index_select(0, idx).contiguous() for b in bs]
    t_f = timed(lambda i: ops.linear_geglu(xs[i], wp[i], bp[i]))
    t_c = timed(lambda i: ops.geglu(F.linear(xs[i], ws[i], bs[i])))
    t_g = timed(lambda i: F.linear(xs[i], ws[i], bs[i]))
    y0 = ops.linear_geglu(xs[0], wp[0], bp[0]).float()
    h = F.linear(xs[0].float(), ws[0].float(), bs[0].float())
    ref = h[:, :D] * F.gelu(h[:, D:])
    err = ((y0 - ref).abs().max() / ref.abs().max()).item()
    fl = 2.0 * M * K * 2 * D
    print(f"M={M:5d} K={K:5d} d={D:5d}: fused {t_f:6.2f} us ({fl / t_f * 1e-6:6.1f} TF/s)  cuBLAS+geglu {t_c:6.2f} us  "
          f"(cuBLAS alone {t_g:6.2f})  max rel err {err:.2e}")
