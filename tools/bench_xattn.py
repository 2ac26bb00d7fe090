"""slider_cross_attention against the library SDPA routes the UNet used before, at the attn2 shapes of SD-1.4 (batch 2,
8 heads, 77 keys): each timed as the average over a CUDA graph of 16 launches on rotating buffers."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops, unet

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


CASES = [(2, 4096, 77, 8, 40), (2, 1024, 77, 8, 80), (2, 256, 77, 8, 160), (2, 64, 77, 8, 160), (6, 4096, 77, 8, 40),
         (2, 4096, 77, 10, 64), (2, 64, 64, 8, 160)]
for (B, L, S, H, d) in CASES:
    qs = [torch.randn(B, L, H * d, device="cuda").bfloat16() for _ in range(ROT)]
    ks = [torch.randn(B, S, H * d, device="cuda").bfloat16() for _ in range(ROT)]
    vs = [torch.randn(B, S, H * d, device="cuda").bfloat16() for _ in range(ROT)]

    def lib(i):
        q, k, v = (t.view(B, -1, H, d).transpose(1, 2) for t in (qs[i], ks[i], vs[i]))
        return unet._sdpa(q, k, v).transpose(1, 2).reshape(B, L, H * d)
    with torch.no_grad():
        t_o = timed(lambda i: ops.attention(qs[i], ks[i], vs[i], H, kv_static=S <= 80))
        t_l = timed(lib)
    print(f"B={B} L={L:5d} S={S:5d} heads={H:2d} d={d:3d}: libslider {t_o:7.2f} us   library SDPA route (+ output transpose copy) {t_l:7.2f} us")
