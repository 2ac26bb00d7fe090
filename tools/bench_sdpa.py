"""SDPA backend timing at the UNet's attention shapes (B=2 CFG batch, SD-1.4): which library kernel to ask for."""
import sys
import torch
import torch.nn.functional as F
from torch.nn.attention import SDPBackend, sdpa_kernel

shapes = [(2, 8, 4096, 4096, 40), (2, 8, 1024, 1024, 80), (2, 8, 256, 256, 160), (2, 8, 64, 64, 160),
          (2, 8, 4096, 77, 40), (2, 8, 1024, 77, 80), (2, 8, 256, 77, 160), (2, 8, 64, 77, 160),
          (6, 8, 4096, 4096, 40)]
for (B, H, L, S, D) in shapes:
    q = torch.randn(B, L, H, D, device="cuda", dtype=torch.bfloat16).transpose(1, 2)
    k = torch.randn(B, S, H, D, device="cuda", dtype=torch.bfloat16).transpose(1, 2)
    v = torch.randn(B, S, H, D, device="cuda", dtype=torch.bfloat16).transpose(1, 2)
    res = {}
    for name, be in (("default", None), ("cudnn", SDPBackend.CUDNN_ATTENTION), ("flash", SDPBackend.FLASH_ATTENTION),
                     ("efficient", SDPBackend.EFFICIENT_ATTENTION)):
        try:
            def run():
                if be is None:
                    return F.scaled_dot_product_attention(q, k, v)
                with sdpa_kernel(be):
                    return F.scaled_dot_product_attention(q, k, v)
            with torch.no_grad():
                for _ in range(3):
                    run()
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    for _ in range(10):
                        run()
                g.replay(); torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(5):
                    g.replay()
                e1.record(); torch.cuda.synchronize()
                res[name] = round(e0.elapsed_time(e1) * 1e3 / 50, 1)
        except Exception as e:
            res[name] = "n/a: " + str(e)[:40]
    print((B, H, L, S, D), res, flush=True)
