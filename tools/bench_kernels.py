"""Micro-benchmark of the C-ABI kernels vs PyTorch eager (cuBLAS/cuDNN) on the shapes of SURVEY §8a.
Dev tool: prints one line per shape. Not the headline bench (see bench.py)."""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.nn.functional as F
from sliders_conceptmod_b200 import ops


def timeit(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1e-3


SHAPES = [
    # B,H,W,C,N,k
    (1, 1, 8192, 320, 320, 1), (1, 1, 2048, 640, 640, 1), (1, 1, 512, 1280, 1280, 1),
    (1, 1, 2048, 1280, 1280, 1), (1, 1, 8192, 640, 640, 1), (1, 1, 65536, 640, 640, 1), (1, 1, 16384, 1280, 1280, 1),
    (2, 64, 64, 320, 320, 3), (2, 32, 32, 640, 640, 3), (2, 16, 16, 1280, 1280, 3), (2, 8, 8, 1280, 1280, 3),
    (1, 1, 2, 1280, 1280, 1), (1, 1, 2, 1280, 320, 1), (2, 8, 8, 2560, 1280, 3),
    (2, 64, 64, 640, 320, 3), (2, 16, 16, 2560, 1280, 3), (16, 128, 128, 320, 320, 3), (16, 64, 64, 640, 640, 3),
]


def main():
    dt = torch.bfloat16
    r, s = 4, 0.25
    for (B, H, W, C, N, k) in SHAPES:
        x = torch.randn(B, H, W, C, device="cuda", dtype=dt)
        w = (torch.randn(N, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
        a = (torch.randn(r, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
        b = (torch.randn(N, r, device="cuda") * 0.02).to(dt)
        bias = torch.zeros(N, device="cuda", dtype=dt)
        xin = x.reshape(W, C) if (k == 1 and B == 1 and H == 1) else x
        M = B * H * W
        flops = 2.0 * M * N * C * k * k + 2.0 * M * r * (C * k * k + N)
        t_f = timeit(lambda: ops.lora_forward(xin, w, bias, a, b, s, ksize=k))
        t_0 = timeit(lambda: ops.lora_forward(xin, w, bias, a, b, 0.0, ksize=k))
        # torch eager reference op sequence (lora.py:134-138) with cuBLAS/cuDNN
        if k == 1:
            x2 = x.reshape(M, C); w2 = w.reshape(N, C); a2 = a.reshape(r, C)
            ref = lambda: F.linear(x2, w2, bias) + F.linear(F.linear(x2, a2), b) * 1.0 * s
            base = lambda: F.linear(x2, w2, bias)
        else:
            xc = x.permute(0, 3, 1, 2)  # channels_last view
            wc = w.reshape(N, k, k, C).permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
            ac = a.reshape(r, k, k, C).permute(0, 3, 1, 2).contiguous(memory_format=torch.channels_last)
            bc = b[:, :, None, None].contiguous()
            ref = lambda: F.conv2d(xc, wc, bias, padding=1) + F.conv2d(F.conv2d(xc, ac, padding=1), bc) * 1.0 * s
            base = lambda: F.conv2d(xc, wc, bias, padding=1)
        t_r = timeit(ref)
        t_b = timeit(base)
        print(json.dumps({"shape": [B, H, W, C, N, k], "fused_us": round(t_f * 1e6, 1), "fused_off_us": round(t_0 * 1e6, 1),
                          "torch_lora_us": round(t_r * 1e6, 1), "torch_base_us": round(t_b * 1e6, 1),
                          "fused_tflops": round(flops / t_f / 1e12, 1), "torch_base_tflops": round(2.0 * M * N * C * k * k / t_b / 1e12, 1)}))


if __name__ == "__main__":
    main()
