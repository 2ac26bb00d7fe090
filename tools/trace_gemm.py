"""Per-CTA clock64() timeline of lora_gemm_kernel (slider_debug_gemm_trace) for given shapes."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops, _lib

shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:] if "," in a] or [(1, 1, 8192, 320, 320, 1)]
dt, r, s = torch.bfloat16, 4, 0.25
if "--nosplit" in sys.argv:
    _lib.lib().slider_debug_splitk_depth(1000000)
names = ["setup", "first_tma_issue", "first_full", "last_mma_commit", "acc_visible", "epi_done", "exit",
         "prod_ready", "tmem_in_regs", "chunk0_math_sts", "chunk0_fence", "chunk0_group_bar"]
for (B, H, W, C, N, k) in shapes:
    x = torch.randn(B, H, W, C, device="cuda", dtype=dt)
    w = (torch.randn(N, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
    a = (torch.randn(r, k * k, C, device="cuda") / (C * k * k) ** 0.5).to(dt)
    b = (torch.randn(N, r, device="cuda") * 0.02).to(dt)
    bias = torch.zeros(N, device="cuda", dtype=dt)
    xin = x.reshape(W, C) if (k == 1 and B == 1 and H == 1) else x
    for _ in range(3):
        ops.lora_forward(xin, w, bias, a, b, s, ksize=k)
    trace = torch.zeros(148 * 16, dtype=torch.int64, device="cuda")
    _lib.lib().slider_debug_gemm_trace(trace.data_ptr())
    ops.lora_forward(xin, w, bias, a, b, s, ksize=k)
    torch.cuda.synchronize()
    _lib.lib().slider_debug_gemm_trace(None)
    t = trace.view(148, 16).cpu()
    t = t[t[:, 0] > 0]
    d = (t[:, 1:13] - t[:, :1]).double()
    print(f"shape {(B, H, W, C, N, k)} ctas {t.shape[0]}: cycles since CTA entry (median / max over CTAs)")
    for i, n in enumerate(names):
        col = d[:, i]
        print(f"   {n:18s} {col.median().item():9.0f} {col.max().item():9.0f}")
