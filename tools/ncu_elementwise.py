"""ncu target: the bandwidth-class kernels of the slider path, each launched twice at n = 2^26 (HBM regime, distinct
buffers per launch) and twice at the SD-1.4 native size (16 384-element latents; 2.9 M LoRA parameters for the
optimizer): guidance_loss_kernel, cfg_step_kernel, adamw_flat_kernel, clip_coef_kernel, tc_colreduce_kernel (one 1x1 and
one 3x3 layer through slider_lora_backward_1x1 / slider_lora_grad), gn_group_kernel / gn_bwd_group_kernel.  Run plain first, then under
  ncu --set full --clock-control none -k regex:'guidance_loss|cfg_step|adamw|clip_coef|colreduce|gn_group|gn_bwd' ..."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops

dt = torch.bfloat16
for n in (1 << 26, 16384):
    for rep in range(2):
        t, p, nn, q = (torch.randn(n, device="cuda").to(dt) for _ in range(4))
        ops.guidance_loss(t, p, nn, q, 4.0, 1)
        x, eu, ec = (torch.randn(n, device="cuda").to(dt) for _ in range(3))
        ops.cfg_step(x, eu, ec, None, guidance=3.0, c_x=0.9, c_eps=-0.1)
for n in (1 << 26, 2906880):
    for rep in range(2):
        prm = torch.randn(n, device="cuda").to(dt)
        g, m, v = (torch.randn(n, device="cuda") for _ in range(3))
        v.abs_()
        ops.adamw_flat(prm, g, m, v, 1e-4, 0.9, 0.999, 1e-8, 1e-2, 3)
        ops.clip_coef(g, 0.2)
# reductions of the low-rank backward: level-0 1x1 layer (fused path) and level-0 3x3 convolution
r = 4
for rep in range(2):
    M, C, N = 8192, 320, 320
    x = torch.randn(M, C, device="cuda").to(dt); gy = torch.randn(M, N, device="cuda").to(dt)
    w = (torch.randn(N, 1, C, device="cuda") * 0.05).to(dt); a = (torch.randn(r, 1, C, device="cuda") * 0.05).to(dt)
    b = (torch.randn(N, r, device="cuda") * 0.05).to(dt); h = torch.randn(M, r, device="cuda")
    dA = torch.zeros(r, 1, C, device="cuda"); dB = torch.zeros(N, r, device="cuda")
    ops.lora_backward_1x1(gy, x, h, ops.pack_wt(w, 1), a, b, 0.25, dA, dB)
    x4 = torch.randn(2, 64, 64, C, device="cuda").to(dt); gy4 = torch.randn(2, 64, 64, N, device="cuda").to(dt)
    dA9 = torch.zeros(r, 9, C, device="cuda")
    ops.lora_grad(gy4, x4, h, b, 0.25, 3, dA9, dB)
# GroupNorm (+SiLU, + per-channel pre-add) training forward and its single-launch backward at the level-0 size
for rep in range(2):
    xg = torch.randn(2, 64, 64, 320, device="cuda").to(dt); gyg = torch.randn(2, 64, 64, 320, device="cuda").to(dt)
    gam = torch.randn(320, device="cuda").to(dt); bet = torch.randn(320, device="cuda").to(dt)
    add = torch.randn(2, 320, device="cuda").to(dt)
    yg, st = ops.groupnorm_nhwc(xg, 32, gam, bet, 1e-5, True, add, want_stats=True)
    ops.groupnorm_nhwc_backward(gyg, xg, st, 32, gam, bet, True, add, want_dadd=True)
torch.cuda.synchronize()
print("done")
