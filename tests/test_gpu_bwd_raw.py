"""GPU parity of the backward entry points (lora_grad, pack_*, lora_input_grad) against PyTorch
autograd of the same op in fp32."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

CASES = [
    (1, 1, 256, 128, 160, 1, 4),
    (1, 1, 2048, 640, 640, 1, 4),
    (2, 16, 16, 64, 160, 3, 4),
    (2, 8, 8, 128, 320, 3, 4),
    (1, 24, 40, 64, 64, 3, 8),
    (2, 32, 32, 320, 640, 1, 16),
    (2, 8, 8, 1280, 640, 3, 4),      # split-K in the extended-K input-gradient GEMM
    (1, 1, 256, 1280, 1280, 1, 4),
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_backward_matches_autograd(case, dtype):
    from sliders_conceptmod_b200 import ops
    B, H, W, C, N, k, r = case
    dev = "cuda"
    gen = torch.Generator(device=dev).manual_seed(4321)
    x = torch.randn(B, H, W, C, device=dev, generator=gen).to(dtype)
    w = ((torch.rand(N, k * k, C, device=dev, generator=gen) * 2 - 1) / (C * k * k) ** 0.5).to(dtype)
    a = ((torch.rand(r, k * k, C, device=dev, generator=gen) * 2 - 1) / (C * k * k) ** 0.5).to(dtype)
    b = (torch.randn(N, r, device=dev, generator=gen) * 0.2).to(dtype)
    gy = torch.randn(B, H, W, N, device=dev, generator=gen).to(dtype)
    s = 0.5
    # fp32 autograd reference
    xf = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    af = a.float().requires_grad_(True)
    bf = b.float().requires_grad_(True)
    wc = w.float().reshape(N, k, k, C).permute(0, 3, 1, 2)
    ac = af.reshape(r, k, k, C).permute(0, 3, 1, 2)
    pad = (k - 1) // 2
    hh = F.conv2d(xf, ac, padding=pad)
    y = F.conv2d(xf, wc, padding=pad) + s * F.conv2d(hh, bf[:, :, None, None])
    y.backward(gy.float().permute(0, 3, 1, 2))
    dx_ref = xf.grad.permute(0, 2, 3, 1)
    # CUDA path
    _, h = ops.lora_forward(x, w, None, a, b, s, ksize=k, want_h=True)
    dA = torch.empty(r, k * k, C, device=dev)
    dB = torch.empty(N, r, device=dev)
    dh = ops.lora_grad(gy, x, h, b, s, k, dA, dB)
    wt = ops.pack_wt(w, k)
    at = ops.pack_at(a, k)
    dx = ops.lora_input_grad(gy, wt, dh, at, C, k, r)
    torch.cuda.synchronize()
    tol = 3e-2 if dtype == torch.bfloat16 else 2e-4

    def close(got, ref, name):
        err = (got.float() - ref).abs().max().item()
        sc = ref.abs().max().item()
        cos = F.cosine_similarity(got.float().flatten(), ref.flatten(), dim=0).item()
        assert err <= tol * sc and cos > 0.999, f"{name}: err {err} scale {sc} cos {cos}"

    close(dB, bf.grad, "dB")
    close(dA, af.grad, "dA")
    close(dx, dx_ref, "dx")
    # accumulate flag doubles the result
    ops.lora_grad(gy, x, h, b, s, k, dA, dB, accumulate=True)
    close(dB, 2 * bf.grad, "dB accumulate")
    # base-only input gradient
    dx0 = ops.lora_input_grad(gy, wt, None, None, C, k, 0)
    y0 = F.conv2d(xf, wc, padding=pad)
    g0, = torch.autograd.grad(y0, xf, gy.float().permute(0, 3, 1, 2))
    close(dx0, g0.permute(0, 2, 3, 1), "dx base")
