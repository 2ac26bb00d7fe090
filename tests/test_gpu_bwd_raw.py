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


ONE_BY_ONE = [(1, 1, 8192, 320, 320, 4), (1, 1, 2048, 640, 640, 4), (1, 1, 512, 1280, 1280, 4), (1, 1, 300, 64, 72, 4),
              (2, 32, 32, 320, 640, 16), (1, 1, 128, 1280, 320, 8), (1, 1, 6, 320, 1280, 4)]


@pytest.mark.parametrize("case", ONE_BY_ONE)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_fused_1x1_backward_matches_autograd(case, dtype):
    """slider_lora_backward_1x1: dx and g = gy.B from ONE tcgen05 launch, dA and dB from one reduction launch;
    against fp32 autograd of y = x W^T + s (x A^T) B^T, with and without a pre-packed B^T, frozen-adapter form,
    accumulate flag, and bit-identical gradients on a second run (deterministic reductions)."""
    from sliders_conceptmod_b200 import ops
    B, H, W, C, N, r = case
    dev = "cuda"
    gen = torch.Generator(device=dev).manual_seed(99)
    lin = B == 1 and H == 1
    shp = (W,) if lin else (B, H, W)
    x = torch.randn(*shp, C, device=dev, generator=gen).to(dtype)
    w = ((torch.rand(N, 1, C, device=dev, generator=gen) * 2 - 1) / C ** 0.5).to(dtype)
    a = ((torch.rand(r, 1, C, device=dev, generator=gen) * 2 - 1) / C ** 0.5).to(dtype)
    b = (torch.randn(N, r, device=dev, generator=gen) * 0.2).to(dtype)
    gy = torch.randn(*shp, N, device=dev, generator=gen).to(dtype)
    s = 0.5
    xf = x.float().reshape(-1, C).requires_grad_(True)
    af = a.float()[:, 0].requires_grad_(True)
    bf = b.float().requires_grad_(True)
    y = xf @ w.float()[:, 0].t() + s * (xf @ af.t()) @ bf.t()
    y.backward(gy.float().reshape(-1, N))
    _, h = ops.lora_forward(x, w, None, a, b, s, want_h=True)
    wt = ops.pack_wt(w, 1)
    dA = torch.empty(r, 1, C, device=dev)
    dB = torch.empty(N, r, device=dev)
    dx = ops.lora_backward_1x1(gy, x, h, wt, a, b, s, dA, dB)
    torch.cuda.synchronize()
    tol = 3e-2 if dtype == torch.bfloat16 else 2e-4

    def close(got, ref, name):
        err = (got.float().reshape(ref.shape) - ref).abs().max().item()
        sc = ref.abs().max().item()
        cos = F.cosine_similarity(got.float().flatten(), ref.flatten(), dim=0).item()
        assert err <= tol * sc and cos > 0.999, f"{name}: err {err} scale {sc} cos {cos}"

    close(dx, xf.grad, "dx")
    close(dA[:, 0], af.grad, "dA")
    close(dB, bf.grad, "dB")
    # pre-packed B^T (what LoRANetwork.pack_transposes provides): same bits
    dA2, dB2 = torch.empty_like(dA), torch.empty_like(dB)
    dx2 = ops.lora_backward_1x1(gy, x, h, wt, a, b, s, dA2, dB2, bt=b.t().contiguous())
    assert torch.equal(dx2, dx) and torch.equal(dA2, dA) and torch.equal(dB2, dB)
    # accumulate adds to the current content
    ops.lora_backward_1x1(gy, x, h, wt, a, b, s, dA2, dB2, accumulate=True)
    close(dA2[:, 0], 2 * af.grad, "dA accumulate")
    close(dB2, 2 * bf.grad, "dB accumulate")
    # frozen adapter: dx only (still with the s*dh.A term)
    dx3 = ops.lora_backward_1x1(gy, None, None, wt, a, b, s, None, None)
    assert torch.equal(dx3, dx)
    # the generic entry points agree
    dA4, dB4 = torch.empty_like(dA), torch.empty_like(dB)
    dh = ops.lora_grad(gy, x, h, b, s, 1, dA4, dB4)
    dx4 = ops.lora_input_grad(gy, wt, dh, ops.pack_at(a, 1), C, 1, r)
    close(dx4, xf.grad, "dx generic")
    close(dA4[:, 0], af.grad, "dA generic")
    close(dB4, bf.grad, "dB generic")


@pytest.mark.parametrize("shape", [(2, 64, 64, 320, 320), (2, 16, 16, 1280, 1280), (1, 17, 23, 64, 96)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_stride2_backward_matches_autograd(shape, dtype):
    """Low-rank gradients of a stride-2 3x3 adapter layer: gy and h live on the OUTPUT grid, x on the input grid."""
    from sliders_conceptmod_b200 import ops
    B, H, W, C, N = shape
    r = 4
    dev = "cuda"
    gen = torch.Generator(device=dev).manual_seed(5)
    x = torch.randn(B, H, W, C, device=dev, generator=gen).to(dtype)
    a = ((torch.rand(r, 9, C, device=dev, generator=gen) * 2 - 1) / (9 * C) ** 0.5).to(dtype)
    w = ((torch.rand(N, 9, C, device=dev, generator=gen) * 2 - 1) / (9 * C) ** 0.5).to(dtype)
    b = (torch.randn(N, r, device=dev, generator=gen) * 0.2).to(dtype)
    Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    gy = torch.randn(B, Ho, Wo, N, device=dev, generator=gen).to(dtype)
    s = 0.5
    xf = x.float().permute(0, 3, 1, 2)
    af = a.float().requires_grad_(True)
    bf = b.float().requires_grad_(True)
    hh = F.conv2d(xf, af.reshape(r, 3, 3, C).permute(0, 3, 1, 2), stride=2, padding=1)
    y = s * F.conv2d(hh, bf[:, :, None, None])
    y.backward(gy.float().permute(0, 3, 1, 2))
    _, h = ops.lora_forward(x, w, None, a, b, s, ksize=3, want_h=True, stride=2)
    dA = torch.empty(r, 9, C, device=dev)
    dB = torch.empty(N, r, device=dev)
    dh = ops.lora_grad(gy, x, h, b, s, 3, dA, dB, stride=2)
    torch.cuda.synchronize()
    tol = 3e-2 if dtype == torch.bfloat16 else 2e-4
    for got, ref, name in ((dA, af.grad, "dA"), (dB, bf.grad, "dB")):
        err = (got - ref).abs().max().item()
        cos = F.cosine_similarity(got.flatten(), ref.flatten(), dim=0).item()
        assert err <= tol * ref.abs().max().item() and cos > 0.999, f"{name}: err {err} cos {cos}"
    assert dh.shape == (B * Ho * Wo, 16)
    dA1, dB1 = torch.empty_like(dA), torch.empty_like(dB)
    ops.lora_grad(gy, x, h, b, s, 3, dA1, dB1, stride=2)
    assert torch.equal(dA1, dA) and torch.equal(dB1, dB), "LoRA gradients must be bit-reproducible"
