"""GPU parity of the row-wise transformer-block kernels (LayerNorm, GEGLU) against PyTorch fp32."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,C", [(8192, 320), (2048, 640), (512, 1280), (77, 64), (3, 2560)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_layernorm(rows, C, dtype):
    from sliders_conceptmod_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    x = (torch.randn(rows, C, device="cuda", generator=g) * 2 + 0.5).to(dtype)
    w = (1 + 0.1 * torch.randn(C, device="cuda", generator=g)).to(dtype)
    b = (0.1 * torch.randn(C, device="cuda", generator=g)).to(dtype)
    y = ops.layernorm(x, w, b, 1e-5)
    yr = F.layer_norm(x.float(), (C,), w.float(), b.float(), 1e-5)
    tol = 3e-2 if dtype == torch.bfloat16 else 2e-5
    assert (y.float() - yr).abs().max().item() <= tol * max(1.0, yr.abs().max().item())


@pytest.mark.parametrize("rows,D", [(8192, 1280), (2048, 2560), (512, 5120), (5, 64)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_geglu(rows, D, dtype):
    from sliders_conceptmod_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(4)
    x = torch.randn(rows, 2 * D, device="cuda", generator=g).to(dtype)
    y = ops.geglu(x)
    a, gate = x.float().chunk(2, dim=-1)
    yr = a * F.gelu(gate)
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-5
    assert (y.float() - yr).abs().max().item() <= tol * max(1.0, yr.abs().max().item())


@pytest.mark.parametrize("shape", [(2, 32, 32, 640), (2, 8, 8, 1280), (1, 3, 5, 8), (6, 16, 16, 1280)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_upsample_nearest2x_nhwc_is_bit_exact(shape, dtype):
    """slider_upsample_nearest2x_nhwc vs F.interpolate(scale_factor=2, mode="nearest") (diffusers Upsample2D): a copy."""
    from sliders_conceptmod_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(*shape, device="cuda", generator=g).to(dtype)
    guard = torch.full((shape[0] * 4 * shape[1] * shape[2] * shape[3] + 64,), 7.0, device="cuda", dtype=dtype)
    y = ops.upsample_nearest2x_nhwc(x)
    ref = F.interpolate(x.permute(0, 3, 1, 2), scale_factor=2.0, mode="nearest").permute(0, 2, 3, 1)
    assert y.shape == ref.shape and torch.equal(y, ref.contiguous())
    assert bool((guard == 7.0).all())


@pytest.mark.parametrize("shape", [(2, 64, 64, 320, 320), (2, 16, 16, 1280, 640), (1, 3, 5, 8, 24), (6, 8, 8, 1280, 1280)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_concat_channels_nhwc_is_bit_exact(shape, dtype):
    """slider_concat_channels_nhwc vs torch.cat (the skip connections of the up path): a copy."""
    from sliders_conceptmod_b200 import ops
    B, H, W, ca, cb = shape
    g = torch.Generator(device="cuda").manual_seed(6)
    a = torch.randn(B, H, W, ca, device="cuda", generator=g).to(dtype)
    b = torch.randn(B, H, W, cb, device="cuda", generator=g).to(dtype)
    guard = torch.full((B * H * W * (ca + cb) + 64,), 7.0, device="cuda", dtype=dtype)
    y = ops.concat_channels_nhwc(a, b)
    assert torch.equal(y, torch.cat([a, b], dim=-1))
    assert bool((guard == 7.0).all())


@pytest.mark.parametrize("rows,C", [(8192, 320), (512, 1280), (77, 64), (3, 2560)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_layernorm_train_and_backward_match_autograd(rows, C, dtype):
    """Training form (stats out) + backward for frozen gamma / beta vs fp32 autograd of F.layer_norm."""
    from sliders_conceptmod_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(13)
    x = (torch.randn(rows, C, device="cuda", generator=g) * 2 + 0.5).to(dtype)
    w = (1 + 0.1 * torch.randn(C, device="cuda", generator=g)).to(dtype)
    b = (0.1 * torch.randn(C, device="cuda", generator=g)).to(dtype)
    gy = torch.randn(rows, C, device="cuda", generator=g).to(dtype)
    y, stats = ops.layernorm_train(x, w, b, 1e-5)
    assert torch.equal(y, ops.layernorm(x, w, b, 1e-5))
    xf = x.float().requires_grad_(True)
    yr = F.layer_norm(xf, (C,), w.float(), b.float(), 1e-5)
    yr.backward(gy.float())
    assert (stats[:, 0] - x.float().mean(dim=1)).abs().max().item() <= 1e-4
    dx = ops.layernorm_backward(gy, x, w, stats)
    tol = 3e-2 if dtype == torch.bfloat16 else 2e-5
    err = (dx.float() - xf.grad).abs().max().item()
    assert err <= tol * max(1.0, xf.grad.abs().max().item()), f"dx max-abs {err}"


@pytest.mark.parametrize("rows,D", [(8192, 1280), (512, 5120), (5, 64)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_geglu_backward_matches_autograd(rows, D, dtype):
    from sliders_conceptmod_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(14)
    x = torch.randn(rows, 2 * D, device="cuda", generator=g).to(dtype)
    gy = torch.randn(rows, D, device="cuda", generator=g).to(dtype)
    xf = x.float().requires_grad_(True)
    a, gate = xf.chunk(2, dim=-1)
    (a * F.gelu(gate)).backward(gy.float())
    dx = ops.geglu_backward(gy, x)
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-5
    err = (dx.float() - xf.grad).abs().max().item()
    assert err <= tol * max(1.0, xf.grad.abs().max().item()), f"dx max-abs {err}"


def test_transformer_block_autograd_functions():
    """The autograd wrappers the grad-recording UNet forward uses (unet._LayerNormFn / _GEGLUFn) against torch autograd."""
    from sliders_conceptmod_b200 import unet as U
    g = torch.Generator(device="cuda").manual_seed(15)
    norm = torch.nn.LayerNorm(320).cuda().to(torch.bfloat16).requires_grad_(False)
    with torch.no_grad():
        norm.weight.add_(0.1 * torch.randn(320, device="cuda", generator=g).to(torch.bfloat16))
    x = torch.randn(2, 64, 320, device="cuda", generator=g).to(torch.bfloat16).requires_grad_(True)
    h = torch.randn(2, 64, 640, device="cuda", generator=g).to(torch.bfloat16).requires_grad_(True)
    y = U.layer_norm_fused(norm, x)
    assert y.grad_fn is not None and "LayerNormFn" in type(y.grad_fn).__name__
    z = U._GEGLUFn.apply(h)
    (y.float().square().mean() + z.float().square().mean()).backward()
    xr = x.detach().float().requires_grad_(True)
    hr = h.detach().float().requires_grad_(True)
    a, gate = hr.chunk(2, dim=-1)
    (F.layer_norm(xr, (320,), norm.weight.float(), norm.bias.float(), norm.eps).square().mean()
     + (a * F.gelu(gate)).square().mean()).backward()
    for got, ref in ((x.grad, xr.grad), (h.grad, hr.grad)):
        assert (got.float() - ref).abs().max().item() <= 4e-2 * ref.abs().max().item()


@pytest.mark.parametrize("M,K,D", [(8192, 320, 1280), (2048, 640, 2560), (512, 1280, 5120), (128, 1280, 5120), (300, 64, 80)])
@pytest.mark.parametrize("dtype,tol", [(torch.bfloat16, 2e-2), (torch.float32, 1e-4)])
def test_linear_geglu_fused(M, K, D, dtype, tol):
    """slider_linear_geglu (feed-forward projection with the GEGLU gate in the GEMM's store) vs
    F.linear -> chunk -> a * gelu(g) in fp32, and vs the two-launch product path (GEMM, then slider_geglu)."""
    from sliders_conceptmod_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(21)
    x = torch.randn(M, K, device="cuda", generator=g).to(dtype)
    w = (torch.randn(2 * D, K, device="cuda", generator=g) / K ** 0.5).to(dtype)
    b = (0.2 * torch.randn(2 * D, device="cuda", generator=g)).to(dtype)
    idx = ops.geglu_permutation(D, "cuda")
    assert sorted(idx.tolist()) == list(range(2 * D))
    y = ops.linear_geglu(x, w.index_select(0, idx).contiguous(), b.index_select(0, idx).contiguous())
    h = F.linear(x.float(), w.float(), b.float())
    a, gate = h.chunk(2, dim=-1)
    ref = a * F.gelu(gate)
    err = (y.float() - ref).abs().max().item()
    assert y.shape == (M, D) and err <= tol * ref.abs().max().item(), f"max-abs {err:.3e} vs {ref.abs().max().item():.3e}"
    two, _ = ops.lora_forward(x, w.view(2 * D, 1, K), b, None, None, 0.0)
    two = ops.geglu(two)
    assert (y.float() - two.float()).abs().max().item() <= tol * ref.abs().max().item()
    # no bias
    y0 = ops.linear_geglu(x, w.index_select(0, idx).contiguous(), None)
    a0, g0 = F.linear(x.float(), w.float()).chunk(2, dim=-1)
    assert (y0.float() - a0 * F.gelu(g0)).abs().max().item() <= tol * ref.abs().max().item()
