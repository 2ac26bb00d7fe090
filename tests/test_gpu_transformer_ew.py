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
