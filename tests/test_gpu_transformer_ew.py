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
