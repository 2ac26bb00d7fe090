"""GPU parity of the NHWC GroupNorm(+pre-add, +SiLU) kernels against torch.nn.functional.group_norm in fp32."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

CASES = [  # B, H, W, C, G
    (2, 64, 64, 320, 32), (2, 32, 32, 640, 32), (2, 16, 16, 1280, 32), (2, 8, 8, 2560, 32), (2, 16, 16, 1920, 32),
    (6, 32, 32, 960, 32), (1, 16, 16, 32, 8), (3, 5, 7, 64, 8), (2, 128, 128, 320, 32),
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("silu,with_add", [(True, False), (True, True), (False, False)])
def test_groupnorm_nhwc_matches_torch(case, dtype, silu, with_add):
    from sliders_conceptmod_b200 import ops
    B, H, W, C, G = case
    gen = torch.Generator(device="cuda").manual_seed(7)
    x = (torch.randn(B, H, W, C, device="cuda", generator=gen) * 1.7 + 0.3).to(dtype)
    gamma = (1 + 0.2 * torch.randn(C, device="cuda", generator=gen)).to(dtype)
    beta = (0.1 * torch.randn(C, device="cuda", generator=gen)).to(dtype)
    add = torch.randn(B, C, device="cuda", generator=gen).to(dtype) if with_add else None
    for _ in range(2):      # twice: the ticket words must be left re-armed
        y = ops.groupnorm_nhwc(x, G, gamma, beta, 1e-5, silu, add)
    torch.cuda.synchronize()
    xr = x.float().permute(0, 3, 1, 2)
    if add is not None:
        xr = xr + add.float()[:, :, None, None]
    yr = F.group_norm(xr, G, gamma.float(), beta.float(), 1e-5)
    if silu:
        yr = F.silu(yr)
    yr = yr.permute(0, 2, 3, 1)
    err = (y.float() - yr).abs().max().item()
    tol = 3e-2 if dtype == torch.bfloat16 else 2e-5
    assert err <= tol * max(1.0, yr.abs().max().item()), f"max-abs {err}"


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("silu,with_add", [(True, False), (True, True), (False, False)])
def test_groupnorm_nhwc_backward_matches_torch_autograd(case, dtype, silu, with_add):
    """Training form: forward with saved {mean, rstd} + fused backward (dx, and d chan_add) vs autograd of the fp32
    torch expression act(group_norm(x + add))."""
    from sliders_conceptmod_b200 import ops
    B, H, W, C, G = case
    gen = torch.Generator(device="cuda").manual_seed(11)
    x = (torch.randn(B, H, W, C, device="cuda", generator=gen) * 1.7 + 0.3).to(dtype)
    gy = torch.randn(B, H, W, C, device="cuda", generator=gen).to(dtype)
    gamma = (1 + 0.2 * torch.randn(C, device="cuda", generator=gen)).to(dtype)
    beta = (0.1 * torch.randn(C, device="cuda", generator=gen)).to(dtype)
    add = torch.randn(B, C, device="cuda", generator=gen).to(dtype) if with_add else None
    y, stats = ops.groupnorm_nhwc(x, G, gamma, beta, 1e-5, silu, add, want_stats=True)
    dx, dadd = ops.groupnorm_nhwc_backward(gy, x, stats, G, gamma, beta, silu, add, want_dadd=with_add)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    ar = add.float().requires_grad_(True) if with_add else None
    z = xr + ar[:, :, None, None] if with_add else xr
    yr = F.group_norm(z, G, gamma.float(), beta.float(), 1e-5)
    if silu:
        yr = F.silu(yr)
    yr.backward(gy.float().permute(0, 3, 1, 2))
    assert (y.float() - yr.detach().permute(0, 2, 3, 1)).abs().max().item() <= (3e-2 if dtype == torch.bfloat16 else 2e-5) * max(1.0, yr.abs().max().item())
    ref = xr.grad.permute(0, 2, 3, 1)
    err = (dx.float() - ref).abs().max().item()
    tol = 2e-2 if dtype == torch.bfloat16 else 5e-5
    assert err <= tol * max(1e-3, ref.abs().max().item()), f"dx max-abs {err} vs {ref.abs().max().item()}"
    if with_add:
        err = (dadd.float() - ar.grad).abs().max().item()
        # a sum over H*W pixels of bf16-rounded terms in the reference vs fp32 accumulation here
        assert err <= (3e-2 if dtype == torch.bfloat16 else 2e-4) * max(1e-3, ar.grad.abs().max().item()), f"dadd max-abs {err}"
    else:
        assert dadd is None


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("slab_kb", [4, 16, 48, 128])
def test_groupnorm_group_form_matches_two_kernel_form(dtype, slab_kb):
    """The default single-launch form (one CTA or a cluster of 2 / 4 / 8 per (image, group), slider_debug_gn_group) vs
    the two-kernel form it replaces, over cluster sizes (slab_kb = bytes of a group one CTA may hold) and the saved
    statistics of the training form."""
    import ctypes
    from sliders_conceptmod_b200 import ops, _lib
    fn = _lib.lib().slider_debug_gn_group
    fn.argtypes, fn.restype = [ctypes.c_int, ctypes.c_int], ctypes.c_int
    gen = torch.Generator(device="cuda").manual_seed(13)
    tol = 2e-2 if dtype == torch.bfloat16 else 2e-5
    for (B, H, W, C, G) in [(2, 64, 64, 320, 32), (2, 64, 64, 960, 32), (2, 32, 32, 1920, 32), (2, 8, 8, 1280, 32),
                            (3, 5, 7, 64, 8), (1, 6, 6, 48, 8), (2, 16, 16, 2560, 32), (1, 4, 4, 40, 8), (2, 128, 128, 320, 32)]:
        x = (torch.randn(B, H, W, C, device="cuda", generator=gen) * 1.7 + 0.3).to(dtype)
        gamma = (1 + 0.2 * torch.randn(C, device="cuda", generator=gen)).to(dtype)
        beta = (0.1 * torch.randn(C, device="cuda", generator=gen)).to(dtype)
        add = torch.randn(B, C, device="cuda", generator=gen).to(dtype)
        try:
            fn(0, 0)
            ref, ref_stats = ops.groupnorm_nhwc(x, G, gamma, beta, 1e-5, True, add, want_stats=True)
            fn(1, slab_kb)
            y, stats = ops.groupnorm_nhwc(x, G, gamma, beta, 1e-5, True, add, want_stats=True)
            y2 = ops.groupnorm_nhwc(x, G, gamma, beta, 1e-5, True, add)
            torch.cuda.synchronize()
        finally:
            fn(1, 48)
        assert torch.equal(y, y2)       # deterministic
        assert (y.float() - ref.float()).abs().max().item() <= tol * ref.float().abs().max().item()
        assert (stats - ref_stats).abs().max().item() <= 1e-4 * max(1.0, ref_stats.abs().max().item())
