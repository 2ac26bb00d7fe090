"""Cross-attention core for short key sets (slider_cross_attention) against fp32 softmax attention."""
import pytest
import torch

pytestmark = pytest.mark.gpu

# (batch, queries, keys, heads, head_dim): the attn2 layers of SD-1.4 (8 heads, d = 40 / 80 / 160), SD-XL (d = 64), ragged
CASES = [(2, 4096, 77, 8, 40), (2, 1024, 77, 8, 80), (2, 256, 77, 8, 160), (2, 64, 77, 8, 160), (6, 1024, 77, 8, 80),
         (2, 4096, 77, 10, 64), (2, 1024, 77, 20, 64), (1, 100, 77, 4, 48), (3, 65, 5, 2, 8), (1, 1, 80, 1, 128),
         (2, 300, 1, 3, 24)]


def _ref(q, k, v, heads):
    B, L, HD = q.shape
    d = HD // heads
    qf, kf, vf = (t.float().view(B, -1, heads, d).transpose(1, 2) for t in (q, k, v))
    p = torch.softmax(qf @ kf.transpose(-1, -2) / d ** 0.5, dim=-1)
    return (p @ vf).transpose(1, 2).reshape(B, L, HD)


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("kv_static", [False, True])
def test_cross_attention_matches_fp32_softmax_attention(case, dtype, kv_static):
    from sliders_conceptmod_b200 import ops
    B, L, S, H, d = case
    g = torch.Generator(device="cuda").manual_seed(L * 7 + S + d)
    q = (torch.randn(B, L, H * d, device="cuda", generator=g) * 1.5).to(dtype)
    k = (torch.randn(B, S, H * d, device="cuda", generator=g) * 1.5).to(dtype)
    v = torch.randn(B, S, H * d, device="cuda", generator=g).to(dtype)
    o = ops.cross_attention(q, k, v, H, kv_static=kv_static)
    ref = _ref(q, k, v, H)
    assert o.shape == q.shape and o.dtype == dtype
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3            # P and o are rounded to the storage type
    err = (o.float() - ref).abs().max().item()
    assert err <= tol * max(1.0, ref.abs().max().item()), f"max err {err}"
    # against the library kernel at the same precision: same order of error
    lib = torch.nn.functional.scaled_dot_product_attention(
        *(t.view(B, -1, H, d).transpose(1, 2) for t in (q, k, v))).transpose(1, 2).reshape(B, L, H * d)
    assert err <= 2.0 * (lib.float() - ref).abs().max().item() + 1e-3


def test_peaked_softmax_and_masked_tail():
    """One key dominates (large logits): no overflow, and the zero rows behind kv_len take no weight."""
    from sliders_conceptmod_b200 import ops
    B, L, S, H, d = 1, 70, 77, 2, 40
    q = torch.zeros(B, L, H * d, device="cuda", dtype=torch.bfloat16)
    k = torch.zeros(B, S, H * d, device="cuda", dtype=torch.bfloat16)
    v = torch.randn(B, S, H * d, device="cuda").bfloat16()
    q[..., 0] = 30.0; q[..., d] = 30.0
    k[:, 5, 0] = 30.0; k[:, 9, d] = 30.0                       # head 0 attends key 5, head 1 key 9
    o = ops.cross_attention(q, k, v, H)
    assert torch.isfinite(o.float()).all()
    assert torch.allclose(o[0, :, :d].float(), v[0, 5, :d].float().expand(L, d), atol=1e-2)
    assert torch.allclose(o[0, :, d:].float(), v[0, 9, d:].float().expand(L, d), atol=1e-2)
    # uniform attention (all logits equal) = the mean of the LIVE keys only
    q.zero_()
    o = ops.cross_attention(q, k, v, H)
    assert torch.allclose(o[0, 0].float(), v[0].float().mean(0), atol=2e-2)


def test_limits_are_reported():
    from sliders_conceptmod_b200 import ops
    q = torch.zeros(1, 8, 64, device="cuda", dtype=torch.bfloat16)
    kv = torch.zeros(1, 81, 64, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="keys"):
        ops.cross_attention(q, kv, kv, 2)
    with pytest.raises(Exception):
        ops.cross_attention(q.float(), kv[:, :77].float(), kv[:, :77].float(), 2)


def test_unet_attention_layer_uses_it_and_matches_the_library_path():
    """Attention.forward with a 77-token context on a no-grad bf16 forward: same result as the SDPA route."""
    import importlib
    from sliders_conceptmod_b200 import ops, unet
    torch.manual_seed(0)
    att = unet.Attention(320, 768, 8, 40).cuda().bfloat16().requires_grad_(False)
    x = torch.randn(2, 1024, 320, device="cuda").bfloat16()
    ctx = torch.randn(2, 77, 768, device="cuda").bfloat16()
    calls = []
    real = ops.attention
    ops.attention = lambda *a, **k: (calls.append(1), real(*a, **k))[1]
    try:
        with torch.no_grad():
            y = att(x, ctx)
    finally:
        ops.attention = real
    assert calls, "the attn2 layer did not take the short-key cross-attention kernel"
    unet._DISABLED.add("xattn")
    try:
        with torch.no_grad():
            y_lib = att(x, ctx)
    finally:
        unet._DISABLED.discard("xattn")
    assert (y.float() - y_lib.float()).abs().max().item() <= 3e-2 * max(1.0, y_lib.float().abs().max().item())



def test_strided_views_of_a_fused_projection():
    """The 64-token self-attention of the 8 x 8 level: q / k / v are last-dim slices of one [B, L, 3 * H * d] tensor."""
    from sliders_conceptmod_b200 import ops
    B, L, H, d = 2, 64, 8, 160
    qkv = torch.randn(B, L, 3 * H * d, device="cuda").bfloat16()
    q, k, v = qkv.split(H * d, dim=-1)
    assert ops.attention_usable(q, k, v, H)
    o = ops.attention(q, k, v, H)
    ref = _ref(q, k, v, H)
    assert o.is_contiguous() and (o.float() - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    assert not ops.attention_usable(qkv[:, :, :H * d], torch.zeros(B, 81, H * d, device="cuda").bfloat16(),
                                    torch.zeros(B, 81, H * d, device="cuda").bfloat16(), H)


def test_self_attention_layer_routes_by_key_count():
    """Attention.forward without context: 64 keys -> libslider, 256 keys -> library SDPA."""
    from sliders_conceptmod_b200 import ops, unet
    torch.manual_seed(1)
    att = unet.Attention(1280, None, 8, 160).cuda().bfloat16().requires_grad_(False)
    calls = []
    real = ops.attention
    ops.attention = lambda *a, **k: (calls.append(a[1].shape[1]), real(*a, **k))[1]
    try:
        with torch.no_grad():
            x = torch.randn(2, 64, 1280, device="cuda").bfloat16()
            y = att(x)
            unet._DISABLED.add("xattn")
            y_lib = att(x)
            unet._DISABLED.discard("xattn")
            att(torch.randn(2, 256, 1280, device="cuda").bfloat16())
    finally:
        ops.attention = real
        unet._DISABLED.discard("xattn")
    assert calls == [64]
    assert (y.float() - y_lib.float()).abs().max().item() <= 3e-2 * max(1.0, y_lib.float().abs().max().item())


@pytest.mark.parametrize("case", [(2, 4096, 77, 8, 40), (2, 1024, 77, 8, 80), (2, 256, 77, 8, 160), (2, 64, 77, 8, 160),
                                  (2, 1024, 77, 20, 64), (1, 100, 77, 4, 48), (3, 65, 5, 2, 8), (1, 70, 80, 1, 128)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_dq_backward_matches_fp32_autograd(case, dtype):
    """slider_cross_attention_backward (frozen keys / values: dq only) against autograd of fp32 softmax attention."""
    from sliders_conceptmod_b200 import ops
    B, L, S, H, d = case
    g = torch.Generator(device="cuda").manual_seed(L + S + d)
    q = (torch.randn(B, L, H * d, device="cuda", generator=g) * 1.2).to(dtype)
    k = (torch.randn(B, S, H * d, device="cuda", generator=g) * 1.2).to(dtype)
    v = torch.randn(B, S, H * d, device="cuda", generator=g).to(dtype)
    do = torch.randn(B, L, H * d, device="cuda", generator=g).to(dtype)
    qf = q.float().requires_grad_(True)
    _ref(qf, k, v, H).backward(do.float())
    dq = ops.cross_attention_backward(do, q, k, v, H)
    tol = 3e-2 if dtype == torch.bfloat16 else 6e-3
    err = (dq.float() - qf.grad).abs().max().item()
    assert err <= tol * max(1.0, qf.grad.abs().max().item()), f"max err {err} scale {qf.grad.abs().max().item()}"
    cos = torch.nn.functional.cosine_similarity(dq.float().flatten(), qf.grad.flatten(), dim=0).item()
    assert cos > 0.999, cos


def test_attn2_layer_trains_through_the_kernel_pair():
    """Attention.forward with a frozen 77-token context in grad mode: output and input gradient follow the SDPA route."""
    from sliders_conceptmod_b200 import ops, unet
    torch.manual_seed(0)
    att = unet.Attention(640, 768, 8, 80).cuda().bfloat16().requires_grad_(False)
    ctx = torch.randn(2, 77, 768, device="cuda").bfloat16()
    x = torch.randn(2, 1024, 640, device="cuda").bfloat16()
    go = torch.randn(2, 1024, 640, device="cuda").bfloat16()
    calls = []
    real = ops.cross_attention_backward
    ops.cross_attention_backward = lambda *a, **k: (calls.append(1), real(*a, **k))[1]
    try:
        x1 = x.clone().requires_grad_(True)
        att(x1, ctx).backward(go)
    finally:
        ops.cross_attention_backward = real
    assert calls, "the grad-recording attn2 layer did not take the libslider kernel pair"
    unet._DISABLED.add("xattn")
    try:
        x2 = x.clone().requires_grad_(True)
        att(x2, ctx).backward(go)
    finally:
        unet._DISABLED.discard("xattn")
    cos = torch.nn.functional.cosine_similarity(x1.grad.float().flatten(), x2.grad.float().flatten(), dim=0).item()
    assert cos > 0.999, cos
    assert (x1.grad.float() - x2.grad.float()).abs().max().item() <= 4e-2 * x2.grad.float().abs().max().item()
