"""Tensor-level wrappers over the C ABI (one function per exported kernel entry point).

All tensors are CUDA tensors; activations of conv layers are NHWC-contiguous
(`[B, H, W, C]`), linear activations are `[M, K]`.  Nothing here computes on the host.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import LayerGeom, check, dtype_code, lib, ptr, require_cuda, stream_ptr


def make_geom(dtype: torch.dtype, batch: int, height: int, width: int, cin: int, cout: int,
              ksize: int = 1, rank: int = 0, stride: int = 1) -> LayerGeom:
    return LayerGeom(dtype_code(dtype), batch, height, width, cin, cout, ksize, ksize, (ksize - 1) // 2, rank, stride)


def _act_dims(x: torch.Tensor) -> tuple[int, int, int]:
    if x.dim() == 4:
        return x.shape[0], x.shape[1], x.shape[2]
    if x.dim() == 2:
        return 1, 1, x.shape[0]
    raise ValueError(f"expected [B,H,W,C] or [M,K] activations, got {tuple(x.shape)}")


_lane = 0


class workspace_lane:
    """Scratch buffers (split-K partials, GroupNorm partials, ...) are shared by all calls ordered on ONE stream.
    Work issued concurrently on another stream (a second branch of a captured graph) must use its own set:
    `with ops.workspace_lane(1): ...` around that branch."""

    def __init__(self, lane: int):
        self.lane, self.prev = lane, 0

    def __enter__(self):
        global _lane
        self.prev, _lane = _lane, self.lane

    def __exit__(self, *exc):
        global _lane
        _lane = self.prev


def current_lane() -> int:
    return _lane


def _ws_key(device: torch.device):
    return (device.index if device.index is not None else torch.cuda.current_device(), _lane)


_gemm_ws: dict = {}
# Scratch buffers that were outgrown.  A captured CUDA graph bakes the raw pointer of the buffer it was recorded with,
# so a buffer that has ever been handed to a launch is never returned to the allocator: it is parked here and the
# graph that references it keeps replaying over valid memory (a few MB per growth step, a handful of steps per process).
_retired: list = []


def _grow(table: dict, key, need: int, slack: int, zero: bool, device) -> torch.Tensor:
    ws = table.get(key)
    if ws is None or ws.numel() < need:
        if ws is not None:
            _retired.append(ws)
        make = torch.zeros if zero else torch.empty
        ws = make(max(need + slack, 1 << 20), dtype=torch.uint8, device=device)
        table[key] = ws
    return ws


def _gemm_workspace(device: torch.device) -> torch.Tensor:
    """Per-device split-K scratch (contents don't-care; every call on the stream may share it)."""
    key = _ws_key(device)
    ws = _gemm_ws.get(key)
    if ws is None:
        ws = torch.empty(lib().slider_gemm_workspace_bytes(), dtype=torch.uint8, device=device)
        _gemm_ws[key] = ws
    return ws


def lora_forward(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None, a: torch.Tensor | None,
                 b: torch.Tensor | None, s: float, ksize: int = 1, want_h: bool = False,
                 residual: torch.Tensor | None = None, stride: int = 1):
    """y = conv(x, w) + bias + s * (conv(x, a) @ b^T).   (lora.py:134-138)

    x [B,H,W,Cin] | [M,Cin];  w [Cout, k*k, Cin];  a [r, k*k, Cin];  b [Cout, r].
    Returns (y, h) with h = conv(x, a) in fp32 ([output pixels, r]) when want_h else None.
    `residual` (contiguous, shaped like y): the result is added INTO it by the store and it is returned as y.
    `stride` 2 (3x3 only): y is [B, (H-1)//2+1, (W-1)//2+1, Cout].
    """
    require_cuda(x, w, bias, a, b, residual)
    assert x.is_contiguous() and w.is_contiguous()
    B, H, W = _act_dims(x)
    cin, cout = x.shape[-1], w.shape[0]
    rank = 0 if a is None else a.shape[0]
    g = make_geom(x.dtype, B, H, W, cin, cout, ksize, rank, stride)
    Ho, Wo = ((H - 1) // 2 + 1, (W - 1) // 2 + 1) if stride == 2 else (H, W)
    if residual is not None:
        if not residual.is_contiguous() or residual.dtype != x.dtype or residual.numel() != B * Ho * Wo * cout:
            raise ValueError("lora_forward: residual must be contiguous, of the activation dtype and shaped like y")
        y = residual
    elif stride == 2:
        y = torch.empty(B, Ho, Wo, cout, device=x.device, dtype=x.dtype)
    else:
        y = torch.empty(*x.shape[:-1], cout, device=x.device, dtype=x.dtype)
    h = torch.empty(B * Ho * Wo, rank, device=x.device, dtype=torch.float32) if (want_h and rank > 0) else None
    ws = _gemm_workspace(x.device)
    if residual is not None:
        check(lib().slider_lora_forward_ex(C.byref(g), ptr(y), ptr(x), ptr(w), ptr(bias), ptr(a), ptr(b), ptr(h),
                                           float(s), 1, ptr(ws), ws.numel(), stream_ptr()))
    else:
        check(lib().slider_lora_forward(C.byref(g), ptr(y), ptr(x), ptr(w), ptr(bias), ptr(a), ptr(b), ptr(h),
                                        float(s), ptr(ws), ws.numel(), stream_ptr()))
    return y, h


def lora_forward_grouped(x: torch.Tensor, w_cat: torch.Tensor, a_cat: torch.Tensor, b_cat: torch.Tensor, groups: int,
                         s: float) -> torch.Tensor:
    """Several Linear adapters on one input in ONE launch (self-attention to_q/to_k/to_v):
    x [M, K]; w_cat [groups*Ng, 1, K]; a_cat [groups*r, 1, K]; b_cat [groups*Ng, r] -> y [M, groups*Ng]."""
    require_cuda(x, w_cat, a_cat, b_cat)
    N, _, K = w_cat.shape
    r = a_cat.shape[0] // groups
    assert x.dim() == 2 and x.is_contiguous() and w_cat.is_contiguous() and a_cat.is_contiguous() and b_cat.is_contiguous()
    g = make_geom(x.dtype, 1, 1, x.shape[0], K, N, 1, r)
    y = torch.empty(x.shape[0], N, device=x.device, dtype=x.dtype)
    ws = _gemm_workspace(x.device)
    check(lib().slider_lora_forward_grouped(C.byref(g), ptr(y), ptr(x), ptr(w_cat), None, ptr(a_cat), ptr(b_cat),
                                            int(groups), float(s), ptr(ws), ws.numel(), stream_ptr()))
    return y


def geglu_permutation(d: int, device=None) -> torch.Tensor:
    """Row order of the GEGLU projection that `linear_geglu` expects: every 160-row block = [a(40) | g(40) | a(40) | g(40)]
    (a = rows [0, d) of the projection, g = rows [d, 2d))."""
    assert d % 80 == 0
    t = torch.arange(d // 80).view(-1, 1, 1, 1)
    h = torch.arange(2).view(1, -1, 1, 1)
    s = torch.arange(2).view(1, 1, -1, 1)
    i = torch.arange(40).view(1, 1, 1, -1)
    return (s * d + 80 * t + 40 * h + i).reshape(-1).to(device)


def linear_geglu(x: torch.Tensor, w_perm: torch.Tensor, bias_perm: torch.Tensor | None) -> torch.Tensor:
    """(x @ Wa^T + ba) * gelu(x @ Wg^T + bg) in ONE tensor-core launch; w_perm [2d, K] / bias_perm [2d] are the GEGLU
    projection's rows in `geglu_permutation` order.  x [M, K] -> [M, d]."""
    require_cuda(x, w_perm, bias_perm)
    assert x.dim() == 2 and x.is_contiguous() and w_perm.is_contiguous()
    n2, K = w_perm.shape[0], w_perm.shape[-1]
    g = make_geom(x.dtype, 1, 1, x.shape[0], K, n2, 1, 0)
    y = torch.empty(x.shape[0], n2 // 2, device=x.device, dtype=x.dtype)
    check(lib().slider_linear_geglu(C.byref(g), ptr(y), ptr(x), ptr(w_perm), ptr(bias_perm), stream_ptr()))
    return y


class SmallMultiPlan:
    """Pointer tables of one `lora_forward_small_multi` call, built once per (layer set, input buffer): the frozen
    weights, the flat LoRA parameter buffer and the output buffer never move, so a captured launch stays valid."""

    def __init__(self, x_rows: int, k: int, dtype: torch.dtype, device, layers):
        # layers: list of (w [N,1,K] | [N,K], bias [N] | None, a [r,1,K] | None, b [N,r] | None)
        self.n = [int(l[0].shape[0]) for l in layers]
        self.rank = max([0] + [int(l[2].shape[0]) for l in layers if l[2] is not None])
        self.m, self.k, self.dtype = int(x_rows), int(k), dtype
        self.out = torch.empty(self.m * sum(self.n), device=device, dtype=dtype)
        self.views, off = [], 0
        for n in self.n:
            self.views.append(self.out[off * self.m:(off + n) * self.m].view(self.m, n))
            off += n
        L = len(layers)
        vp = C.c_void_p * L
        self.keep = layers
        self.W = vp(*[l[0].data_ptr() for l in layers])
        self.bias = vp(*[(l[1].data_ptr() if l[1] is not None else None) for l in layers])
        self.A = vp(*[(l[2].data_ptr() if l[2] is not None else None) for l in layers])
        self.B = vp(*[(l[3].data_ptr() if l[3] is not None else None) for l in layers])
        self.Y = vp(*[v.data_ptr() for v in self.views])
        self.N = (C.c_int * L)(*self.n)
        self.L = L


def lora_forward_small_multi(x: torch.Tensor, plan: SmallMultiPlan, scales) -> list:
    """All small-M adapters of `plan` on the same rows x [M <= 16, K] in ONE launch (the time_emb_proj layers of a
    UNet forward); returns the per-layer outputs (views of the plan's buffer)."""
    require_cuda(x)
    assert x.dim() == 2 and x.is_contiguous() and x.shape == (plan.m, plan.k) and x.dtype == plan.dtype
    s = (C.c_float * plan.L)(*[float(v) for v in scales])
    check(lib().slider_lora_forward_small_multi(dtype_code(x.dtype), plan.L, plan.m, plan.k, plan.rank, ptr(x),
                                                plan.Y, plan.W, plan.bias, plan.A, plan.B, plan.N, s, stream_ptr()))
    return plan.views


def lora_input_grad(gy: torch.Tensor, wt: torch.Tensor, dh: torch.Tensor | None, at: torch.Tensor | None,
                    cin: int, ksize: int = 1, rank: int = 0) -> torch.Tensor:
    """dx = convT(gy, W) + convT(dh, A);  wt [Cin, k*k, Cout] flipped taps, at [Cin, k*k, 16]."""
    require_cuda(gy, wt, dh, at)
    assert gy.is_contiguous() and wt.is_contiguous()
    B, H, W = _act_dims(gy)
    g = make_geom(gy.dtype, B, H, W, cin, gy.shape[-1], ksize, rank)
    dx = torch.empty(*gy.shape[:-1], cin, device=gy.device, dtype=gy.dtype)
    ws = _gemm_workspace(gy.device)
    check(lib().slider_lora_input_grad(C.byref(g), ptr(dx), ptr(gy), ptr(wt), ptr(dh), ptr(at), ptr(ws), ws.numel(),
                                       stream_ptr()))
    return dx


_loss_ws: dict[int, torch.Tensor] = {}


def _reduce_ws(device: torch.device) -> torch.Tensor:
    key = _ws_key(device)
    ws = _loss_ws.get(key)
    if ws is None:
        ws = torch.zeros(lib().slider_loss_workspace_bytes(), dtype=torch.uint8, device=device)
        _loss_ws[key] = ws
    return ws


def guidance_loss(t: torch.Tensor, p: torch.Tensor | None, n: torch.Tensor, q: torch.Tensor | None,
                  guidance: float, sign: int, want_grad: bool = True):
    """loss = mean((t - (n + sign*g*(p - q)))^2), grad = dloss/dt  (prompt_util.py:108-148).

    Returns (loss fp32 0-dim, grad like t or None).  p is None => plain MSE(t, n).
    """
    require_cuda(t, p, n, q)
    ts = [x.contiguous() for x in (t, n)] + [None if x is None else x.contiguous() for x in (p, q)]
    t_, n_, p_, q_ = ts
    for x in (n_, p_, q_):
        if x is not None and (x.shape != t_.shape or x.dtype != t_.dtype):
            raise ValueError("guidance_loss: all latents must share shape and dtype")
    loss = torch.empty((), device=t.device, dtype=torch.float32)
    grad = torch.empty_like(t_) if want_grad else None
    ws = _reduce_ws(t.device)
    check(lib().slider_guidance_loss(ptr(loss), ptr(grad), ptr(t_), ptr(p_), ptr(n_), ptr(q_), t_.numel(),
                                     float(guidance), int(sign), dtype_code(t_.dtype), ptr(ws), ws.numel(),
                                     stream_ptr()))
    return loss, grad


def cfg_step(x: torch.Tensor | None, eps_u: torch.Tensor | None, eps_c: torch.Tensor | None,
             noise: torch.Tensor | None = None, guidance: float = 1.0, c_x: float = 1.0, c_eps: float = 1.0,
             c_noise: float = 0.0, dev_coef: torch.Tensor | None = None, out: torch.Tensor | None = None):
    """out = c_x*x + c_eps*(eps_u + g*(eps_c - eps_u)) + c_noise*noise  (train_util.py:297-300,324)."""
    ref = x if x is not None else eps_u
    require_cuda(x, eps_u, eps_c, noise, dev_coef)
    for tns in (x, eps_u, eps_c, noise):
        if tns is not None and not tns.is_contiguous():
            raise ValueError("cfg_step: tensors must be contiguous")
    if out is None:
        out = torch.empty_like(ref)
    check(lib().slider_cfg_step(ptr(out), ptr(x), ptr(eps_u), ptr(eps_c), ptr(noise), ref.numel(), float(guidance),
                                float(c_x), float(c_eps), float(c_noise), ptr(dev_coef), dtype_code(ref.dtype),
                                stream_ptr()))
    return out


def scale_by_device_scalar(x: torch.Tensor, scalar: torch.Tensor) -> torch.Tensor:
    require_cuda(x, scalar)
    assert x.is_contiguous() and scalar.dtype == torch.float32
    check(lib().slider_scale_by_device_scalar(ptr(x), ptr(scalar), x.numel(), dtype_code(x.dtype), stream_ptr()))
    return x


_gn_ws: dict[int, torch.Tensor] = {}


def groupnorm_nhwc(x: torch.Tensor, groups: int, gamma: torch.Tensor | None, beta: torch.Tensor | None,
                   eps: float = 1e-5, silu: bool = False, chan_add: torch.Tensor | None = None,
                   want_stats: bool = False):
    """y = act(GroupNorm(x + chan_add[:, None, :])) on NHWC data.  x [B, ..., C] contiguous (C innermost),
    chan_add [B, C] (the time-embedding projection of ResnetBlock2D) or None."""
    require_cuda(x, gamma, beta, chan_add)
    assert x.is_contiguous()
    B, Cc = x.shape[0], x.shape[-1]
    hw = x.numel() // (B * Cc)
    ws = _grow(_gn_ws, _ws_key(x.device), lib().slider_groupnorm_workspace_bytes(int(B), int(groups)), 0, True, x.device)
    if chan_add is not None:
        chan_add = chan_add.contiguous()
        assert chan_add.shape == (B, Cc) and chan_add.dtype == x.dtype
    y = torch.empty_like(x)
    if want_stats:
        stats = torch.empty(B, groups, 2, device=x.device, dtype=torch.float32)
        check(lib().slider_groupnorm_nhwc_train(ptr(y), ptr(stats), ptr(x), ptr(chan_add), ptr(gamma), ptr(beta), int(B),
                                                int(hw), int(Cc), int(groups), float(eps), int(bool(silu)),
                                                dtype_code(x.dtype), ptr(ws), ws.numel(), stream_ptr()))
        return y, stats
    check(lib().slider_groupnorm_nhwc(ptr(y), ptr(x), ptr(chan_add), ptr(gamma), ptr(beta), int(B), int(hw), int(Cc),
                                      int(groups), float(eps), int(bool(silu)), dtype_code(x.dtype), ptr(ws),
                                      ws.numel(), stream_ptr()))
    return y


_gn_bwd_ws: dict[int, torch.Tensor] = {}


def groupnorm_nhwc_backward(gy: torch.Tensor, x: torch.Tensor, stats: torch.Tensor, groups: int,
                            gamma: torch.Tensor | None, beta: torch.Tensor | None, silu: bool,
                            chan_add: torch.Tensor | None = None, want_dadd: bool = False):
    """Backward of groupnorm_nhwc for frozen gamma / beta: returns (dx like x, dadd [B, C] or None)."""
    require_cuda(gy, x, stats, gamma, beta, chan_add)
    assert x.is_contiguous() and gy.is_contiguous() and gy.shape == x.shape and gy.dtype == x.dtype
    B, Cc = x.shape[0], x.shape[-1]
    hw = x.numel() // (B * Cc)
    ws = _grow(_gn_bwd_ws, _ws_key(x.device), lib().slider_groupnorm_bwd_workspace_bytes(int(B), int(Cc), int(groups)),
               0, False, x.device)
    dx = torch.empty_like(x)
    dadd = torch.empty(B, Cc, device=x.device, dtype=x.dtype) if (want_dadd and chan_add is not None) else None
    check(lib().slider_groupnorm_nhwc_backward(ptr(dx), ptr(dadd), ptr(gy), ptr(x), ptr(chan_add), ptr(gamma), ptr(beta),
                                               ptr(stats), int(B), int(hw), int(Cc), int(groups), int(bool(silu)),
                                               dtype_code(x.dtype), ptr(ws), ws.numel(), stream_ptr()))
    return dx, dadd


def layernorm(x: torch.Tensor, gamma: torch.Tensor | None, beta: torch.Tensor | None, eps: float = 1e-5) -> torch.Tensor:
    """LayerNorm over the last dimension (diffusers BasicTransformerBlock.norm1/2/3)."""
    require_cuda(x, gamma, beta)
    assert x.is_contiguous()
    C_ = x.shape[-1]
    y = torch.empty_like(x)
    check(lib().slider_layernorm(ptr(y), ptr(x), ptr(gamma), ptr(beta), x.numel() // C_, int(C_), float(eps),
                                 dtype_code(x.dtype), stream_ptr()))
    return y


def layernorm_train(x: torch.Tensor, gamma, beta, eps: float = 1e-5):
    """layernorm + the per-row {mean, rstd} (fp32 [rows, 2]) its backward takes back."""
    require_cuda(x, gamma, beta)
    assert x.is_contiguous()
    C_ = x.shape[-1]
    y = torch.empty_like(x)
    stats = torch.empty(x.numel() // C_, 2, device=x.device, dtype=torch.float32)
    check(lib().slider_layernorm_train(ptr(y), ptr(stats), ptr(x), ptr(gamma), ptr(beta), x.numel() // C_, int(C_),
                                       float(eps), dtype_code(x.dtype), stream_ptr()))
    return y, stats


def layernorm_backward(gy: torch.Tensor, x: torch.Tensor, gamma, stats: torch.Tensor) -> torch.Tensor:
    """Input gradient of LayerNorm with frozen gamma / beta."""
    require_cuda(gy, x, gamma, stats)
    assert gy.is_contiguous() and x.is_contiguous() and gy.shape == x.shape and gy.dtype == x.dtype
    C_ = x.shape[-1]
    dx = torch.empty_like(x)
    check(lib().slider_layernorm_backward(ptr(dx), ptr(gy), ptr(x), ptr(gamma), ptr(stats), x.numel() // C_, int(C_),
                                          dtype_code(x.dtype), stream_ptr()))
    return dx


def geglu_backward(gy: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    """Gradient of geglu(x) w.r.t. x [..., 2d] given gy [..., d]."""
    require_cuda(gy, x)
    assert gy.is_contiguous() and x.is_contiguous() and x.shape[-1] == 2 * gy.shape[-1]
    d = gy.shape[-1]
    dx = torch.empty_like(x)
    check(lib().slider_geglu_backward(ptr(dx), ptr(gy), ptr(x), gy.numel() // d, int(d), dtype_code(x.dtype), stream_ptr()))
    return dx


def geglu(x: torch.Tensor) -> torch.Tensor:
    """x[..., :d] * gelu(x[..., d:]) for x [..., 2d] (diffusers GEGLU after its projection)."""
    require_cuda(x)
    assert x.is_contiguous() and x.shape[-1] % 2 == 0
    d = x.shape[-1] // 2
    y = torch.empty(*x.shape[:-1], d, device=x.device, dtype=x.dtype)
    check(lib().slider_geglu(ptr(y), ptr(x), x.numel() // (2 * d), int(d), dtype_code(x.dtype), stream_ptr()))
    return y


def upsample_nearest2x_nhwc(x: torch.Tensor) -> torch.Tensor:
    """F.interpolate(scale_factor=2, mode="nearest") on NHWC data: x [B, H, W, C] contiguous -> [B, 2H, 2W, C]."""
    require_cuda(x)
    assert x.dim() == 4 and x.is_contiguous()
    B, H, W, Cc = x.shape
    y = torch.empty(B, 2 * H, 2 * W, Cc, device=x.device, dtype=x.dtype)
    check(lib().slider_upsample_nearest2x_nhwc(ptr(y), ptr(x), int(B), int(H), int(W), int(Cc), dtype_code(x.dtype),
                                               stream_ptr()))
    return y


def concat_channels_nhwc(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """torch.cat([a, b], dim=-1) for NHWC tensors a [..., Ca], b [..., Cb] (same leading shape, contiguous)."""
    require_cuda(a, b)
    assert a.is_contiguous() and b.is_contiguous() and a.shape[:-1] == b.shape[:-1] and a.dtype == b.dtype
    ca, cb = a.shape[-1], b.shape[-1]
    y = torch.empty(*a.shape[:-1], ca + cb, device=a.device, dtype=a.dtype)
    check(lib().slider_concat_channels_nhwc(ptr(y), ptr(a), ptr(b), a.numel() // ca, int(ca), int(cb),
                                            dtype_code(a.dtype), stream_ptr()))
    return y


def cross_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int, kv_static: bool = False) -> torch.Tensor:
    """softmax(q k^T / sqrt(d)) v per head for q [B, L, H*d], k / v [B, S <= 80, H*d] (contiguous, 16-bit): the attn2
    core with the 77 prompt tokens resident in shared memory; returns o [B, L, H*d] (no head transposes)."""
    require_cuda(q, k, v)
    assert q.is_contiguous() and k.is_contiguous() and v.is_contiguous() and q.dtype == k.dtype == v.dtype
    B, L, HD = q.shape
    S = k.shape[1]
    assert k.shape == v.shape == (B, S, HD) and HD % heads == 0
    o = torch.empty_like(q)
    check(lib().slider_cross_attention(ptr(o), ptr(q), ptr(k), ptr(v), int(B), int(L), int(S), int(heads), int(HD // heads),
                                       dtype_code(q.dtype), int(bool(kv_static)), stream_ptr()))
    return o


def _rows_ld(t: torch.Tensor):
    """Row stride (elements) of a [B, len, C] tensor whose rows are dense and whose batch items follow each other at
    len * stride(1) — a dense tensor or a last-dim slice of one; None otherwise."""
    if t.dim() != 3 or t.stride(2) != 1 or t.stride(1) % 8 or (t.shape[0] > 1 and t.stride(0) != t.shape[1] * t.stride(1)):
        return None
    return int(t.stride(1))


def attention_usable(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int, max_keys: int = 80) -> bool:
    if not (q.is_cuda and q.dtype in (torch.bfloat16, torch.float16) and q.dtype == k.dtype == v.dtype):
        return False
    d = q.shape[-1] // heads
    return (d % 8 == 0 and 8 <= d <= 160 and q.shape[-1] == heads * d and k.shape[1] <= min(max_keys, 80) and q.shape[0] * heads <= 65535
            and all(_rows_ld(t) is not None and t.data_ptr() % 16 == 0 for t in (q, k, v)))


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int, kv_static: bool = False) -> torch.Tensor:
    """softmax(q k^T / sqrt(d)) v per head for q [B, L, H*d], k / v [B, S <= 80, H*d] (dense, or the views of one fused
    q/k/v projection), every key resident in shared memory.  Returns a dense o [B, L, H*d]."""
    require_cuda(q, k, v)
    B, L, HD = q.shape
    S = k.shape[1]
    assert k.shape == v.shape == (B, S, HD) and HD % heads == 0
    o = torch.empty(B, L, HD, device=q.device, dtype=q.dtype)
    if S <= 80 and q.is_contiguous() and k.is_contiguous() and v.is_contiguous():
        check(lib().slider_cross_attention(ptr(o), ptr(q), ptr(k), ptr(v), int(B), int(L), int(S), int(heads), int(HD // heads),
                                           dtype_code(q.dtype), int(bool(kv_static)), stream_ptr()))
        return o
    lds = [_rows_ld(t) for t in (q, k, v)]
    if any(ld is None for ld in lds):
        raise ValueError("attention: q / k / v must be [B, len, C] with dense rows (a dense tensor or a last-dim slice of one)")
    check(lib().slider_cross_attention_strided(ptr(o), ptr(q), ptr(k), ptr(v), int(B), int(L), int(S), int(heads),
                                               int(HD // heads), lds[0], lds[1], lds[2], int(HD), dtype_code(q.dtype),
                                               int(bool(kv_static)), stream_ptr()))
    return o


def cross_attention_backward(dout: torch.Tensor, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int) -> torch.Tensor:
    """dq of attention(q, k, v) for frozen k / v (dense [B, len, H*d] tensors, S <= 80 keys)."""
    require_cuda(dout, q, k, v)
    assert all(t.is_contiguous() for t in (dout, q, k, v)) and dout.shape == q.shape and q.dtype == k.dtype == v.dtype == dout.dtype
    B, L, HD = q.shape
    dq = torch.empty_like(q)
    check(lib().slider_cross_attention_backward(ptr(dq), ptr(dout), ptr(q), ptr(k), ptr(v), int(B), int(L), int(k.shape[1]),
                                                int(heads), int(HD // heads), dtype_code(q.dtype), stream_ptr()))
    return dq


def launch_count() -> int:
    return int(lib().slider_launch_count())


def pack_wt(w: torch.Tensor, ksize: int = 1) -> torch.Tensor:
    """Wt[c][t][n] = W[n][T-1-t][c] — one-time re-layout of a frozen base weight [Cout, k*k, Cin]."""
    require_cuda(w)
    cout, taps, cin = w.shape
    g = make_geom(w.dtype, 1, 1, 1, cin, cout, ksize, 0)
    wt = torch.empty(cin, taps, cout, device=w.device, dtype=w.dtype)
    check(lib().slider_pack_wt(C.byref(g), ptr(wt), ptr(w.contiguous()), stream_ptr()))
    return wt


def pack_at(a: torch.Tensor, ksize: int = 1) -> torch.Tensor:
    """At[c][t][j] = A[j][T-1-t][c], rank padded to 16 (operand of the input-gradient GEMM)."""
    require_cuda(a)
    rank, taps, cin = a.shape
    g = make_geom(a.dtype, 1, 1, 1, cin, 8, ksize, rank)
    at = torch.empty(cin, taps, 16, device=a.device, dtype=a.dtype)
    check(lib().slider_lora_pack_at(C.byref(g), ptr(at), ptr(a.contiguous()), stream_ptr()))
    return at


_grad_ws: dict[int, torch.Tensor] = {}


def _grad_workspace(device: torch.device, need: int) -> torch.Tensor:
    # zero-filled at allocation: the first 4 KB are the reduction tickets, which the kernels leave zero
    return _grow(_grad_ws, _ws_key(device), need, 1 << 20, True, device)


def lora_grad(gy: torch.Tensor, x: torch.Tensor, h: torch.Tensor | None, b: torch.Tensor, s: float, ksize: int,
              dA: torch.Tensor | None, dB: torch.Tensor | None, accumulate: bool = False, stride: int = 1,
              dA_chan_scale: torch.Tensor | None = None) -> torch.Tensor:
    """Low-rank backward (autograd of lora.py:134-138): fills dA [r,k*k,Cin], dB [Cout,r] (fp32; both None: skipped)
    and returns dh [output pixels,16] = s * gy @ B (element dtype, zero padded).  x [B,H,W,Cin] is the layer input,
    gy the gradient of its output ([B,H/stride,W/stride,Cout])."""
    require_cuda(gy, x, h, b, dA, dB)
    assert gy.is_contiguous() and x.is_contiguous() and b.is_contiguous()
    if dA is not None:
        assert h is not None and h.is_contiguous()
        assert dA.dtype == torch.float32 and dB.dtype == torch.float32 and dA.is_contiguous() and dB.is_contiguous()
    B, H, W = _act_dims(x)
    rank = b.shape[1]
    g = make_geom(x.dtype, B, H, W, x.shape[-1], gy.shape[-1], ksize, rank, stride)
    dh = torch.empty(gy.numel() // gy.shape[-1], 16, device=x.device, dtype=x.dtype)
    ws = _grad_workspace(x.device, lib().slider_lora_grad_workspace_bytes(C.byref(g)))
    if dA_chan_scale is not None:
        assert dA_chan_scale.dtype == torch.float32 and dA_chan_scale.numel() == x.shape[-1]
    check(lib().slider_lora_grad_ex(C.byref(g), ptr(dA), ptr(dB), ptr(dh), ptr(gy), ptr(x), ptr(h), ptr(b), float(s),
                                    int(bool(accumulate)), ptr(dA_chan_scale), ptr(ws), ws.numel(), stream_ptr()))
    return dh


def lora_backward_1x1(gy: torch.Tensor, x: torch.Tensor | None, h: torch.Tensor | None, wt: torch.Tensor,
                      a: torch.Tensor, b: torch.Tensor, s: float, dA: torch.Tensor | None, dB: torch.Tensor | None,
                      accumulate: bool = False, bt: torch.Tensor | None = None, return_g: bool = False):
    """Whole backward of a 1x1 / Linear adapter layer: returns dx = gy @ W + s*(gy @ B) @ A and fills dA / dB (fp32;
    both None: frozen adapter, only dx).  One fused tcgen05 launch + one reduction launch.
    gy [.., Cout]; x [.., Cin]; h [M, r] fp32 from the forward; wt [Cin, 1, Cout]; a [r, 1, Cin]; b [Cout, r];
    bt [r, Cout] = b^T when the caller has it (LoRANetwork.pack_transposes), else it is packed here."""
    require_cuda(gy, x, h, wt, a, b, dA, dB, bt)
    assert gy.is_contiguous() and wt.is_contiguous() and a.is_contiguous() and b.is_contiguous()
    B, H, W = _act_dims(gy)
    rank, cin, cout = b.shape[1], wt.shape[0], gy.shape[-1]
    g = make_geom(gy.dtype, B, H, W, cin, cout, 1, rank)
    dx = torch.empty(*gy.shape[:-1], cin, device=gy.device, dtype=gy.dtype)
    gbuf = torch.empty(B * H * W, rank, device=gy.device, dtype=torch.float32)
    if dA is not None:
        assert x is not None and h is not None and x.is_contiguous() and h.is_contiguous()
        assert dA.dtype == torch.float32 and dB.dtype == torch.float32 and dA.is_contiguous() and dB.is_contiguous()
    ws = _grad_workspace(gy.device, lib().slider_lora_grad_workspace_bytes(C.byref(g)))
    check(lib().slider_lora_backward_1x1(C.byref(g), ptr(dx), ptr(dA), ptr(dB), ptr(gbuf), ptr(gy), ptr(x), ptr(h),
                                         ptr(wt), ptr(bt), ptr(a), ptr(b), float(s), int(bool(accumulate)), ptr(ws),
                                         ws.numel(), stream_ptr()))
    return (dx, gbuf) if return_g else dx


def dora_prepare(w: torch.Tensor, a: torch.Tensor, b: torch.Tensor, m: torch.Tensor, s: float, ksize: int,
                 w_eff: torch.Tensor | None = None, a_s: torch.Tensor | None = None):
    """DoRA operands of one wrapped layer (dora.py:122-131): returns (W_eff = W*v, A_s = A*(s c), c, v, 1/n) with
    c = m / ||W + B.A|| per input channel, v = 1 - s + s*c.  w [Cout,k*k,Cin], a [r,k*k,Cin], b [Cout,r], m [Cin]."""
    require_cuda(w, a, b, m, w_eff, a_s)
    assert w.is_contiguous() and a.is_contiguous() and b.is_contiguous() and m.is_contiguous() and m.dtype == w.dtype
    cout, taps, cin = w.shape
    g = make_geom(w.dtype, 1, 1, 1, cin, cout, ksize, a.shape[0])
    w_eff = torch.empty_like(w) if w_eff is None else w_eff
    a_s = torch.empty_like(a) if a_s is None else a_s
    c, v, ninv = (torch.empty(cin, device=w.device, dtype=torch.float32) for _ in range(3))
    ws = _grad_workspace(w.device, lib().slider_lora_grad_workspace_bytes(C.byref(g)))
    check(lib().slider_dora_prepare(C.byref(g), ptr(w_eff), ptr(a_s), ptr(c), ptr(v), ptr(ninv), ptr(w), ptr(a), ptr(b),
                                    ptr(m), float(s), ptr(ws), ws.numel(), stream_ptr()))
    return w_eff, a_s, c, v, ninv


def dora_finish(u: torch.Tensor, t1: torch.Tensor | None, x: torch.Tensor | None, gbuf: torch.Tensor | None,
                a: torch.Tensor | None, c: torch.Tensor, v: torch.Tensor, ninv: torch.Tensor, s: float,
                want_dx: bool = True, dm: torch.Tensor | None = None, accumulate: bool = False):
    """Tail of the DoRA backward: dx = v*T1 + s*c*T2 (T1 = gy.W, T2 = u - T1) and dm_i (+)= s/n_i * sum_p x*u.
    u, t1, x [.., Cin]; gbuf [M, r] fp32 and a [r, 1, Cin] replace t1 for 1x1 layers.  Returns dx (or None)."""
    require_cuda(u, t1, x, gbuf, a, c, v, ninv, dm)
    assert u.is_contiguous() and (t1 is None or t1.is_contiguous()) and (x is None or x.is_contiguous())
    cin = u.shape[-1]
    M = u.numel() // cin
    rank = 0 if a is None else a.shape[0]
    g = make_geom(u.dtype, 1, 1, M, cin, 8, 1 if t1 is None else 3, rank)
    dx = torch.empty_like(u) if want_dx else None
    ws = _grad_workspace(u.device, lib().slider_lora_grad_workspace_bytes(C.byref(g)))
    check(lib().slider_dora_finish(C.byref(g), ptr(dx), ptr(dm), ptr(u), ptr(t1), ptr(x), ptr(gbuf), ptr(a), ptr(c),
                                   ptr(v), ptr(ninv), float(s), int(bool(accumulate)), ptr(ws), ws.numel(), stream_ptr()))
    return dx


def transpose_batch(table: torch.Tensor, entries: int, max_elems: int, dtype: torch.dtype) -> None:
    """One launch of `entries` small transposes described by a device table of slider_transpose_entry records."""
    require_cuda(table)
    check(lib().slider_transpose_batch(ptr(table), int(entries), int(max_elems), dtype_code(dtype), stream_ptr()))


def adamw_flat(p: torch.Tensor, grad: torch.Tensor, m: torch.Tensor, v: torch.Tensor, lr: float, beta1: float,
               beta2: float, eps: float, weight_decay: float, step: int,
               grad_scale: torch.Tensor | None = None) -> None:
    """In-place AdamW over flat buffers (train_lora.py:80-89; train_lora_xl.py:104-105)."""
    require_cuda(p, grad, m, v, grad_scale)
    assert p.is_contiguous() and grad.dtype == m.dtype == v.dtype == torch.float32
    check(lib().slider_adamw_flat(ptr(p), ptr(grad), ptr(m), ptr(v), p.numel(), float(lr), float(beta1),
                                  float(beta2), float(eps), float(weight_decay), int(step), ptr(grad_scale),
                                  dtype_code(p.dtype), stream_ptr()))


def clip_coef(grad: torch.Tensor, max_norm: float, pre_scale: float = 1.0) -> torch.Tensor:
    """Device scalar min(1, max_norm/(||grad||*pre_scale+1e-6))*pre_scale (train_lora_xl.py:349)."""
    require_cuda(grad)
    assert grad.dtype == torch.float32 and grad.is_contiguous()
    out = torch.empty((), device=grad.device, dtype=torch.float32)
    ws = _reduce_ws(grad.device)
    check(lib().slider_clip_coef(ptr(out), ptr(grad), grad.numel(), float(max_norm), float(pre_scale), ptr(ws),
                                 ws.numel(), stream_ptr()))
    return out
