"""DoRAModule / DoRANetwork — host-side mirror of conceptmod/textsliders/dora.py:52-330 (`--peft_type dora` of
train_lora_xl.py:87-90), with the arithmetic on libslider_b200.

The reference forms W' = (W + B.A) * c with ONE factor per input channel, c_i = dora_scale_i / ||(W + B.A)[:, i]||
(norm over outputs and kernel taps, detached), and returns org(x) + s * ops(x, W' - W), s = scale * multiplier: a
second dense GEMM per wrapped module plus, in backward, a dense out x in weight gradient.  Because c scales input
channels,

    y = x . (W * v)^T + bias + (x . (A * s*c)^T) . B^T,        v = 1 - s + s*c,

i.e. ONE launch of the fused base + low-rank kernel on the column-scaled operands (`ops.dora_prepare`: two
bandwidth passes over the frozen weight whenever A, B, dora_scale or s changed).  Backward: the low-rank gradients are
the LoRA kernels' (with the per-channel factor s*c on dA), dx and the dora_scale gradient come from the same
input-gradient GEMM u = gy.W + (gy.B).A the LoRA path runs, finished by one elementwise / column-sum kernel
(`ops.dora_finish`) — the dense weight gradient is never formed.
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch
import torch.nn as nn

from . import ops
from .lora import DEFAULT_TARGET_REPLACE, LORA_PREFIX_UNET, TRAINING_METHODS, LoRANetwork


# ---- host algebra (CPU-checkable against the reference's golden vectors, tests/test_oracle_golden.py) -----------
def column_norm(weight: torch.Tensor, lora_down_w: torch.Tensor, lora_up_w: torch.Tensor) -> torch.Tensor:
    """||(W + B.A)[:, i]|| per input channel i (over outputs and taps), shape [in] (dora.py:124-131)."""
    lw = (lora_up_w.flatten(start_dim=1) @ lora_down_w.flatten(start_dim=1)).reshape(weight.shape)
    w = weight + lw
    return w.transpose(0, 1).reshape(w.shape[1], -1).norm(dim=1)


def dora_fold(weight: torch.Tensor, lora_down_w: torch.Tensor, lora_up_w: torch.Tensor, dora_scale: torch.Tensor,
              s: float):
    """Activation-side form of the identity: y = (x*v).W^T + ((x*v).A'^T).B^T with A'[:, i] = A[:, i]*s*c_i/v_i.
    Returns (v [in], A').  Kept as the CPU cross-check of the algebra; the product scales the operands instead
    (`dora_operands`), which needs no division by v."""
    c = dora_scale.reshape(-1) / column_norm(weight, lora_down_w, lora_up_w)
    v = 1.0 - s + s * c
    if bool((v == 0).any()):
        raise ZeroDivisionError("dora_fold: 1 - s + s*c vanishes for an input channel")
    ratio = (s * c / v).reshape(1, -1, *[1] * (lora_down_w.dim() - 2))
    return v, lora_down_w * ratio


def dora_operands(weight: torch.Tensor, lora_down_w: torch.Tensor, lora_up_w: torch.Tensor, dora_scale: torch.Tensor,
                  s: float):
    """Operand-side form used by the CUDA path: (W_eff = W*v, A_s = A*(s c), c, v); torch restatement of
    `ops.dora_prepare` for tests."""
    c = dora_scale.reshape(-1) / column_norm(weight, lora_down_w, lora_up_w)
    v = 1.0 - s + s * c
    shp = (1, -1) + (1,) * (weight.dim() - 2)
    return weight * v.reshape(shp), lora_down_w * (s * c).reshape(shp), c, v


class _DoRAFn(torch.autograd.Function):
    """y = org(x) + s * ops(x, W' - W) on [B,H,W,C] / [M,K] activations (dora.py:145-163)."""

    @staticmethod
    def forward(ctx, x, a_w, b_w, m_w, mod: "DoRAModule"):
        s = float(mod.multiplier) * float(mod.scale)
        k = mod._ksize
        w_eff, a_s, c, v, ninv = mod._prepare(s, a_w, b_w, m_w)
        b_p = mod._packed_up(b_w)
        train = any(ctx.needs_input_grad[1:4])
        y, h = ops.lora_forward(x, w_eff, mod._bias(), a_s, b_p, 1.0, ksize=k, want_h=train, stride=mod._stride)
        ctx.mod, ctx.s, ctx.train, ctx.x_shape = mod, s, train, tuple(x.shape)
        ctx.save_for_backward(x if train else None, h, mod._packed_down(a_w), b_p, c, v, ninv)
        return y

    @staticmethod
    def backward(ctx, gy):
        mod, s = ctx.mod, ctx.s
        x, h, a_p, b_p, c, v, ninv = ctx.saved_tensors
        gy = gy.contiguous()
        k, r, stride = mod._ksize, mod.lora_dim, mod._stride
        need_x = ctx.needs_input_grad[0]
        dx = ga = gb = gm = None
        dh = None
        if ctx.train:
            # low-rank gradients: dB = gy^T . h' (h' = x.(A s c)^T from the forward), dA = (s c) * (gy.B)^T (*) x
            dA = torch.empty(a_p.shape, device=gy.device, dtype=torch.float32)
            dB = torch.empty(b_p.shape, device=gy.device, dtype=torch.float32)
            dh = ops.lora_grad(gy, x, h, b_p, 1.0, k, dA, dB, stride=stride, dA_chan_scale=(s * c).contiguous())
            ga = mod._unpack_down_grad(dA)
            gb = dB.to(b_p.dtype).reshape(mod.lora_up.weight.shape)
        need_m = ctx.train and ctx.needs_input_grad[3]
        if need_x or need_m:
            dm = torch.empty(mod._cin, device=gy.device, dtype=torch.float32) if need_m else None
            wt = mod._packed_base_t()
            if k == 1:
                # u = gy.W + (gy.B).A and g = gy.B from ONE launch; T2 = g.A is formed inside the finish kernel
                u, gbuf = ops.lora_backward_1x1(gy, None, None, wt, a_p, b_p, 1.0, None, None, return_g=True)
                dx = ops.dora_finish(u, None, x, gbuf, a_p, c, v, ninv, s, want_dx=need_x, dm=dm)
            else:
                if dh is None:
                    dh = ops.lora_grad(gy, gy.new_empty(ctx.x_shape), None, b_p, 1.0, k, None, None, stride=stride)
                if stride == 2:     # transposed stride-2 convolution through the zero-stuffed output gradient
                    Bn, H, W = ctx.x_shape[0], ctx.x_shape[1], ctx.x_shape[2]
                    gz = gy.new_zeros(Bn, H, W, gy.shape[-1]); gz[:, ::2, ::2] = gy
                    dz = dh.new_zeros(Bn, H, W, 16); dz[:, ::2, ::2] = dh.view(Bn, gy.shape[1], gy.shape[2], 16)
                    gy, dh = gz, dz.view(-1, 16)
                t1 = ops.lora_input_grad(gy, wt, None, None, mod._cin, k, 0)                       # gy.W
                u = ops.lora_input_grad(gy, wt, dh, ops.pack_at(a_p, k), mod._cin, k, r)          # + (gy.B).A
                dx = ops.dora_finish(u, t1, x, None, None, c, v, ninv, s, want_dx=need_x, dm=dm)
            if need_m:
                gm = dm.to(mod.dora_scale.dtype).reshape(mod.dora_scale.shape)
        return dx, ga, gb, gm, None


class DoRAModule(nn.Module):
    """replaces forward method of the original Linear / Conv2d (reference dora.py:52-163).  Constructor arguments in
    the reference's order: (lora_name, org_module, multiplier, lora_dim, prefix, alpha)."""

    def __init__(self, lora_name, org_module: nn.Module, multiplier=1.0, lora_dim=4, prefix: str = LORA_PREFIX_UNET,
                 alpha=1):
        """if alpha == 0 or None, alpha is rank (no scaling)."""
        super().__init__()
        self.lora_name = lora_name
        self.lora_dim = lora_dim
        cls = org_module.__class__.__name__
        if "Linear" in cls:
            in_dim, out_dim = org_module.in_features, org_module.out_features
            self.lora_down = nn.Linear(in_dim, lora_dim, bias=False)
            self.lora_up = nn.Linear(lora_dim, out_dim, bias=False)
            self._ksize, self._stride, self._is_conv = 1, 1, False
        elif "Conv" in cls:
            in_dim, out_dim = org_module.in_channels, org_module.out_channels
            self.lora_dim = min(self.lora_dim, in_dim, out_dim)
            if self.lora_dim != lora_dim:
                print(f"{lora_name} dim (rank) is changed to: {self.lora_dim}")
            ks, st, pd = org_module.kernel_size, org_module.stride, org_module.padding
            self.stride, self.padding = st, pd
            self.lora_down = nn.Conv2d(in_dim, self.lora_dim, ks, st, pd, bias=False)
            self.lora_up = nn.Conv2d(self.lora_dim, out_dim, (1, 1), (1, 1), bias=False)
            ok = (ks[0] == ks[1] and ks[0] in (1, 3) and st[0] == st[1] and st[0] in (1, 2)
                  and tuple(pd) == ((ks[0] - 1) // 2,) * 2 and org_module.groups == 1
                  and tuple(org_module.dilation) == (1, 1) and (st[0] == 1 or ks[0] == 3))
            if not ok:
                raise NotImplementedError(
                    f"{lora_name}: conv geometry k={ks} s={st} p={pd} is outside the fused kernel's coverage")
            self._ksize, self._stride, self._is_conv = ks[0], st[0], True
        else:
            raise NotImplementedError(f"{lora_name}: cannot wrap {cls}")
        self._cin, self._cout = in_dim, out_dim
        org_weight = org_module.weight
        self.dora_norm_dims = org_weight.dim() - 1
        # per INPUT channel norm of the frozen weight, shaped [1, in, 1, 1] / [1, in] (dora.py:98-107)
        self.dora_scale = nn.Parameter(
            torch.norm(org_weight.transpose(1, 0).reshape(org_weight.shape[1], -1), dim=1, keepdim=True)
            .reshape(org_weight.shape[1], *[1] * self.dora_norm_dims).transpose(1, 0).detach().clone())
        if type(alpha) == torch.Tensor:
            alpha = alpha.detach().numpy()
        alpha = lora_dim if alpha is None or alpha == 0 else alpha
        self.scale = alpha / self.lora_dim
        self.register_buffer("alpha", torch.tensor(alpha))
        nn.init.kaiming_uniform_(self.lora_down.weight, a=1)
        nn.init.zeros_(self.lora_up.weight)
        self.multiplier = multiplier
        self.org_module = org_module
        self._cache = {}
        self._prep = None

    def apply_to(self):
        self.org_forward = self.org_module.forward
        self.org_module.forward = self.forward
        self.org_module_weight = self.org_module.weight.detach()
        self.__dict__["_org"] = self.org_module
        del self.org_module

    # operand views shared with LoRAModule (same layouts)
    _bias = lambda self: self._org.bias

    def _packed_base(self):
        from .lora import LoRAModule
        return LoRAModule._packed_base(self)

    def _packed_base_t(self):
        from .lora import LoRAModule
        return LoRAModule._packed_base_t(self)

    def _packed_down(self, a_w):
        from .lora import LoRAModule
        return LoRAModule._packed_down(self, a_w).contiguous()

    def _packed_up(self, b_w):
        return b_w.detach().reshape(self._cout, self.lora_dim)

    def _unpack_down_grad(self, dA):
        from .lora import LoRAModule
        return LoRAModule._unpack_down_grad(self, dA)

    def _prepare(self, s: float, a_w, b_w, m_w):
        """(W_eff, A_s, c, v, 1/n) for the current parameters and scale; recomputed when any of them changed.  W_eff and
        A_s live in per-module buffers (only the forward launch that follows reads them); c, v, 1/n are fresh small
        tensors per preparation because a backward may need those of an EARLIER forward (two-sided objectives)."""
        key = (s, a_w.data_ptr(), a_w._version, b_w._version, m_w._version, a_w.dtype)
        if self._prep is not None and self._prep[0] == key:
            return self._prep[1]
        w = self._packed_base()
        bufs = self._cache.get("eff")
        if bufs is None or bufs[0].dtype != w.dtype or bufs[0].device != w.device:
            bufs = (torch.empty_like(w), torch.empty(self.lora_dim, w.shape[1], self._cin, device=w.device, dtype=w.dtype))
            self._cache["eff"] = bufs
        out = ops.dora_prepare(w, self._packed_down(a_w), self._packed_up(b_w),
                               m_w.detach().reshape(-1).contiguous(), s, self._ksize, w_eff=bufs[0], a_s=bufs[1])
        self._prep = (key, out)
        return out

    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError("DoRAModule.forward runs on sm_100a only (libslider_b200); got a CPU tensor")
        a_w, b_w, m_w = self.lora_down.weight, self.lora_up.weight, self.dora_scale
        if a_w.dtype != x.dtype:
            raise TypeError(f"{self.lora_name}: activation dtype {x.dtype} != DoRA dtype {a_w.dtype}")
        s = float(self.multiplier) * float(self.scale)
        if self._is_conv:
            xn = x.permute(0, 2, 3, 1)
            if not xn.is_contiguous():
                xn = xn.contiguous()
            if s == 0.0:        # adapter off (outside `with network:`): W' - W is multiplied by zero
                y, _ = ops.lora_forward(xn, self._packed_base(), self._bias(), None, None, 0.0, ksize=self._ksize,
                                        stride=self._stride)
            else:
                y = _DoRAFn.apply(xn, a_w, b_w, m_w, self)
            return y.permute(0, 3, 1, 2)
        lead = x.shape[:-1]
        x2 = x.reshape(-1, self._cin)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        if s == 0.0:
            y, _ = ops.lora_forward(x2, self._packed_base(), self._bias(), None, None, 0.0)
        else:
            y = _DoRAFn.apply(x2, a_w, b_w, m_w, self)
        return y.reshape(*lead, self._cout)


class DoRANetwork(LoRANetwork):
    """Reference conceptmod/textsliders/dora.py:166-330.  Same module selection, naming and save layout as
    LoRANetwork; the constructor keeps the reference's argument ORDER (rank, multiplier, delimiter, alpha, prefix,
    target_replace, train_method) and its create_modules call, which passes `alpha` into DoRAModule's `prefix` slot, so
    every module runs with alpha = 1 (scale = 1 / rank) whatever `alpha` says."""

    def __init__(self, unet, rank: int = 4, multiplier: float = 1.0, delimiter: str = "_", alpha: float = 1.0,
                 prefix: str = LORA_PREFIX_UNET, target_replace=DEFAULT_TARGET_REPLACE,
                 train_method: TRAINING_METHODS = "full") -> None:
        nn.Module.__init__(self)
        self.lora_scale = 1
        self.multiplier = multiplier
        self.lora_dim = rank
        self.alpha = alpha
        self.module = lambda name, child, mult, r, a: DoRAModule(name, child, mult, r, a)     # `a` lands in `prefix`
        self.unet_loras = self.create_modules(prefix, unet, target_replace, delimiter, self.lora_dim, self.multiplier,
                                              train_method=train_method)
        lora_names = set()
        for lora in self.unet_loras:
            assert lora.lora_name not in lora_names, f"duplicated lora name: {lora.lora_name}. {lora_names}"
            lora_names.add(lora.lora_name)
        for lora in self.unet_loras:
            lora.apply_to()
            self.add_module(lora.lora_name, lora)
        self.flat_param = None
        self.flat_grad = None
        del unet

    def flatten(self):
        raise NotImplementedError("DoRANetwork trains through torch optimizers over its parameters (the flat-buffer "
                                  "engine covers LoRANetwork)")

    def get_state_dict(self, dtype=None):           # dora.py:293-301 (no .contiguous(), no cpu move when dtype is None)
        state_dict = self.state_dict()
        if dtype is not None:
            for key in list(state_dict.keys()):
                state_dict[key] = state_dict[key].detach().clone().to("cpu").to(dtype)
        return state_dict

    def save_weights(self, file, dtype=None, metadata: Optional[dict] = None):
        state_dict = self.get_state_dict(dtype)
        if os.path.splitext(file)[1] == ".safetensors":
            from safetensors.torch import save_file
            save_file({k: v.contiguous() for k, v in state_dict.items()}, file, metadata)
        else:
            torch.save(state_dict, file)

    def __enter__(self):                             # dora.py:324-326: no slider scale on this class
        for lora in self.unet_loras:
            lora.multiplier = 1.0
