"""LoRAModule / LoRANetwork — host-side mirror of the reference adapter API.

Reference: conceptmod/textsliders/lora.py:76-301 (fork) and
trainscripts/imagesliders/lora.py:49-256 (upstream).  Same constructor arguments, attribute
names (`lora_down`, `lora_up`, `scale`, `multiplier`, buffer `alpha`), forward-patching semantics,
naming of the saved keys, `with network:` / `set_lora_slider` behaviour.  The arithmetic of
`LoRAModule.forward` and its autograd is NOT done with torch ops: it is one call into
libslider_b200 (fused base+low-rank tcgen05 GEMM) and, in backward, the low-rank gradient kernels
plus the extended-K input-gradient GEMM.
"""
from __future__ import annotations

import contextlib
import math
import os
import types
from typing import List, Literal, Optional

import torch
import torch.nn as nn

from . import ops

UNET_TARGET_REPLACE_MODULE_TRANSFORMER = ["Attention"]
UNET_TARGET_REPLACE_MODULE_CONV = ["ResnetBlock2D", "Downsample2D", "Upsample2D", "DownBlock2D", "UpBlock2D"]
LORA_PREFIX_UNET = "lora_unet"
DEFAULT_TARGET_REPLACE = UNET_TARGET_REPLACE_MODULE_TRANSFORMER

TRAINING_METHODS = Literal[
    "noxattn", "innoxattn", "selfattn", "xattn", "full", "xattn-strict", "noxattn-hspace", "noxattn-hspace-last",
]


class _FusedLoRAFn(torch.autograd.Function):
    """y = conv(x, W) + b + s*(conv(x, A) . B^T) on [B,H,W,C] / [M,K] activations (lora.py:134-138)."""

    @staticmethod
    def forward(ctx, x, a_w, b_w, mod: "LoRAModule"):
        s = float(mod.multiplier) * float(mod.scale)
        k = mod._ksize
        a_p, b_p = mod._packed_down(a_w), mod._packed_up(b_w)
        need_a, need_b = ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        train_lora = (need_a or need_b) and s != 0.0
        y, h = ops.lora_forward(x, mod._packed_base(), mod._bias(), a_p, b_p, s, ksize=k, want_h=train_lora,
                                stride=mod._stride)
        ctx.mod, ctx.s, ctx.train_lora, ctx.x_shape = mod, s, train_lora, tuple(x.shape)
        ctx.save_for_backward(x if train_lora else None, h, a_p, b_p)
        return y

    @staticmethod
    def backward(ctx, gy):
        mod, s = ctx.mod, ctx.s
        x, h, a_p, b_p = ctx.saved_tensors
        gy = gy.contiguous()
        k, r, stride = mod._ksize, mod.lora_dim, mod._stride
        need_x = ctx.needs_input_grad[0]
        dx = ga = gb = dA = dB = None
        flat = mod._flat_grads
        if ctx.train_lora:
            if flat is not None:
                dA, dB = flat
            else:
                dA = torch.empty(a_p.shape, device=gy.device, dtype=torch.float32)
                dB = torch.empty(b_p.shape, device=gy.device, dtype=torch.float32)
        acc = ctx.train_lora and flat is not None
        if k == 1 and need_x and s != 0.0:
            # 1x1 / Linear: dx and the rank-r gradient g = gy.B from ONE launch, dA and dB from one more
            dx = ops.lora_backward_1x1(gy, x, h, mod._packed_base_t(), a_p, b_p, s, dA, dB, accumulate=acc,
                                       bt=mod._bt_ready())
        else:
            dh = None
            if s != 0.0 and (ctx.train_lora or need_x):      # a frozen adapter still contributes s*dh.A to dx
                xs = x if x is not None else gy.new_empty(ctx.x_shape)       # shape carrier only (dA is None)
                dh = ops.lora_grad(gy, xs, h, b_p, s, k, dA, dB, accumulate=acc, stride=stride)
            if need_x:
                if stride == 2:
                    # transposed stride-2 convolution = stride-1 input gradient of the zero-stuffed output gradient
                    # (three Downsample2D layers, once per training step)
                    Bn, H, W = ctx.x_shape[0], ctx.x_shape[1], ctx.x_shape[2]
                    Ho, Wo = gy.shape[1], gy.shape[2]
                    gz = gy.new_zeros(Bn, H, W, gy.shape[-1])
                    gz[:, ::2, ::2] = gy
                    gy = gz
                    if dh is not None:
                        dz = dh.new_zeros(Bn, H, W, 16)
                        dz[:, ::2, ::2] = dh.view(Bn, Ho, Wo, 16)
                        dh = dz.view(-1, 16)
                at = ops.pack_at(a_p, k) if dh is not None else None
                dx = ops.lora_input_grad(gy, mod._packed_base_t(), dh, at, mod._cin, k, r if dh is not None else 0)
        if ctx.train_lora and flat is None:
            ga = mod._unpack_down_grad(dA)
            gb = dB.to(b_p.dtype).reshape(mod.lora_up.weight.shape)
        return dx, ga, gb, None


class LoRAModule(nn.Module):
    """replaces forward method of the original Linear / Conv2d, instead of replacing the module
    (reference lora.py:76-138)."""

    def __init__(self, lora_name, org_module: nn.Module, multiplier=1.0, lora_dim=4, alpha=None):
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

        if type(alpha) == torch.Tensor:
            alpha = alpha.detach().numpy()
        alpha = lora_dim if alpha is None or alpha == 0 else alpha
        self.scale = alpha / self.lora_dim
        self.register_buffer("alpha", torch.tensor(alpha))

        nn.init.kaiming_uniform_(self.lora_down.weight, a=1)   # fork: a=1; upstream: a=sqrt(5)
        nn.init.zeros_(self.lora_up.weight)

        self.multiplier = multiplier
        self.org_module = org_module  # removed in apply_to (reference keeps it out of state_dict)
        self._flat_grads = None
        self._cache = {}
        self._bt = None        # (state, view [r, Cout]) of lora_up^T, filled by LoRANetwork.pack_transposes()

    def _bt_ready(self):
        """lora_up^T for the fused 1x1 backward when the network has packed it for the current parameters."""
        if self._bt is not None and self._bt[0].valid:
            return self._bt[1]
        return None

    def apply_to(self):
        self.org_forward = self.org_module.forward
        self.org_module.forward = self.forward
        self.__dict__["_org"] = self.org_module       # plain reference, not a registered submodule
        self.org_module.__dict__["_lora_module"] = self   # lets the UNet ask for the residual-fused form
        del self.org_module

    # ---- operand views / one-time re-layouts ------------------------------------------------
    def _bias(self):
        return self._org.bias

    def _packed_base(self) -> torch.Tensor:
        """Frozen base weight as [Cout, k*k, Cin] (tap-major).  Conv weights are switched to
        channels_last storage in place once, so this is a view, not a copy."""
        w = self._org.weight
        key = ("w", w.data_ptr(), w._version, w.dtype)
        hit = self._cache.get("w")
        if hit is not None and hit[0] == key:
            return hit[1]
        if self._is_conv and self._ksize > 1:
            if not w.is_contiguous(memory_format=torch.channels_last):
                w.data = w.data.contiguous(memory_format=torch.channels_last)
            p = w.detach().permute(0, 2, 3, 1).reshape(self._cout, self._ksize * self._ksize, self._cin)
            key = ("w", w.data_ptr(), w._version, w.dtype)
        else:
            p = w.detach().reshape(self._cout, 1, self._cin)
        assert p.is_contiguous()
        self._cache["w"] = (key, p)
        self._cache.pop("wt", None)
        return p

    def _packed_base_t(self) -> torch.Tensor:
        w = self._packed_base()
        hit = self._cache.get("wt")
        if hit is None:
            hit = ops.pack_wt(w, self._ksize)
            self._cache["wt"] = hit
        return hit

    def _packed_down(self, a_w: torch.Tensor) -> torch.Tensor:
        r, T = self.lora_dim, self._ksize * self._ksize
        if not self._is_conv or self._ksize == 1:
            return a_w.detach().reshape(r, 1, self._cin)
        a = a_w.detach().permute(0, 2, 3, 1)
        return (a if a.is_contiguous() else a.contiguous()).reshape(r, T, self._cin)

    def _packed_up(self, b_w: torch.Tensor) -> torch.Tensor:
        return b_w.detach().reshape(self._cout, self.lora_dim)

    def _unpack_down_grad(self, dA: torch.Tensor) -> torch.Tensor:
        w = self.lora_down.weight
        if not self._is_conv:
            return dA.reshape(w.shape).to(w.dtype)
        g = dA.reshape(self.lora_dim, self._ksize, self._ksize, self._cin).permute(0, 3, 1, 2)
        return g.to(w.dtype)

    # ---- forward ----------------------------------------------------------------------------
    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError("LoRAModule.forward runs on sm_100a only (libslider_b200); got a CPU tensor")
        a_w, b_w = self.lora_down.weight, self.lora_up.weight
        if a_w.dtype != x.dtype:
            raise TypeError(f"{self.lora_name}: activation dtype {x.dtype} != LoRA dtype {a_w.dtype}")
        if self._is_conv:
            xn = x.permute(0, 2, 3, 1)
            if not xn.is_contiguous():
                xn = xn.contiguous()
            y = _FusedLoRAFn.apply(xn, a_w, b_w, self)      # stride 2: native (TMA element strides), y is [B,H/2,W/2,C]
            return y.permute(0, 3, 1, 2)
        lead = x.shape[:-1]
        x2 = x.reshape(-1, self._cin)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        y = _FusedLoRAFn.apply(x2, a_w, b_w, self)
        return y.reshape(*lead, self._cout)


def residual_forward(mod: LoRAModule, x: torch.Tensor, residual: torch.Tensor):
    """`residual + mod.forward(x)` with the sum formed by the GEMM's store, IN PLACE of `residual` (no-grad
    forwards only; the caller guarantees nobody else needs `residual`).  Returns None when the layouts do not allow
    it (the caller then adds with torch)."""
    a_w, b_w = mod.lora_down.weight, mod.lora_up.weight
    if a_w.dtype != x.dtype or residual.dtype != x.dtype or mod._stride != 1:
        return None
    s = float(mod.multiplier) * float(mod.scale)
    if mod._is_conv:
        xn, rn = x.permute(0, 2, 3, 1), residual.permute(0, 2, 3, 1)
        if not rn.is_contiguous():
            return None
        if not xn.is_contiguous():
            xn = xn.contiguous()
        y, _ = ops.lora_forward(xn, mod._packed_base(), mod._bias(), mod._packed_down(a_w), mod._packed_up(b_w), s,
                                ksize=mod._ksize, residual=rn)
        return y.permute(0, 3, 1, 2)
    if not residual.is_contiguous() or residual.shape[:-1] != x.shape[:-1]:
        return None
    x2 = x.reshape(-1, mod._cin)
    if not x2.is_contiguous():
        x2 = x2.contiguous()
    if x2.shape[0] <= 16:
        return None
    y, _ = ops.lora_forward(x2, mod._packed_base(), mod._bias(), mod._packed_down(a_w), mod._packed_up(b_w), s,
                            residual=residual.view(-1, mod._cout))
    return y.view(residual.shape)


class _QKVGroup:
    """to_q / to_k / to_v adapters of one self-attention, issued as ONE grouped launch on the no-grad forwards."""

    def __init__(self, loras, a_cat: torch.Tensor, b_cat: torch.Tensor):
        self.loras, self.a_cat, self.b_cat = loras, a_cat, b_cat
        self._w_cat, self._w_key = None, None

    def usable(self, x: torch.Tensor) -> bool:
        q = self.loras[0]
        return (x.is_cuda and not torch.is_grad_enabled() and x.dtype == self.a_cat.dtype
                and x.numel() // x.shape[-1] > 16 and all(m.multiplier == q.multiplier for m in self.loras))

    def forward(self, x: torch.Tensor):
        q = self.loras[0]
        # frozen weights: concatenated once, rebuilt when a base weight is replaced or modified in place
        # (load_state_dict / .to() on the UNet after the first grouped forward)
        key = tuple((m._org.weight.data_ptr(), m._org.weight._version, m._org.weight.dtype) for m in self.loras)
        if self._w_cat is None or self._w_key != key:
            if self._w_cat is not None:
                ops._retired.append(self._w_cat)      # a captured graph may still reference the old buffer
            self._w_cat = torch.cat([m._packed_base() for m in self.loras]).contiguous()
            self._w_key = key
        x2 = x.reshape(-1, q._cin)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        s = float(q.multiplier) * float(q.scale)
        if s == 0.0:        # adapter off (outside `with network:`): one plain GEMM over the concatenated weights
            y, _ = ops.lora_forward(x2, self._w_cat, None, None, None, 0.0)
        else:
            y = ops.lora_forward_grouped(x2, self._w_cat, self.a_cat, self.b_cat, 3, s)
        return y.reshape(*x.shape[:-1], 3 * q._cout).split(q._cout, dim=-1)


class LoRANetwork(nn.Module):
    """Reference conceptmod/textsliders/lora.py:141-301; also accepts the upstream positional form
    `LoRANetwork(unet, rank, multiplier, alpha, train_method)` (trainscripts/imagesliders/lora.py:114-121)."""

    def __init__(self, unet, rank: int = 4, multiplier: float = 1.0, delimiter: str = "_", alpha: float = 1.0,
                 target_replace=DEFAULT_TARGET_REPLACE, prefix=LORA_PREFIX_UNET,
                 train_method: TRAINING_METHODS = "full") -> None:
        super().__init__()
        if not isinstance(delimiter, str):          # upstream call order: (unet, rank, multiplier, alpha, method)
            if isinstance(alpha, str):
                train_method = alpha
            alpha, delimiter = delimiter, "_"
        self.lora_scale = 1
        self.multiplier = multiplier
        self.lora_dim = rank
        self.alpha = alpha
        self.module = LoRAModule
        self.unet_loras = self.create_modules(prefix, unet, target_replace, delimiter, self.lora_dim,
                                              self.multiplier, train_method=train_method)
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

    def create_modules(self, prefix: str, root_module: nn.Module, target_replace_modules: List[str],
                       delimiter: str, rank: int, multiplier: float, train_method: TRAINING_METHODS) -> list:
        loras, names = [], []
        for name, module in root_module.named_modules():
            if train_method in ("noxattn", "noxattn-hspace", "noxattn-hspace-last"):
                if "attn2" in name or "time_embed" in name:
                    continue
            elif train_method == "innoxattn":
                if "attn2" in name:
                    continue
            elif train_method == "selfattn":
                if "attn1" not in name:
                    continue
            elif train_method in ("xattn", "xattn-strict"):
                if "attn2" not in name:
                    continue
            elif train_method == "full":
                pass
            else:
                raise NotImplementedError(f"train_method: {train_method} is not implemented.")
            if module.__class__.__name__ in target_replace_modules:
                for child_name, child_module in module.named_modules():
                    if "add_" in child_name:
                        continue
                    if child_module.__class__.__name__ in ["Linear", "Conv2d", "LoRACompatibleLinear",
                                                           "LoRACompatibleConv"]:
                        if train_method == "xattn-strict" and "out" in child_name:
                            continue
                        if train_method == "noxattn-hspace" and "mid_block" not in name:
                            continue
                        if train_method == "noxattn-hspace-last":
                            if "mid_block" not in name or ".1" not in name or "conv2" not in child_name:
                                continue
                        lora_name = (prefix + "." + name + "." + child_name).replace(".", delimiter)
                        # constructed even when dropped as a duplicate: keeps the RNG stream of the reference
                        lora = self.module(lora_name, child_module, multiplier, rank, self.alpha)
                        if lora_name not in names:
                            loras.append(lora)
                            names.append(lora_name)
        return loras

    def prepare_optimizer_params(self):
        all_params = []
        if self.unet_loras:
            params = []
            [params.extend(lora.parameters()) for lora in self.unet_loras]
            all_params.append({"params": params})
        return all_params

    def get_state_dict(self, dtype=None):
        state_dict = self.state_dict()
        for key in list(state_dict.keys()):
            v = state_dict[key].detach().clone().to("cpu")
            state_dict[key] = v.to(dtype).contiguous() if dtype is not None else v.contiguous()
        return state_dict

    def save_weights(self, file, dtype=None, metadata: Optional[dict] = None):
        state_dict = self.get_state_dict(dtype)
        if os.path.splitext(file)[1] == ".safetensors":
            from safetensors.torch import save_file
            save_file(state_dict, file, metadata)
        else:
            torch.save(state_dict, file)

    def set_lora_slider(self, scale):
        self.lora_scale = scale

    def __enter__(self):
        for lora in self.unet_loras:
            lora.multiplier = 1.0 * self.lora_scale

    def __exit__(self, exc_type, exc_value, tb):
        for lora in self.unet_loras:
            lora.multiplier = 0

    # ---- B200 extension: one contiguous parameter buffer + one fp32 gradient buffer ------------
    def flatten(self):
        """Re-home every LoRA factor in ONE flat parameter buffer and allocate ONE flat fp32 gradient
        buffer with the same element order.  The gradient kernels then accumulate straight into that
        buffer (no per-tensor .grad), so data-parallel training needs a single all-reduce and a single
        fused AdamW launch.  Conv `lora_down` factors are stored tap-major
        (channels_last strides): state_dict()/save_weights() are unchanged.  Call after `.to(device, dtype)`.
        """
        params = [p for lora in self.unet_loras for p in (lora.lora_down.weight, lora.lora_up.weight)]
        dev, dt = params[0].device, params[0].dtype
        total = sum(p.numel() for p in params)
        flat = torch.empty(total, device=dev, dtype=dt)
        gflat = torch.zeros(total, device=dev, dtype=torch.float32)
        # self-attention q/k/v adapters read the same input: lay their factors out as [A_q;A_k;A_v][B_q;B_k;B_v] so
        # that the concatenated operands of the grouped launch (ops.lora_forward_grouped) are plain views
        by_name = {lora.lora_name: lora for lora in self.unet_loras}
        groups, grouped = [], set()
        for lora in self.unet_loras:
            if lora.lora_name.endswith("attn1_to_q"):
                stem = lora.lora_name[:-1]
                trio = [by_name.get(stem + c) for c in "qkv"]
                if all(m is not None and not m._is_conv and m._cin == lora._cin and m._cout == lora._cout
                       and m.lora_dim == lora.lora_dim and m._bias() is None for m in trio) \
                        and 3 * lora.lora_dim <= 16 and lora._cout % 160 == 0:
                    groups.append(trio)
                    grouped.update(m.lora_name for m in trio)
        units = [[m] for m in self.unet_loras if m.lora_name not in grouped] + groups
        off = 0

        def home(lora, which, start):
            w = lora.lora_down.weight if which == "a" else lora.lora_up.weight
            n = w.numel()
            f = flat[start:start + n]
            if which == "a" and lora._is_conv:
                f4 = f.view(lora.lora_dim, lora._ksize, lora._ksize, lora._cin).permute(0, 3, 1, 2)
                f4.copy_(w.data)
                w.data = f4
            else:
                f.view_as(w).copy_(w.data)
                w.data = f.view_as(w)
            g = gflat[start:start + n]
            return n, (g.view(lora.lora_dim, lora._ksize * lora._ksize, lora._cin) if which == "a"
                       else g.view(lora._cout, lora.lora_dim))

        for unit in units:
            start = off
            ga = []
            for lora in unit:
                n, g = home(lora, "a", off)
                ga.append(g); off += n
            mid = off
            for lora, g_a in zip(unit, ga):
                n, g_b = home(lora, "b", off)
                lora._flat_grads = (g_a, g_b); off += n
            if len(unit) == 3:
                r, K, N = unit[0].lora_dim, unit[0]._cin, unit[0]._cout
                group = _QKVGroup(unit, flat[start:mid].view(3 * r, 1, K), flat[mid:off].view(3 * N, r))
                unit[0]._org._qkv_group = group
        self.flat_param, self.flat_grad = flat, gflat
        # lora_up^T of every 1x1 / Linear adapter (the "lora_down" operand of its fused backward launch) lives in ONE
        # buffer that ONE batched-transpose launch rebuilds from the flat parameters (pack_transposes)
        import numpy as np
        ones = [m for m in self.unet_loras if m._ksize == 1]
        self._bt_state = types.SimpleNamespace(valid=False)
        self._bt_buf = torch.empty(sum(m.lora_dim * m._cout for m in ones), device=dev, dtype=dt)
        rec = np.zeros(len(ones), dtype=[("src", "<u8"), ("dst", "<u8"), ("rows", "<i4"), ("cols", "<i4")])
        off = 0
        for i, m in enumerate(ones):
            n = m.lora_dim * m._cout
            view = self._bt_buf[off:off + n].view(m.lora_dim, m._cout)
            m._bt = (self._bt_state, view)
            rec[i] = (m.lora_up.weight.data_ptr(), view.data_ptr(), m._cout, m.lora_dim)
            off += n
        self._bt_entries = len(ones)
        self._bt_max = max([m.lora_dim * m._cout for m in ones] + [1])
        self._bt_table = torch.from_numpy(rec.view(np.uint8).copy()).to(dev) if ones else None
        return flat, gflat

    def pack_transposes(self):
        """Rebuild every 1x1 adapter's lora_up^T from the flat parameter buffer: one launch."""
        if self._bt_table is not None:
            ops.transpose_batch(self._bt_table, self._bt_entries, self._bt_max, self._bt_buf.dtype)

    @contextlib.contextmanager
    def packed_transposes(self):
        """Scope of ONE forward/backward over fixed parameters (a training-step body): the transposes are packed once
        on entry and the per-layer backward launches use them instead of packing their own."""
        if self.flat_param is None or self._bt_table is None:
            yield
            return
        self.pack_transposes()
        self._bt_state.valid = True
        try:
            yield
        finally:
            self._bt_state.valid = False
