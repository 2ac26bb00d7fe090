"""DoRA (reference conceptmod/textsliders/dora.py:52-163) folded onto the fused base + low-rank kernel — host algebra
of a "next" row; `--peft_type dora` is not wired to it yet.

The reference forms W' = (W + B.A) * c with ONE factor per input channel, c_i = dora_scale_i / ||(W + B.A)[:, i]||
(norm over outputs and kernel taps, detached), and returns org(x) + s * ops(x, W' - W), s = scale * multiplier: a second
dense GEMM per wrapped module.  Because c scales input channels,

    y = (x * v) . W^T + bias + ((x * v) . A'^T) . B^T,      v = 1 - s + s*c,      A'[:, i] = A[:, i] * s*c_i / v_i

i.e. `slider_lora_forward` runs DoRA unchanged on the pre-scaled input x * v with the rescaled down-projection A' and
adapter scale 1 (the r x K rescale is per step; x * v folds into the gamma / beta of the norm that feeds the module).
"""
from __future__ import annotations

import torch


def column_norm(weight: torch.Tensor, lora_down_w: torch.Tensor, lora_up_w: torch.Tensor) -> torch.Tensor:
    """||(W + B.A)[:, i]|| per input channel i (over outputs and taps), shape [in] (dora.py:124-131)."""
    lw = (lora_up_w.flatten(start_dim=1) @ lora_down_w.flatten(start_dim=1)).reshape(weight.shape)
    w = weight + lw
    return w.transpose(0, 1).reshape(w.shape[1], -1).norm(dim=1)


def dora_fold(weight: torch.Tensor, lora_down_w: torch.Tensor, lora_up_w: torch.Tensor, dora_scale: torch.Tensor,
              s: float):
    """Returns (v [in], A' shaped like lora_down_w) of the identity in the module docstring.  Raises where some
    v_i == 0 (the fold divides by it; the two-GEMM form has to be used for that step)."""
    c = dora_scale.reshape(-1) / column_norm(weight, lora_down_w, lora_up_w)
    v = 1.0 - s + s * c
    if bool((v == 0).any()):
        raise ZeroDivisionError("dora_fold: 1 - s + s*c vanishes for an input channel")
    ratio = (s * c / v).reshape(1, -1, *[1] * (lora_down_w.dim() - 2))
    return v, lora_down_w * ratio
