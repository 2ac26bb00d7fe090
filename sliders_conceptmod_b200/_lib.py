"""ctypes binding of libslider_b200.so (the C ABI declared in include/slider_b200.h).

There is no CPU or PyTorch fallback: if the shared library is missing or a call fails,
an exception is raised.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch

PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = PKG_DIR / "libslider_b200.so"

SLIDER_F32, SLIDER_BF16, SLIDER_F16 = 0, 1, 2


class LayerGeom(C.Structure):
    """Mirror of `slider_layer_geom`."""

    _fields_ = [(n, C.c_int32) for n in
                ("dtype", "batch", "height", "width", "cin", "cout", "kh", "kw", "pad", "rank", "stride")]


_vp, _fp, _i64, _f, _i, _sz = C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_int, C.c_size_t
_G = C.POINTER(LayerGeom)

_SIGNATURES = {
    "slider_version": (C.c_int, []),
    "slider_last_error": (C.c_char_p, []),
    "slider_launch_count": (C.c_int64, []),
    "slider_gemm_workspace_bytes": (_sz, []),
    "slider_debug_gemm_trace": (C.c_int, [_vp]),
    "slider_lora_forward": (C.c_int, [_G, _vp, _vp, _vp, _vp, _vp, _vp, _fp, _f, _vp, _sz, _vp]),
    "slider_lora_forward_ex": (C.c_int, [_G, _vp, _vp, _vp, _vp, _vp, _vp, _fp, _f, C.c_uint, _vp, _sz, _vp]),
    "slider_linear_geglu": (C.c_int, [_G, _vp, _vp, _vp, _vp, _vp]),
    "slider_lora_forward_grouped": (C.c_int, [_G, _vp, _vp, _vp, _vp, _vp, _vp, _i, _f, _vp, _sz, _vp]),
    "slider_lora_forward_small_multi": (C.c_int, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "slider_lora_grad_workspace_bytes": (_sz, [_G]),
    "slider_lora_grad": (C.c_int, [_G, _fp, _fp, _vp, _vp, _vp, _fp, _vp, _f, _i, _vp, _sz, _vp]),
    "slider_lora_grad_ex": (C.c_int, [_G, _fp, _fp, _vp, _vp, _vp, _fp, _vp, _f, _i, _fp, _vp, _sz, _vp]),
    "slider_dora_prepare": (C.c_int, [_G, _vp, _vp, _fp, _fp, _fp, _vp, _vp, _vp, _vp, _f, _vp, _sz, _vp]),
    "slider_dora_finish": (C.c_int, [_G, _vp, _fp, _vp, _vp, _vp, _fp, _vp, _fp, _fp, _fp, _f, _i, _vp, _sz, _vp]),
    "slider_lora_backward_1x1": (C.c_int, [_G, _vp, _fp, _fp, _fp, _vp, _vp, _fp, _vp, _vp, _vp, _vp, _f, _i, _vp, _sz, _vp]),
    "slider_transpose_batch": (C.c_int, [_vp, _i, _i, _i, _vp]),
    "slider_lora_pack_at": (C.c_int, [_G, _vp, _vp, _vp]),
    "slider_lora_input_grad": (C.c_int, [_G, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "slider_pack_wt": (C.c_int, [_G, _vp, _vp, _vp]),
    "slider_groupnorm_workspace_bytes": (_sz, [_i, _i]),
    "slider_groupnorm_nhwc_train": (C.c_int, [_vp, _fp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _i, _i, _vp, _sz, _vp]),
    "slider_groupnorm_bwd_workspace_bytes": (_sz, [_i, _i, _i]),
    "slider_groupnorm_nhwc_backward": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _fp, _i, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "slider_groupnorm_nhwc": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _i, _i, _vp, _sz, _vp]),
    "slider_layernorm": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i, _f, _i, _vp]),
    "slider_geglu": (C.c_int, [_vp, _vp, _i64, _i, _i, _vp]),
    "slider_layernorm_train": (C.c_int, [_vp, _fp, _vp, _vp, _vp, _i64, _i, _f, _i, _vp]),
    "slider_layernorm_backward": (C.c_int, [_vp, _vp, _vp, _vp, _fp, _i64, _i, _i, _vp]),
    "slider_geglu_backward": (C.c_int, [_vp, _vp, _vp, _i64, _i, _i, _vp]),
    "slider_upsample_nearest2x_nhwc": (C.c_int, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "slider_concat_channels_nhwc": (C.c_int, [_vp, _vp, _vp, _i64, _i, _i, _i, _vp]),
    "slider_cross_attention": (C.c_int, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "slider_cross_attention_backward": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "slider_cross_attention_strided": (C.c_int, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "slider_loss_workspace_bytes": (_sz, []),
    "slider_guidance_loss": (C.c_int, [_fp, _vp, _vp, _vp, _vp, _vp, _i64, _f, _i, _i, _vp, _sz, _vp]),
    "slider_guidance_loss_noalloc": (C.c_int, [_fp, _vp, _vp, _vp, _vp, _vp, _i64, _f, _i, _i, _vp]),
    "slider_guidance_loss_noalloc_max_elems": (_i64, []),
    "slider_cfg_step": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _f, _f, _f, _f, _fp, _i, _vp]),
    "slider_scale_by_device_scalar": (C.c_int, [_vp, _fp, _i64, _i, _vp]),
    "slider_adamw_flat": (C.c_int, [_vp, _fp, _fp, _fp, _i64, _f, _f, _f, _f, _f, _i, _fp, _i, _vp]),
    "slider_clip_coef": (C.c_int, [_fp, _fp, _i64, _f, _f, _vp, _sz, _vp]),
}

EXPORTS = tuple(_SIGNATURES)

_lib = None


def lib() -> C.CDLL:
    """Load the library (once). Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -m sliders_conceptmod_b200.build` "
                "(there is no fallback path)")
        h = C.CDLL(str(LIB_PATH))
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(h, name)
            fn.restype, fn.argtypes = res, args
        _lib = h
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise RuntimeError(f"libslider_b200 error {rc}: {lib().slider_last_error().decode()}")


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.float32:
        return SLIDER_F32
    if dt == torch.bfloat16:
        return SLIDER_BF16
    if dt == torch.float16:      # forward / inference entry points only; the training entry points reject it
        return SLIDER_F16
    raise TypeError(f"libslider_b200 supports float32, bfloat16 and (inference) float16, got {dt}")


def ptr(t: torch.Tensor | None):
    return None if t is None else t.data_ptr()


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def require_cuda(*tensors: torch.Tensor | None) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("sliders_conceptmod_b200 kernels need CUDA tensors (sm_100a); no CPU path exists")
