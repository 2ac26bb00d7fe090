"""Merge a UNet / transformer slider LoRA with up to two text-encoder LoRAs into one checkpoint — drop-in for
conceptmod/notrigger/combine_loras.py:17-88 (same positional CLI, same key renaming, same scaling rule).

Host-side file utility (the one edge that needs no pretrained weights): no GPU work.  Per source file and
target family `t`:
  'transformer'  "lora_unet-" -> "transformer.", "_down" -> "_A", "_up" -> "_B", "-" -> "."   (combine_loras.py:28-32)
  'CLIP'         "-" -> "_"                                                                   (:33-34)
  'T5'           "-" -> "_", "lora_te2_" -> "lora_te3_"                                       (:35-37)
and per tensor: `alpha` / `dora_scale` / up-projection (`_up`, `_A`) tensors are copied, down-projection tensors
(`_down`, `_B`) are multiplied by the family's strength (:47-62; the test order is the reference's, so a renamed
"..._A" key that came from "_down" is still classified by its ORIGINAL name).
"""
from __future__ import annotations

import argparse
from typing import Dict, Optional, Tuple

import torch

_RENAMES = {
    "transformer": (("lora_unet-", "transformer."), ("_down", "_A"), ("_up", "_B"), ("-", ".")),
    "CLIP": (("-", "_"),),
    "T5": (("-", "_"), ("lora_te2_", "lora_te3_")),
}


class BinDataHandler:
    """`get_tensor` view over a torch.load()-ed state dict, so `.bin` files read like safetensors handles."""

    def __init__(self, data):
        self.data = data

    def get_tensor(self, key):
        return self.data[key]


def read_tensors(file_path: str):
    """(handle with .get_tensor(key), set of keys) for a .safetensors or .bin file; (None, None) otherwise."""
    if file_path.endswith(".safetensors"):
        from safetensors.torch import safe_open
        f = safe_open(file_path, framework="pt", device="cpu")
        return f, set(f.keys())
    if file_path.endswith(".bin"):
        data = torch.load(file_path, map_location=torch.device("cpu"))
        return BinDataHandler(data), set(data.keys())
    return None, None


def rename_key(key: str, t: str) -> str:
    for old, new in _RENAMES.get(t, ()):
        key = key.replace(old, new)
    return key


def merge(tensor_map: Dict[str, torch.Tensor], tensor_map1: Tuple, strength: float, t: str, verbose: bool = False) -> None:
    """Adds every tensor of `tensor_map1 = (handle, keys)` to `tensor_map` under its renamed key."""
    handle, keys = tensor_map1
    for k in keys:
        k2 = rename_key(k, t)
        value = handle.get_tensor(k)
        if verbose:
            print("BEFORE", k, k2, value.abs().sum())
        if "alpha" in k or "dora_scale" in k or "_up" in k or "_A" in k:
            tensor_map[k2] = value.clone()
        elif "_down" in k or "_B" in k:
            tensor_map[k2] = strength * value.clone()
        else:
            raise AssertionError(k + " not supported")


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(description="Merge two safetensor model files.")
    parser.add_argument("unet_model", type=str, help="The unet model safetensor file")
    parser.add_argument("encoder1_model", type=str, help="The encoder1 model safetensor file")
    parser.add_argument("encoder2_model", type=str, help="The encoder2 model safetensor file")
    parser.add_argument("output_model", type=str, help="The output merged model safetensor file")
    parser.add_argument("unet_strength", type=float, default=0.8, help="Strength of the unet")
    parser.add_argument("enc_strength", type=float, default=1, help="Strength of the first encoder")
    parser.add_argument("enc2_strength", type=float, default=1, help="Strength of the second encoder")
    return parser


def combine(unet_model: str, encoder1_model: str, encoder2_model: str, output_model: Optional[str], unet_strength: float,
            enc_strength: float, enc2_strength: float, verbose: bool = False) -> Dict[str, torch.Tensor]:
    tensor_map: Dict[str, torch.Tensor] = {}
    for path, strength, family in ((unet_model, unet_strength, "transformer"), (encoder1_model, enc_strength, "CLIP"),
                                   (encoder2_model, enc2_strength, "T5")):
        merge(tensor_map, read_tensors(path), strength, family, verbose=verbose)
    if output_model is not None:
        from safetensors.torch import save_file
        save_file(tensor_map, output_model)
    return tensor_map


def main(argv=None):
    a = build_parser().parse_args(argv)
    combine(a.unet_model, a.encoder1_model, a.encoder2_model, a.output_model, a.unet_strength, a.enc_strength,
            a.enc2_strength, verbose=True)


if __name__ == "__main__":
    main()
