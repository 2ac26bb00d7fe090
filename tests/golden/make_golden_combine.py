"""Generates tests/golden/reference_combine.pt by EXECUTING THE REFERENCE's merge()
(/root/reference/conceptmod/notrigger/combine_loras.py:26-62) on seeded tensors.  Authoring container only.

    python tests/golden/make_golden_combine.py
"""
import contextlib
import io
import sys
from pathlib import Path

import torch

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
from make_golden import REF, _load  # noqa: E402


def main():
    ref = _load("ref_combine_loras", REF / "conceptmod/notrigger/combine_loras.py")
    g = torch.Generator().manual_seed(77)
    srcs = {
        "transformer": {"lora_unet-transformer_blocks-0-attn-to_q.lora_down.weight": torch.randn(4, 16, generator=g),
                        "lora_unet-transformer_blocks-0-attn-to_q.lora_up.weight": torch.randn(16, 4, generator=g),
                        "lora_unet-transformer_blocks-0-attn-to_q.alpha": torch.tensor(1.0),
                        "lora_unet-transformer_blocks-0-attn-to_q.dora_scale": torch.randn(1, 16, generator=g)},
        "CLIP": {"lora_te1-text_model-encoder-layers-0-self_attn-q_proj.lora_down.weight": torch.randn(4, 8, generator=g),
                 "lora_te1-text_model-encoder-layers-0-self_attn-q_proj.lora_up.weight": torch.randn(8, 4, generator=g),
                 "lora_te1-text_model-encoder-layers-0-self_attn-q_proj.alpha": torch.tensor(4.0)},
        "T5": {"lora_te2-encoder-block-3-layer-0-SelfAttention-q.lora_down.weight": torch.randn(4, 8, generator=g),
               "lora_te2-encoder-block-3-layer-0-SelfAttention-q.lora_up.weight": torch.randn(8, 4, generator=g),
               "lora_te2-encoder-block-3-layer-0-SelfAttention-q.alpha": torch.tensor(2.0)},
    }
    strengths = {"transformer": 0.8, "CLIP": 1.25, "T5": 0.5}
    out = {}
    with contextlib.redirect_stdout(io.StringIO()):
        for fam, sd in srcs.items():
            ref.merge(out, (ref.BinDataHandler(sd), set(sd.keys())), strengths[fam], fam)
    torch.save({"sources": srcs, "strengths": strengths, "merged": out}, HERE / "reference_combine.pt")
    print("wrote", HERE / "reference_combine.pt", len(out), "tensors")


if __name__ == "__main__":
    main()
