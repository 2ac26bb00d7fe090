"""Generates tests/golden/reference_notrigger.pt by RUNNING THE REFERENCE's own text-encoder trainer
(`train()` of /root/reference/conceptmod/notrigger/train_notrigger.py, unmodified, imported read-only) on the CPU against
the random-init CLIP stand-ins.  Run in the authoring container only; the vectors travel, the reference does not.

    python tests/golden/make_golden_notrigger.py

What executes from the reference: train_notrigger.{train, fixed_distance_loss} (the whole loop, :67-486),
conceptmod.textsliders.lora.LoRANetwork / LoRAModule, train_util.text_tokenize, config_util, torch's SGD / LambdaLR /
ConstantLR, transformers.get_linear_schedule_with_warmup.

What is substituted (cannot exist here: no diffusers, no pretrained weights, no wandb):
  model_util.load_models_xl -> (HashTokenizer x2, tiny CLIPTextModel stand-ins x2, None, None), the same seed on both
                               calls (the reference loads the same checkpoint twice, :96-117)
  debug_util / wandb        -> empty stubs

Pinned: the LoRA state dict returned by train(save_file=False) after ITERATIONS iterations for three cases (both
sides, positive only, negative only) on both encoders, fixed_distance_loss on a random triple, and the outputs of the
reference's text_tokenize / encode_prompts / encode_prompts_xl on the stand-ins.
"""
import importlib.util
import sys
import types
from pathlib import Path

import torch

REF = Path("/root/reference")
HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))

ITERATIONS = 12
LR = 40.0            # the warm-up schedule runs iteration i at lr * i / 100; a large base rate makes 12 iterations move
SEED_MODEL = 7
CASES = {
    "both": dict(positive=["happy", "smiling"], negative=["sad", "frowning"]),
    "positive_only": dict(positive=["old person"], negative=None),
    "negative_only": dict(positive=None, negative=["young person"]),
}


def config_dict() -> dict:
    return {"prompts_file": "unused.yaml",
            "pretrained_model": {"name_or_path": "tiny-clip", "v2": False, "v_pred": False},
            "network": {"type": "c3lier", "rank": 4, "alpha": 1.0, "training_method": "noxattn"},
            "train": {"precision": "float32", "noise_scheduler": "ddim", "iterations": ITERATIONS, "lr": LR,
                      "optimizer": "AdamW", "lr_scheduler": "constant", "max_denoising_steps": 10},
            "save": {"name": "nt", "path": "/tmp/unused", "per_steps": 500, "precision": "float32"},
            "logging": {"use_wandb": False, "verbose": False}, "other": {"use_xformers": True}}


def main():
    from make_golden import _stub_diffusers
    _stub_diffusers()
    sys.modules.setdefault("wandb", types.ModuleType("wandb"))
    from sliders_conceptmod_b200 import text_encoder as te

    sys.path.insert(0, str(REF))
    # conceptmod.textsliders.model_util imports half of diffusers at module level: replaced as a whole
    ref_model_util = types.ModuleType("conceptmod.textsliders.model_util")
    ref_model_util.SDXL_TEXT_ENCODER_TYPE = object

    def load_models_xl(name, scheduler_name="ddim", **kw):
        toks, encs = te.load_text_encoders("SDXL", seed=SEED_MODEL, tiny=True)
        return toks, encs, None, None
    ref_model_util.load_models_xl = load_models_xl
    sys.modules["conceptmod.textsliders.model_util"] = ref_model_util
    import conceptmod.textsliders as ref_pkg
    ref_pkg.model_util = ref_model_util
    import conceptmod.textsliders.config_util as ref_config

    spec = importlib.util.spec_from_file_location("ref_train_notrigger", REF / "conceptmod/notrigger/train_notrigger.py")
    ref = importlib.util.module_from_spec(spec)
    sys.modules["ref_train_notrigger"] = ref
    spec.loader.exec_module(ref)
    ref.flush = lambda: None                    # torch.cuda.empty_cache on a CPU-only box

    out = {"iterations": ITERATIONS, "lr": LR, "seed_model": SEED_MODEL, "config": config_dict(), "cases": {}}
    for name, kw in CASES.items():
        for clip_index in (0, 1):
            config = ref_config.RootConfig(**config_dict())
            torch.manual_seed(1234 + clip_index)                 # governs the LoRA init (kaiming lora_down)
            sd = ref.train(config, [], torch.device("cpu"), on_step_complete=None, save_file=False,
                           clip_index=clip_index, peft_type="lora", rank=4, model="SDXL", **kw)
            out["cases"][f"{name}_te{clip_index + 1}"] = {"kw": kw, "clip_index": clip_index, "torch_seed": 1234 + clip_index,
                                                          "final": {k: v.clone() for k, v in sd.items()}}
            up = torch.cat([v.flatten() for k, v in sd.items() if "lora_up" in k])
            print(name, clip_index, "tensors", len(sd), "lora_up norm %.4e" % up.norm().item())

    g = torch.Generator().manual_seed(5)
    a = torch.randn(1, 77, 32, generator=g, requires_grad=True)
    b = torch.randn(1, 77, 32, generator=g)
    d = torch.tensor(0.37)
    val = ref.fixed_distance_loss(a, b, d)
    grad, = torch.autograd.grad(val, a)
    out["fixed_distance"] = {"a": a.detach(), "b": b, "d": d, "value": val.detach(), "grad": grad}
    # prompt encoding (conceptmod/textsliders/train_util.py:108-154, 224-247) by the reference's own functions
    toks, encs = te.load_text_encoders("SDXL", seed=SEED_MODEL, tiny=True)
    for e in encs:
        e.eval()
    prompts = ["a photo of a person", "old person, wrinkles"]
    with torch.no_grad():
        xl_embeds, xl_pooled = ref.train_util.encode_prompts_xl(toks, encs, prompts, num_images_per_prompt=2)
        sd_embeds = ref.train_util.encode_prompts(toks[0], encs[0], prompts)
    out["encode"] = {"prompts": prompts, "tokens": ref.train_util.text_tokenize(toks[0], prompts), "xl_embeds": xl_embeds,
                     "xl_pooled": xl_pooled, "sd_embeds": sd_embeds}
    print("encode_prompts_xl", tuple(xl_embeds.shape), tuple(xl_pooled.shape), "encode_prompts", tuple(sd_embeds.shape))
    torch.save(out, HERE / "reference_notrigger.pt")
    print("wrote", HERE / "reference_notrigger.pt")


if __name__ == "__main__":
    sys.path.insert(0, str(HERE))
    main()
