"""Generates tests/golden/reference_minitrain.pt by RUNNING THE REFERENCE's own training loop
(`train()` of /root/reference/conceptmod/textsliders/train_lora.py, unmodified, imported read-only) on the CPU
against the random-init UNet stand-in.  Run in the authoring container only; the vectors travel, the reference
does not.

    python tests/golden/make_golden_minitrain.py

What executes from the reference: train_lora.train (the loop, :32-340), lora.LoRANetwork / LoRAModule (fork),
train_util.{predict_noise, diffusion, concat_embeddings, get_initial_latents, get_optimizer, get_lr_scheduler},
prompt_util.{PromptSettings, PromptEmbedsPair, PromptEmbedsCache, load_prompts_from_yaml} (the upstream dialect in
trainscripts/imagesliders — the fork's train_lora.py only works with that one), config_util.

What is substituted (cannot exist here: no diffusers, no pretrained weights):
  model_util.load_models -> the tiny UNet stand-in (sliders_conceptmod_b200.unet, torch arithmetic on CPU), the
                            synthetic text encoder, and the oracle's DDIM scheduler object
  train_util.encode_prompts -> synthetic encoder
  debug_util.check_training_mode -> hook: gives lora_up N(0, 0.05) values (so the adapter is active from the first
                            iteration), snapshots the initial LoRA state, reseeds torch (LOOP_SEED) so a replay can
                            start from the same generator state

Pinned: per-iteration losses, initial and final LoRA state dicts, a checksum of the UNet parameters.
"""
import importlib.util
import sys
import tempfile
import types
from pathlib import Path

import torch
import yaml

REF = Path("/root/reference")
HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))

LOOP_SEED = 4321
ITERATIONS = 8
MAX_DENOISING_STEPS = 10
LR = 1e-3

PROMPTS = [
    {"target": "person", "positive": "old person", "unconditional": "young person", "neutral": "person",
     "action": "enhance", "guidance_scale": 4, "resolution": 128, "batch_size": 1, "dynamic_resolution": False},
    {"target": "dog", "positive": "big dog", "unconditional": "small dog", "neutral": "dog",
     "action": "erase", "guidance_scale": 2, "resolution": 128, "batch_size": 1, "dynamic_resolution": False},
]


def config_dict(prompts_file: str, save_path: str) -> dict:
    return {"prompts_file": prompts_file,
            "pretrained_model": {"name_or_path": "tiny-sd", "v2": False, "v_pred": False},
            "network": {"type": "c3lier", "rank": 4, "alpha": 1.0, "training_method": "noxattn"},
            "train": {"precision": "float32", "noise_scheduler": "ddim", "iterations": ITERATIONS, "lr": LR,
                      "optimizer": "AdamW", "lr_scheduler": "constant", "max_denoising_steps": MAX_DENOISING_STEPS},
            "save": {"name": "mini", "path": save_path, "per_steps": 500, "precision": "float32"},
            "logging": {"use_wandb": False, "verbose": False}, "other": {"use_xformers": True}}


def unet_checksum(unet) -> float:
    return float(sum(p.detach().double().abs().sum() for p in unet.parameters()))


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def main():
    from make_golden import _stub_diffusers
    _stub_diffusers()
    sys.modules.setdefault("wandb", types.ModuleType("wandb"))
    from oracle.slider_oracle import OracleScheduler
    from sliders_conceptmod_b200 import model_util as my_model_util

    record = {"losses": []}

    # flat module names exactly as train_lora.py imports them (:15-22)
    sys.path.insert(0, str(REF))                                   # for `from conceptmod.textsliders...`
    _load("lora", REF / "conceptmod/textsliders/lora.py")
    up = REF / "trainscripts/imagesliders"
    ref_prompt = _load("prompt_util", up / "prompt_util.py")
    ref_config = _load("config_util", up / "config_util.py")
    sys.path.insert(0, str(up))
    ref_train_util = _load("train_util", up / "train_util.py")
    ref_train_util.encode_prompts = lambda tok, enc, prompts: enc.encode(list(prompts))

    stub_model_util = types.ModuleType("model_util")

    def load_models(name, scheduler_name, v2=False, v_pred=False):
        tok, enc, unet, _ = my_model_util.load_models(name, scheduler_name=scheduler_name, v2=v2, v_pred=v_pred)
        record["unet_checksum"] = unet_checksum(unet)
        return tok, enc, unet, OracleScheduler("ddim")
    stub_model_util.load_models = load_models
    sys.modules["model_util"] = stub_model_util

    stub_debug = types.ModuleType("debug_util")
    stub_debug.check_requires_grad = lambda network: None

    def hook(network):
        g = torch.Generator().manual_seed(99)
        with torch.no_grad():
            for m in network.unet_loras:
                m.lora_up.weight.copy_(torch.randn(m.lora_up.weight.shape, generator=g) * 0.05)
        record["init"] = {k: v.detach().clone() for k, v in network.state_dict().items()}
        torch.manual_seed(LOOP_SEED)
    stub_debug.check_training_mode = hook
    sys.modules["debug_util"] = stub_debug

    orig_loss = ref_prompt.PromptEmbedsPair.loss

    def loss_spy(self, **kw):
        out = orig_loss(self, **kw)
        record["losses"].append(float(out.detach()))
        return out
    ref_prompt.PromptEmbedsPair.loss = loss_spy

    ref_train = _load("ref_train_lora", REF / "conceptmod/textsliders/train_lora.py")

    with tempfile.TemporaryDirectory() as tmp:
        pfile = Path(tmp) / "prompts.yaml"
        pfile.write_text(yaml.safe_dump(PROMPTS))
        cfile = Path(tmp) / "config.yaml"
        cfile.write_text(yaml.safe_dump(config_dict(str(pfile), str(Path(tmp) / "out"))))
        config = ref_config.load_config_from_yaml(str(cfile))
        prompts = ref_prompt.load_prompts_from_yaml(config.prompts_file, [])
        torch.manual_seed(1234)
        ref_train.train(config, prompts, "cpu")
        record["final"] = torch.load(Path(tmp) / "out" / "mini_last.pt")

    out = {"losses": record["losses"], "init": record["init"], "final": record["final"],
           "unet_checksum": record["unet_checksum"], "loop_seed": LOOP_SEED, "prompts": PROMPTS,
           "config": config_dict("prompts.yaml", "out")}
    torch.save(out, HERE / "reference_minitrain.pt")
    delta = torch.cat([(out["final"][k].float() - out["init"][k].float()).flatten() for k in out["init"]])
    print("wrote", HERE / "reference_minitrain.pt")
    print("losses", ["%.6e" % l for l in out["losses"]])
    print("n tensors", len(out["init"]), "delta norm %.4e" % delta.norm().item(), "unet checksum %.6f" % out["unet_checksum"])


if __name__ == "__main__":
    sys.path.insert(0, str(HERE))
    main()
