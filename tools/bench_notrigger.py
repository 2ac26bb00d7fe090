"""Text-encoder LoRA trainer (f4) at the SD-XL encoder sizes, bf16, both sides, random-init stand-ins:
iterations/s of the product path (adapted projections on libslider's fused GEMM + fused backward) next to the same
trainer with the reference's LoRA arithmetic (three torch ops per layer, oracle.attach) in PyTorch eager on the same GPU."""
import json
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import config_util, train_notrigger, ops
from sliders_conceptmod_b200.text_encoder import load_text_encoders
from oracle import slider_oracle

cfg = {"prompts_file": "unused.yaml", "pretrained_model": {"name_or_path": "stand-in", "v2": False, "v_pred": False},
       "network": {"type": "c3lier", "rank": 4, "alpha": 1.0, "training_method": "noxattn"},
       "train": {"precision": "bfloat16", "noise_scheduler": "ddim", "iterations": 1, "lr": 4e-5, "optimizer": "AdamW",
                 "lr_scheduler": "constant", "max_denoising_steps": 10},
       "save": {"name": "nt", "path": "/tmp/unused", "per_steps": 500, "precision": "bfloat16"},
       "logging": {"use_wandb": False, "verbose": False}, "other": {"use_xformers": True}}


def run(clip_index, eager, iters=30, warm=5, engine=False):
    config = config_util.RootConfig(**cfg)
    torch.manual_seed(0)
    t = train_notrigger.NoTriggerTrainer(config, "cuda", clip_index=clip_index, peft_type="lora", rank=4,
                                         encoders=[load_text_encoders("SDXL", seed=1), load_text_encoders("SDXL", seed=1)],
                                         positive=["happy", "smiling"], negative=["sad", "frowning"], engine=engine)
    if eager:
        slider_oracle.attach(t.network)
    for i in range(warm):
        t.iteration(i)
    torch.cuda.synchronize()
    n0 = ops.launch_count()
    t0 = time.perf_counter()
    for i in range(warm, warm + iters):
        t.iteration(i)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return iters / dt, (ops.launch_count() - n0) // iters, len(t.network.unet_loras)


for ci, name in ((0, "CLIP-L 12 layers x 768"), (1, "CLIP-G 32 layers x 1280")):
    ours, launches, mods = run(ci, False)
    eng, _, _ = run(ci, False, iters=100, engine=True)
    eager, _, _ = run(ci, True)
    print(json.dumps({"workload": f"notrigger_te{ci + 1}", "encoder": name, "lora_modules": mods, "dtype": "bf16",
                      "iterations_per_sec_engine": round(eng, 2), "iterations_per_sec_eager_loop": round(ours, 2),
                      "libslider_launches_per_iteration": launches,
                      "torch_eager_reference_arithmetic_it_per_sec": round(eager, 2),
                      "engine_over_reference_arithmetic": round(eng / eager, 2)}))
