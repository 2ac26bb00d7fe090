"""CPU tests of the drop-in trainer entry points: CLI flags, config/prompt wiring, data sources, and that the
product path fails loudly (no CPU fallback) when asked to compute without the CUDA library's device.

Reference surfaces: conceptmod/textsliders/train_lora.py:343-417, train_lora_xl.py:397-514,
trainscripts/imagesliders/train_lora-scale.py:376-499, eval-scripts/generate_images_sd1.py (loop arguments)."""
import importlib.util
import sys
from pathlib import Path

import pytest
import torch
import yaml

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

from sliders_conceptmod_b200 import (config_util, generate_images, prompt_util, train_lora, train_lora_scale,  # noqa: E402
                                     train_lora_xl)
from sliders_conceptmod_b200.slider_loop import TextSliderTrainer  # noqa: E402


def _write_yaml(tmp_path, model="tiny-sd", scheduler="ddim", iterations=3, precision="float32", net_type="c3lier",
                negative=False):
    prompts = [{"target": "person", "positive": "old person", "unconditional": "young person", "neutral": "person",
                "action": "enhance", "guidance_scale": 4, "resolution": 128, "batch_size": 1,
                "dynamic_resolution": False}]
    if negative:
        prompts[0]["negative"] = "young person"
    (tmp_path / "prompts.yaml").write_text(yaml.safe_dump(prompts))
    cfg = {"prompts_file": str(tmp_path / "prompts.yaml"),
           "pretrained_model": {"name_or_path": model, "v2": False, "v_pred": False},
           "network": {"type": net_type, "rank": 4, "alpha": 1.0, "training_method": "noxattn"},
           "train": {"precision": precision, "noise_scheduler": scheduler, "iterations": iterations, "lr": 2e-4,
                     "optimizer": "AdamW", "lr_scheduler": "constant", "max_denoising_steps": 10},
           "save": {"name": "temp", "path": str(tmp_path / "models"), "per_steps": 500, "precision": precision},
           "logging": {"use_wandb": False, "verbose": False}, "other": {"use_xformers": True}}
    (tmp_path / "config.yaml").write_text(yaml.safe_dump(cfg))
    return tmp_path / "config.yaml"


def test_train_lora_cli_flags_match_reference():
    p = train_lora.build_parser()
    with pytest.raises(SystemExit):          # --alpha is required (train_lora.py:378-383)
        p.parse_args([])
    a = p.parse_args(["--alpha", "1.0"])
    assert (a.config_file, a.rank, a.device, a.name, a.attributes) == ("data/config.yaml", 4, 0, None, None)
    a = p.parse_args(["--config_file", "c.yaml", "--alpha", "2", "--rank", "8", "--device", "1", "--name", "age",
                      "--attributes", "male, female"])
    assert (a.config_file, a.alpha, a.rank, a.device, a.name, a.attributes) == ("c.yaml", 2.0, 8, 1, "age", "male, female")


def test_train_lora_xl_cli_and_peft_gate(tmp_path):
    p = train_lora_xl.build_parser()
    with pytest.raises(SystemExit):          # --config_file is required here (train_lora_xl.py:462-466)
        p.parse_args(["--alpha", "1.0"])
    a = p.parse_args(["--config_file", "data/config-xl.yaml", "--alpha", "1.0"])
    assert a.config_file == "data/config-xl.yaml" and a.rank == 4 and a.peft_type == "lora"
    a = p.parse_args(["--config_file", "data/config-xl.yaml", "--alpha", "1.0", "--peft_type", "dora"])
    assert a.peft_type == "dora"            # train_lora_xl.py:87-90: DoRANetwork instead of LoRANetwork


def test_image_slider_cli_and_wrapper():
    p = train_lora_scale.build_parser()
    a = p.parse_args(["--alpha", "1.0", "--folder_main", "/data", "--folders", "low, high", "--scales", "-1, 1"])
    assert a.folder_main == "/data" and a.folders == "low, high" and a.scales == "-1, 1"
    # the hyphenated file name of the reference script resolves to the same module
    spec = importlib.util.spec_from_file_location("train_lora_dash_scale",
                                                  ROOT / "sliders_conceptmod_b200" / "train_lora-scale.py")
    src = (ROOT / "sliders_conceptmod_b200" / "train_lora-scale.py").read_text()
    assert spec is not None and "train_lora_scale" in src
    with pytest.raises(TypeError):           # upstream crashes with a NameError-like failure here
        train_lora_scale.train(None, [], "cpu")


def test_main_applies_cli_overrides_to_save_name(tmp_path, monkeypatch):
    cfg_path = _write_yaml(tmp_path)
    seen = {}

    def fake_train(config, prompts, device, engine=False, seed=None):
        seen.update(config=config, prompts=prompts, device=device)

    monkeypatch.setattr(train_lora, "train", fake_train)
    args = train_lora.build_parser().parse_args(["--config_file", str(cfg_path), "--alpha", "2.0", "--rank", "8",
                                                 "--name", "ageslider", "--attributes", "male, female"])
    train_lora.main(args)
    c = seen["config"]
    assert c.network.alpha == 2.0 and c.network.rank == 8
    assert c.save.name == "ageslider_alpha2.0_rank8_noxattn"            # train_lora.py:355-359
    assert c.save.path.endswith("/ageslider_alpha2.0_rank8_noxattn")
    assert [p.target for p in seen["prompts"]] == ["male person", "female person"]
    assert seen["device"] == torch.device("cuda:0")


def test_paired_latents_sources(tmp_path):
    d = train_lora_scale.PairedLatents(scales=[-2, -1, 1, 2], synthetic=5, latent_hw=8, seed=3)
    assert d.unique_scales() == [1.0, 2.0]
    import random
    random.seed(0)
    lo, hi = d.sample(2.0)
    random.seed(0)
    lo2, hi2 = d.sample(2.0)
    assert torch.equal(lo, lo2) and torch.equal(hi, hi2) and lo.shape == (1, 4, 8, 8) and not torch.equal(lo, hi)
    # folder source: same file name in the -s and +s folders
    for name, sign in (("low", -1.0), ("high", 1.0)):
        (tmp_path / name).mkdir()
        for k in range(3):
            torch.save(torch.full((1, 4, 8, 8), sign * (k + 1)), tmp_path / name / f"{k}.pt")
    f = train_lora_scale.PairedLatents(str(tmp_path), ["low", "high"], [-1, 1])
    lo, hi = f.sample(1.0)
    assert torch.equal(lo, -hi)


def test_trainer_builds_on_cpu_and_compute_fails_loudly(tmp_path):
    cfg = config_util.load_config_from_yaml(str(_write_yaml(tmp_path)))
    prompts = prompt_util.load_prompts_from_yaml(cfg.prompts_file, [])
    tr = TextSliderTrainer(cfg, prompts, "cpu", seed=1)
    assert len(tr.prompt_pairs) == 1 and tr.prompt_pairs[0].action == "enhance"
    assert tr.denoise_guidance == 3 and tr.predict_guidance == 1 and tr.max_grad_norm is None
    assert len(tr.network.unet_loras) > 0 and type(tr.optimizer).__name__ == "AdamW"
    with pytest.raises(RuntimeError):        # no CPU fallback: the kernels need CUDA tensors
        tr.iteration(0)


def test_trainer_seed_governs_lora_init_not_the_unet(tmp_path):
    """The random-init UNet stand-in is a fixed function of its own seed (it replaces a pretrained checkpoint) and is
    built under fork_rng: the caller's seed keeps governing the LoRA initialisation and the loop's random draws."""
    cfg = config_util.load_config_from_yaml(str(_write_yaml(tmp_path)))
    prompts = prompt_util.load_prompts_from_yaml(cfg.prompts_file, [])
    trs = [TextSliderTrainer(cfg, prompts, "cpu", seed=s) for s in (1, 2, 1)]
    flat = [torch.cat([p.detach().flatten() for p in t.network.parameters()]) for t in trs]
    unets = [torch.cat([p.detach().flatten() for p in t.unet.parameters()]) for t in trs]
    assert torch.equal(unets[0], unets[1])                  # same substitute "checkpoint" whatever the seed
    assert not torch.equal(flat[0], flat[1])                # different seeds: different lora_down init
    assert torch.equal(flat[0], flat[2])                    # same seed: reproducible
    draws = []
    for s in (1, 2):
        TextSliderTrainer(cfg, prompts, "cpu", seed=s)
        draws.append(torch.randint(1, 50, (4,)).tolist())   # the loop's timesteps_to draws follow the seed too
    assert draws[0] != draws[1]


def test_xl_trainer_keeps_reference_quirks(tmp_path):
    cfg = config_util.load_config_from_yaml(str(_write_yaml(tmp_path, model="tiny-xl", scheduler="euler_a", negative=True)))
    cfg.train.cfg = 2.0
    prompts = prompt_util.load_prompts_from_yaml(cfg.prompts_file, [])
    tr = TextSliderTrainer(cfg, prompts, "cpu", xl=True, rank=4, seed=1)
    # train_lora_xl.py:92-105,349: attention-only LoRA, hard-coded AdamW / cosine / clip regardless of the config
    assert all("attn" in m.lora_name for m in tr.network.unet_loras)
    g = tr.optimizer.param_groups[0]
    assert (g["lr"], g["weight_decay"]) == (1e-4, 1e-6) and tr.max_grad_norm == 0.2
    assert type(tr.lr_scheduler).__name__ == "CosineAnnealingLR" and tr.lr_scheduler.T_max == 50
    assert tr.denoise_guidance == 2.0 and tr.predict_guidance == 2.0
    assert tr.prompt_pairs[0].negative is not None
    assert tr.unet.config.addition_embed


def test_generate_images_parser_defaults():
    a = generate_images.build_parser().parse_args([])
    assert (a.start_noise, a.guidance_scale, a.ddim_steps, a.scales) == (800, 7.5, 50, "-2,-1,0,1,2")


def test_bench_reference_arm_contract_on_cpu():
    """`bench.py --impl reference` (the oracle port on the host cores): one JSON line with the contract's keys, real
    iterations counted in `steps`; under torchrun every rank but 0 exits 0 without work or output."""
    import json
    import os
    import subprocess
    cmd = [sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--workload", "tiny", "--timesteps-to", "4",
           "--steps", "2", "--warmup", "1"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "slider_train_steps_per_sec" and line["unit"] == "steps/s"
    assert line["steps"] == 2 and line["higher_is_better"] is True and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"] == line["e2e"]["value"]
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert line["config"]["workload"] == "tiny" and "model" not in line["config"]
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    r = subprocess.run(cmd + ["--gpus", "2"], capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""
