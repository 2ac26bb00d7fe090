"""Text-encoder LoRA trainer on the GPU: the adapted CLIP attention projections run through libslider's
fused base + low-rank GEMM and its fused backward; compared with the vectors made by running the reference's own
train_notrigger.train() (tests/golden/make_golden_notrigger.py).  fp32 storage is the error-compensated 3xTF32 path
(~2^-21 per product), bf16 is checked against the same vectors by direction."""
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden" / "reference_notrigger.pt"


def _encoders(seed):
    from sliders_conceptmod_b200.text_encoder import load_text_encoders
    return [load_text_encoders("SDXL", seed=seed, tiny=True), load_text_encoders("SDXL", seed=seed, tiny=True)]


def _run(gold, case, precision, engine=False):
    from sliders_conceptmod_b200 import config_util, ops, train_notrigger
    c = gold["cases"][case]
    cfg = dict(gold["config"])
    cfg["train"] = dict(cfg["train"], precision=precision)
    config = config_util.RootConfig(**cfg)
    torch.manual_seed(c["torch_seed"])
    n0 = ops.launch_count()
    sd = train_notrigger.train(config, [], "cuda", save_file=False, clip_index=c["clip_index"], peft_type="lora", rank=4,
                               model="SDXL", encoders=_encoders(gold["seed_model"]), engine=engine, **c["kw"])
    return sd, c["final"], ops.launch_count() - n0


@pytest.mark.parametrize("case", ["both_te1", "both_te2", "positive_only_te1", "negative_only_te2"])
def test_fp32_final_state_matches_reference(case):
    gold = torch.load(GOLD)
    sd, ref, launches = _run(gold, case, "float32")
    assert list(sd.keys()) == list(ref.keys())
    # 12 iterations x (1 or 2 encoder passes) x 8 adapted projections, forward + backward, all through libslider
    assert launches >= 12 * 8 * 2, f"only {launches} libslider launches: the adapted layers did not take the CUDA path"
    for k in ref:
        scale = ref[k].abs().max().item() + 1e-12
        err = (sd[k].float().cpu() - ref[k]).abs().max().item()
        assert err <= 2e-3 * scale, f"{k}: max err {err} vs scale {scale}"


@pytest.mark.parametrize("case", ["both_te1", "positive_only_te2"])
def test_bf16_moves_the_adapter_the_reference_way(case):
    gold = torch.load(GOLD)
    sd, ref, launches = _run(gold, case, "bfloat16")
    assert launches >= 12 * 8 * 2
    got = torch.cat([sd[k].float().cpu().flatten() for k in ref if "lora_up" in k])
    want = torch.cat([ref[k].flatten() for k in ref if "lora_up" in k])
    cos = torch.dot(got, want) / (got.norm() * want.norm())
    assert cos > 0.97, f"lora_up direction cosine {cos}"        # lora_up starts at zero: its final value IS the update
    assert 0.8 < got.norm() / want.norm() < 1.25


def test_dora_variant_trains(tmp_path):
    """--peft_type dora is the CLI default (train_notrigger.py:589): the DoRA network takes the same path."""
    from sliders_conceptmod_b200 import config_util, train_notrigger
    gold = torch.load(GOLD)
    cfg = dict(gold["config"])
    cfg["train"] = dict(cfg["train"], precision="bfloat16", iterations=4)
    config = config_util.RootConfig(**cfg)
    torch.manual_seed(3)
    sd = train_notrigger.train(config, [], "cuda", save_file=False, clip_index=0, peft_type="dora", rank=4, model="SDXL",
                               encoders=_encoders(gold["seed_model"]), positive=["happy"], negative=["sad"])
    assert any("dora_scale" in k for k in sd) and all(torch.isfinite(v.float()).all() for v in sd.values())
    up = torch.cat([v.float().flatten() for k, v in sd.items() if "lora_up" in k])
    assert up.norm() > 0


@pytest.mark.parametrize("case", ["both_te1", "both_te2", "negative_only_te1"])
def test_engine_graph_replay_matches_reference_and_eager_loop(case):
    """engine=True: flat parameter / fp32 gradient buffers, iterations 1.. replayed as ONE CUDA graph (forward passes,
    objective, backward, clip, SGD) — same final state as the reference run and as the eager loop."""
    gold = torch.load(GOLD)
    sd_e, ref, _ = _run(gold, case, "float32", engine=True)
    sd_p, _, _ = _run(gold, case, "float32", engine=False)
    assert list(sd_e.keys()) == list(ref.keys())
    for k in ref:
        scale = ref[k].abs().max().item() + 1e-12
        assert (sd_e[k].float().cpu() - ref[k]).abs().max().item() <= 2e-3 * scale, k
        assert (sd_e[k].float() - sd_p[k].float()).abs().max().item() <= 1e-3 * scale, k
