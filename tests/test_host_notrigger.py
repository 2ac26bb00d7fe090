"""Text-encoder LoRA trainer on the CPU against vectors made by RUNNING the reference's own
train_notrigger.train() (tests/golden/make_golden_notrigger.py): module selection + key names, the objective, the
curriculum weights, gradient clipping, the SGD / warm-up / ConstantLR objects — everything up to the final LoRA state.  On the CPU the adapted layers
evaluate through the oracle's LoRA arithmetic (oracle.attach), as in the other host-logic replays; the GPU form of the
same comparison is tests/test_gpu_notrigger.py."""
from pathlib import Path

import pytest
import torch

GOLD = Path(__file__).resolve().parent / "golden" / "reference_notrigger.pt"


@pytest.fixture(scope="module")
def gold():
    return torch.load(GOLD)


def _encoders(seed):
    from sliders_conceptmod_b200.text_encoder import load_text_encoders
    return [load_text_encoders("SDXL", seed=seed, tiny=True), load_text_encoders("SDXL", seed=seed, tiny=True)]


def test_fixed_distance_loss_matches_reference(gold):
    from sliders_conceptmod_b200.train_notrigger import fixed_distance_loss
    fd = gold["fixed_distance"]
    a = fd["a"].clone().requires_grad_(True)
    val = fixed_distance_loss(a, fd["b"], fd["d"])
    grad, = torch.autograd.grad(val, a)
    assert torch.equal(val.detach(), fd["value"])
    assert torch.equal(grad, fd["grad"])


@pytest.mark.parametrize("case", ["both_te1", "both_te2", "positive_only_te1", "positive_only_te2", "negative_only_te1",
                                  "negative_only_te2"])
def test_trainer_reproduces_reference_final_state(gold, case, monkeypatch):
    from sliders_conceptmod_b200 import config_util, lora, train_notrigger
    from oracle import slider_oracle

    class OracleLoRANetwork(lora.LoRANetwork):       # CPU: the product's layers refuse to run; the oracle arithmetic stands in
        def __init__(self, *a, **k):
            super().__init__(*a, **k)
            slider_oracle.attach(self)
    monkeypatch.setattr(train_notrigger, "LoRANetwork", OracleLoRANetwork)
    c = gold["cases"][case]
    config = config_util.RootConfig(**gold["config"])
    torch.manual_seed(c["torch_seed"])
    sd = train_notrigger.train(config, [], "cpu", save_file=False, clip_index=c["clip_index"], peft_type="lora", rank=4,
                               model="SDXL", encoders=_encoders(gold["seed_model"]), **c["kw"])
    ref = c["final"]
    assert list(sd.keys()) == list(ref.keys())                       # same modules, same names, same order
    assert all(k.startswith(f"lora_te{c['clip_index'] + 1}_text_model_encoder_layers_") for k in sd)
    for k in ref:
        assert sd[k].shape == ref[k].shape and sd[k].dtype == ref[k].dtype, k
        torch.testing.assert_close(sd[k], ref[k], rtol=2e-5, atol=2e-6, msg=lambda m, k=k: f"{k}: {m}")
    up = torch.cat([v.flatten() for k, v in sd.items() if "lora_up" in k])
    assert up.norm() > 0.5                                            # the run really moved the adapter


def test_cli_flags_match_reference():
    from sliders_conceptmod_b200.train_notrigger import build_parser
    ap = build_parser()
    a = ap.parse_args(["--config_file", "c.yaml", "--alpha", "1.0", "--clip_index", "1", "--positive", "happy", "smiling",
                       "--negative", "sad"])
    assert a.peft_type == "dora" and a.rank == 4 and a.model == "SDXL" and a.positive == ["happy", "smiling"]
    assert a.negative == "sad" and a.clip_index == 1 and a.device == 0
    with pytest.raises(SystemExit):
        ap.parse_args(["--config_file", "c.yaml", "--alpha", "1.0"])          # --clip_index is required


def test_other_model_families_are_rejected():
    from sliders_conceptmod_b200 import config_util, train_notrigger
    gold = torch.load(GOLD)
    with pytest.raises(NotImplementedError):
        train_notrigger.train(config_util.RootConfig(**gold["config"]), [], "cpu", model="FLUX.1", positive=["x"])


def test_prompt_encoding_matches_reference_functions(gold):
    """text_tokenize / encode_prompts / encode_prompts_xl (train_util.py:108-154, 224-247) on the CLIP stand-ins."""
    from sliders_conceptmod_b200 import text_encoder as te
    enc = gold["encode"]
    toks, encs = te.load_text_encoders("SDXL", seed=gold["seed_model"], tiny=True)
    for e in encs:
        e.eval()
    assert torch.equal(te.text_tokenize(toks[0], enc["prompts"]), enc["tokens"])
    with torch.no_grad():
        xl_embeds, xl_pooled = te.encode_prompts_xl(toks, encs, enc["prompts"], num_images_per_prompt=2)
        sd_embeds = te.encode_prompts(toks[0], encs[0], enc["prompts"])
    assert xl_embeds.shape == (4, 77, 64 + 128) and xl_pooled.shape == (4, 128)
    assert torch.equal(xl_embeds, enc["xl_embeds"]) and torch.equal(xl_pooled, enc["xl_pooled"])
    assert torch.equal(sd_embeds, enc["sd_embeds"])


def test_clip_stand_ins_have_the_published_sizes_and_key_layout():
    """The random-init stand-ins are the architectures the reference loads (model_util.py:117-186: CLIP ViT-L/14 text
    model and OpenCLIP bigG/14 text model with projection): same parameter counts as the published checkpoints and the
    transformers key layout that `LoRANetwork(target_replace=[CLIPAttention, CLIPMLP])` walks (train_notrigger.py:134-148)."""
    from sliders_conceptmod_b200 import text_encoder as te
    with torch.device("meta"):
        enc_l = te.CLIPTextModel(te.CLIPTextConfig.clip_l())
        enc_g = te.CLIPTextModelWithProjection(te.CLIPTextConfig.clip_g())
    assert sum(p.numel() for p in enc_l.parameters()) == 123_060_480
    assert sum(p.numel() for p in enc_g.parameters()) == 694_659_840
    keys = list(enc_g.state_dict())
    assert keys[0] == "text_model.embeddings.token_embedding.weight" and keys[-1] == "text_projection.weight"
    assert "text_model.encoder.layers.31.self_attn.out_proj.weight" in keys and "text_model.encoder.layers.31.mlp.fc2.bias" in keys
    per_layer = {k.split("layers.0.")[1] for k in enc_l.state_dict() if ".layers.0." in k}
    assert per_layer == {f"self_attn.{p}_proj.{w}" for p in "qkvo".replace("o", "") for w in ("weight", "bias")} | {
        "self_attn.out_proj.weight", "self_attn.out_proj.bias", "layer_norm1.weight", "layer_norm1.bias",
        "layer_norm2.weight", "layer_norm2.bias", "mlp.fc1.weight", "mlp.fc1.bias", "mlp.fc2.weight", "mlp.fc2.bias"}
