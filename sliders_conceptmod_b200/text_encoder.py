"""CLIP text-encoder stand-in for the text-encoder LoRA trainer (reference call sites
conceptmod/notrigger/train_notrigger.py:178-215,240-262 and conceptmod/textsliders/train_util.py:108-122).

Pretrained CLIP weights cannot be fetched here, so this is the published CLIP text-transformer architecture with
random-init weights: the module tree and CLASS NAMES follow `transformers`' CLIPTextModel (text_model.embeddings,
text_model.encoder.layers.N.{layer_norm1, self_attn.{q,k,v,out}_proj, layer_norm2, mlp.{fc1,fc2}},
text_model.final_layer_norm), because the reference selects the layers to adapt by class name ("CLIPAttention",
train_notrigger.py:187) and derives the saved keys from the module path (`lora_te1_text_model_encoder_layers_0_...`).
`forward(input_ids, output_hidden_states=True).hidden_states[-1]` is the last encoder layer's output BEFORE the final
LayerNorm, as in transformers.

The adapted projections run through libslider's fused base + low-rank GEMM (LoRAModule patches their forward); on CUDA
the frozen LayerNorms use libslider's LayerNorm kernels (forward and backward).  The frozen MLP projections (cuBLAS)
and the 77-token causal attention core (library SDPA) are library calls, as in the UNet stand-in.
"""
from __future__ import annotations

import dataclasses
import re
import zlib
from types import SimpleNamespace

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclasses.dataclass
class CLIPTextConfig:
    vocab_size: int = 49408
    hidden_size: int = 768
    intermediate_size: int = 3072
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    max_position_embeddings: int = 77
    hidden_act: str = "quick_gelu"
    layer_norm_eps: float = 1e-5

    @staticmethod
    def clip_l() -> "CLIPTextConfig":            # SD 1.x text encoder / SD-XL text_encoder (clip_index 0)
        return CLIPTextConfig()

    @staticmethod
    def clip_g() -> "CLIPTextConfig":            # SD-XL text_encoder_2 (clip_index 1)
        return CLIPTextConfig(hidden_size=1280, intermediate_size=5120, num_hidden_layers=32, num_attention_heads=20,
                              hidden_act="gelu")

    @staticmethod
    def tiny(hidden: int = 64) -> "CLIPTextConfig":
        return CLIPTextConfig(vocab_size=1024, hidden_size=hidden, intermediate_size=4 * hidden, num_hidden_layers=2,
                              num_attention_heads=4)


class CLIPTextEmbeddings(nn.Module):
    def __init__(self, cfg: CLIPTextConfig):
        super().__init__()
        self.token_embedding = nn.Embedding(cfg.vocab_size, cfg.hidden_size)
        self.position_embedding = nn.Embedding(cfg.max_position_embeddings, cfg.hidden_size)
        self.register_buffer("position_ids", torch.arange(cfg.max_position_embeddings).unsqueeze(0), persistent=False)

    def forward(self, input_ids):
        return self.token_embedding(input_ids) + self.position_embedding(self.position_ids[:, : input_ids.shape[1]])


class CLIPAttention(nn.Module):
    """Causal multi-head self-attention; the four projections are what the trainer adapts."""

    def __init__(self, cfg: CLIPTextConfig):
        super().__init__()
        self.num_heads = cfg.num_attention_heads
        self.head_dim = cfg.hidden_size // cfg.num_attention_heads
        self.k_proj = nn.Linear(cfg.hidden_size, cfg.hidden_size)
        self.v_proj = nn.Linear(cfg.hidden_size, cfg.hidden_size)
        self.q_proj = nn.Linear(cfg.hidden_size, cfg.hidden_size)
        self.out_proj = nn.Linear(cfg.hidden_size, cfg.hidden_size)

    def forward(self, x):
        B, L, _ = x.shape
        q = self.q_proj(x).view(B, L, self.num_heads, self.head_dim).transpose(1, 2)
        k = self.k_proj(x).view(B, L, self.num_heads, self.head_dim).transpose(1, 2)
        v = self.v_proj(x).view(B, L, self.num_heads, self.head_dim).transpose(1, 2)
        o = F.scaled_dot_product_attention(q, k, v, is_causal=True)      # scale = head_dim ** -0.5, as CLIP
        return self.out_proj(o.transpose(1, 2).reshape(B, L, -1))


class CLIPMLP(nn.Module):
    def __init__(self, cfg: CLIPTextConfig):
        super().__init__()
        self.fc1 = nn.Linear(cfg.hidden_size, cfg.intermediate_size)
        self.fc2 = nn.Linear(cfg.intermediate_size, cfg.hidden_size)
        self.quick = cfg.hidden_act == "quick_gelu"

    def forward(self, x):
        h = self.fc1(x)
        h = h * torch.sigmoid(1.702 * h) if self.quick else F.gelu(h)
        return self.fc2(h)


def _layer_norm(norm: nn.LayerNorm, x: torch.Tensor) -> torch.Tensor:
    if x.is_cuda:
        from .unet import layer_norm_fused          # libslider forward (+ backward behind an autograd function)
        return layer_norm_fused(norm, x)
    return norm(x)


class CLIPEncoderLayer(nn.Module):
    def __init__(self, cfg: CLIPTextConfig):
        super().__init__()
        self.self_attn = CLIPAttention(cfg)
        self.layer_norm1 = nn.LayerNorm(cfg.hidden_size, eps=cfg.layer_norm_eps)
        self.mlp = CLIPMLP(cfg)
        self.layer_norm2 = nn.LayerNorm(cfg.hidden_size, eps=cfg.layer_norm_eps)

    def forward(self, x):
        x = x + self.self_attn(_layer_norm(self.layer_norm1, x))
        return x + self.mlp(_layer_norm(self.layer_norm2, x))


class CLIPEncoder(nn.Module):
    def __init__(self, cfg: CLIPTextConfig):
        super().__init__()
        self.layers = nn.ModuleList([CLIPEncoderLayer(cfg) for _ in range(cfg.num_hidden_layers)])


class CLIPTextTransformer(nn.Module):
    def __init__(self, cfg: CLIPTextConfig):
        super().__init__()
        self.embeddings = CLIPTextEmbeddings(cfg)
        self.encoder = CLIPEncoder(cfg)
        self.final_layer_norm = nn.LayerNorm(cfg.hidden_size, eps=cfg.layer_norm_eps)


class CLIPTextModel(nn.Module):
    def __init__(self, cfg: CLIPTextConfig):
        super().__init__()
        self.config = cfg
        self.text_model = CLIPTextTransformer(cfg)

    @property
    def device(self):
        return self.text_model.final_layer_norm.weight.device

    @property
    def dtype(self):
        return self.text_model.final_layer_norm.weight.dtype

    def _encode(self, input_ids, output_hidden_states):
        tm = self.text_model
        x = tm.embeddings(input_ids)
        hidden = [x]
        for layer in tm.encoder.layers:
            x = layer(x)
            hidden.append(x)
        last = _layer_norm(tm.final_layer_norm, x)
        pooled = last[torch.arange(last.shape[0], device=last.device), input_ids.argmax(dim=-1)]   # EOS = highest id
        return last, pooled, (tuple(hidden) if output_hidden_states else None)

    def forward(self, input_ids, output_hidden_states: bool = False, **_):
        last, pooled, hidden = self._encode(input_ids, output_hidden_states)
        return TextModelOutput(("last_hidden_state", "pooler_output"), last_hidden_state=last, pooler_output=pooled,
                               hidden_states=hidden)


class TextModelOutput(SimpleNamespace):
    """Attribute access plus the positional indexing of transformers' ModelOutput (`out[0]`)."""

    def __init__(self, order, **kw):
        super().__init__(**kw)
        self._order = order

    def __getitem__(self, i):
        return getattr(self, self._order[i])


class CLIPTextModelWithProjection(CLIPTextModel):
    """SD-XL's second encoder: `out[0]` = text_embeds = text_projection(pooled) (train_util.py:147-150 reads it)."""

    def __init__(self, cfg: CLIPTextConfig, projection_dim: int | None = None):
        super().__init__(cfg)
        self.text_projection = nn.Linear(cfg.hidden_size, projection_dim or cfg.hidden_size, bias=False)

    def forward(self, input_ids, output_hidden_states: bool = False, **_):
        last, pooled, hidden = self._encode(input_ids, output_hidden_states)
        return TextModelOutput(("text_embeds", "last_hidden_state"), text_embeds=self.text_projection(pooled),
                               last_hidden_state=last, hidden_states=hidden)


class HashTokenizer:
    """Deterministic stand-in for CLIPTokenizer: lower-cased words and punctuation marks map to `crc32 % (vocab - 2)`,
    BOS = vocab - 2, EOS / padding = vocab - 1 (CLIP pads with EOS, so EOS is the highest id of a row)."""
    model_max_length = 77

    def __init__(self, vocab_size: int = 49408):
        self.vocab_size = vocab_size

    def __call__(self, prompts, padding="max_length", max_length=None, truncation=True, return_tensors="pt", **_):
        if isinstance(prompts, str):
            prompts = [prompts]
        L = max_length or self.model_max_length
        bos, eos = self.vocab_size - 2, self.vocab_size - 1
        rows = []
        for p in prompts:
            words = re.findall(r"[a-z0-9]+|[^\sa-z0-9]", p.lower())
            ids = [bos] + [zlib.crc32(w.encode("utf-8")) % (self.vocab_size - 2) for w in words][: L - 2] + [eos]
            rows.append(ids + [eos] * (L - len(ids)))
        return SimpleNamespace(input_ids=torch.tensor(rows, dtype=torch.long))


def text_tokenize(tokenizer, prompts: list[str]) -> torch.Tensor:
    """conceptmod/textsliders/train_util.py:108-116."""
    return tokenizer(prompts, padding="max_length", max_length=tokenizer.model_max_length, truncation=True,
                     return_tensors="pt").input_ids


def load_text_encoders(model: str = "SDXL", seed: int = 0, tiny: bool = False):
    """(tokenizers, text_encoders) of the family — stand-ins for what model_util.load_models_xl returns
    (train_notrigger.py:96-117).  A fixed function of `seed`; the caller's RNG stream is untouched."""
    cfgs = [CLIPTextConfig.tiny(64), CLIPTextConfig.tiny(128)] if tiny else [CLIPTextConfig.clip_l(), CLIPTextConfig.clip_g()]
    if model not in ("SDXL", "PonyXL"):
        cfgs = cfgs[:1]
    with torch.random.fork_rng(devices=[]):
        torch.manual_seed(seed)
        # SD-XL: text_encoder is a CLIPTextModel, text_encoder_2 a CLIPTextModelWithProjection
        encoders = [CLIPTextModel(c) if i == 0 else CLIPTextModelWithProjection(c) for i, c in enumerate(cfgs)]
    return [HashTokenizer(c.vocab_size) for c in cfgs], encoders


def text_encode(text_encoder, tokens):
    """conceptmod/textsliders/train_util.py:119-120."""
    return text_encoder(tokens.to(text_encoder.device))[0]


def encode_prompts(tokenizer, text_encoder, prompts: list[str]):
    """train_util.py:123-134."""
    return text_encode(text_encoder, text_tokenize(tokenizer, prompts))


def text_encode_xl(text_encoder, tokens, num_images_per_prompt: int = 1):
    """train_util.py:138-154: penultimate hidden state + `out[0]` (the projected pooled embedding for text_encoder_2)."""
    out = text_encoder(tokens.to(text_encoder.device), output_hidden_states=True)
    pooled = out[0]
    embeds = out.hidden_states[-2]
    bs, seq, _ = embeds.shape
    embeds = embeds.repeat(1, num_images_per_prompt, 1).view(bs * num_images_per_prompt, seq, -1)
    return embeds, pooled


def encode_prompts_xl(tokenizers, text_encoders, prompts: list[str], num_images_per_prompt: int = 1):
    """train_util.py:224-247: both encoders' penultimate states concatenated on the feature axis, text_encoder_2's pool."""
    parts, pooled = [], None
    for tokenizer, text_encoder in zip(tokenizers, text_encoders):
        embeds, pooled = text_encode_xl(text_encoder, text_tokenize(tokenizer, prompts), num_images_per_prompt)
        parts.append(embeds)
    bs = pooled.shape[0]
    pooled = pooled.repeat(1, num_images_per_prompt).view(bs * num_images_per_prompt, -1)
    return torch.concat(parts, dim=-1), pooled
