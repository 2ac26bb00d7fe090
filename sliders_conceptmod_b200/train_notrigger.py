"""Text-encoder LoRA trainer ("no trigger word" sliders) — host mirror of
/root/reference/conceptmod/notrigger/train_notrigger.py: same entry points (`train`, `train_lora`,
`main`, the CLI flags), same objective, same optimizer / schedule objects, same saved key layout
(`lora_te1_text_model_encoder_layers_N_self_attn_{q,k,v,out}_proj.*`).

What it does: a LoRA (or DoRA) on the attention projections of one CLIP text encoder is trained so that the encoding
of the EMPTY prompt moves towards the encoding of the positive term(s) when the adapter is applied with multiplier +1
and towards the negative term(s) with multiplier -1 — the resulting text-encoder slider needs no trigger word.  The
hot ops are again LoRA-wrapped Linear layers: on CUDA every adapted projection runs through libslider's fused
base + low-rank GEMM and its fused backward (lora.LoRAModule), M = 77 token rows per prompt.

Pretrained encoders cannot be fetched here: `text_encoder.load_text_encoders` builds random-init stand-ins with the
transformers module tree and class names, so module selection and saved keys are the reference's.  SD3 / FLUX.1
(`--model`) are other model families (out of scope).
"""
from __future__ import annotations

import argparse
from pathlib import Path
from typing import Optional

import torch
import torch.nn.functional as F

from . import config_util
from .config_util import RootConfig
from .dora import DoRANetwork
from .lora import LoRANetwork
from .text_encoder import load_text_encoders, text_tokenize

WARMUP_STEPS = 100          # train_notrigger.py:207-217
DECAY_STEPS = 900
SPLIT = 20                  # :291 — the per-step travel is 1/20 of the initial distance


def fixed_distance_loss(trainable_embs: torch.Tensor, target_embs: torch.Tensor, fixed_distance: torch.Tensor):
    """train_notrigger.py:45-64.  Pull `trainable_embs` a FIXED distance along the straight line to `target_embs`
    (never past it): the regression target is trainable + unit(target - trainable) * clamp(d, -|.|, |.|), built from
    the trainable tensor itself, so the gradient flows through both terms as in the reference."""
    gap = target_embs - trainable_embs
    gap_len = torch.norm(gap, dim=-1, keepdim=True)
    unit = gap / (gap_len + 1e-8)
    travel = torch.clamp(fixed_distance.unsqueeze(-1), -gap_len, gap_len)
    goal = trainable_embs + unit * travel
    return ((trainable_embs - goal) ** 2).mean()


def _cos(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """The reference's call, verbatim in effect (:342-343): both operands get a leading axis and cosine_similarity
    reduces over its DEFAULT dim=1 — which has size 1 for [1, 1, 77, D] operands, so the "cosine" is the elementwise
    sign agreement a*b / (|a| |b|).  Kept as is: it is what the reference optimises."""
    return F.cosine_similarity(a.unsqueeze(0), b.unsqueeze(0)).squeeze()


def direction_regulariser(moved: torch.Tensor, own_target: torch.Tensor, other_target: torch.Tensor,
                          neutral: torch.Tensor) -> torch.Tensor:
    """:339-349 (positive side) / :366-376 (negative side): keep the moved encoding's relation to the OTHER side's
    direction what the own target's is, plus the raw agreement, plus a repulsion 1 / mse(moved, other)."""
    v_moved, v_own, v_other = moved - neutral, own_target - neutral, other_target - neutral
    reg = torch.abs((_cos(v_moved, v_other) - _cos(v_own, v_other)).mean())
    reg = reg + _cos(v_moved, v_other).mean()
    return reg + 1 / (((moved - other_target) ** 2).mean() + 1e-8)


def curriculum(pperc, nperc, similarity):
    """:413-432.  pperc / nperc = remaining fraction of each side's initial distance; the side that lags gets the
    larger weight (both weights share the same scale factor, which therefore cancels — as in the reference), and the
    similarity term fades in as the encodings approach their targets."""
    gap = abs(pperc - nperc)
    boost = 1 + 4 * (1 - torch.exp(-gap / 0.05))
    w_p = (pperc * boost) / (pperc * boost + nperc * boost)
    w_n = (nperc * boost) / (pperc * boost + nperc * boost)
    w_r = min(0.95, (nperc + pperc) / 2)
    return w_p * pperc + w_n * nperc, (1.0 - w_r) * similarity, (w_p, w_n, w_r)


class NoTriggerTrainer:
    """State of train_notrigger.train (:67-465) with the loop body as `iteration(i)`."""

    def __init__(self, config: RootConfig, device, clip_index: int = 0, peft_type: str = "lora", rank: int = 4,
                 model: str = "SDXL", positive=None, negative=None, encoders=None, seed: int = 0, engine: bool = False):
        if model not in ("SDXL", "PonyXL"):
            raise NotImplementedError(f"model {model}: SD3 / FLUX.1 are other model families (out of scope)")
        self.config, self.device, self.index = config, torch.device(device), clip_index
        self.dtype = config_util.parse_precision(config.train.precision)
        # The reference loads the checkpoint twice (:96-117): one copy gets the adapter, the other stays static and
        # supplies the target / neutral encodings.  Here the STATIC encodings come from the adapted copy with the
        # adapter switched off (multiplier 0): the reference's static encoder and its adapter-off adapted encoder are
        # the same arithmetic, and its objective relies on it — at iteration 0 (lora_up = 0) the moved encoding must
        # equal the neutral one EXACTLY, because the sign-agreement term `_cos` is differentiated at v = moved - neutral
        # and a rounding-noise v turns that derivative (analytically 0) into O(1) cancellation noise.  A static copy
        # evaluated by cuBLAS next to an adapted copy evaluated by the fused kernel differs in the last bits; the
        # same kernel with and without the rank-r term does not (s * h . 0 adds exact zeros).
        tok_a, enc_a = encoders[0] if encoders is not None else load_text_encoders(model, seed)
        tok_b = encoders[1][0] if encoders is not None else tok_a
        for enc in list(enc_a):
            enc.to(self.device, dtype=self.dtype)
            enc.requires_grad_(False)
            enc.eval()
        self.prefix = ["lora_te1", "lora_te2"][clip_index]
        self.peft_encoder = enc_a[clip_index]
        net_cls = DoRANetwork if peft_type == "dora" else LoRANetwork
        self.network = net_cls(self.peft_encoder, rank=rank, multiplier=1.0, delimiter="_", alpha=config.network.alpha,
                               target_replace=["CLIPAttention", "CLIPSdpaAttention"], prefix=self.prefix,
                               train_method=config.network.training_method).to(self.device, dtype=self.dtype)
        self.network.requires_grad_(True)
        # the reference's optimizer / schedule OBJECTS in the reference's construction order (:207-221): SGD, a linear
        # warm-up LambdaLR stepped for the first 100 iterations (so iteration 0 runs at lr 0), then a default
        # ConstantLR stepped afterwards (which multiplies the rate by 3 at its fifth step)
        self.optimizer = torch.optim.SGD(self.network.parameters(), lr=config.train.lr)
        total = WARMUP_STEPS + DECAY_STEPS

        def warmup(step: int) -> float:        # transformers.get_linear_schedule_with_warmup, restated
            if step < WARMUP_STEPS:
                return float(step) / float(max(1, WARMUP_STEPS))
            return max(0.0, float(total - step) / float(max(1, total - WARMUP_STEPS)))
        self.scheduler = torch.optim.lr_scheduler.LambdaLR(self.optimizer, warmup)
        self.lr_scheduler = torch.optim.lr_scheduler.ConstantLR(self.optimizer)

        def hidden(enc, tokens):
            return enc(tokens.to(self.device), output_hidden_states=True).hidden_states[-1]
        self._hidden = hidden
        positive = ", ".join(positive) if positive is not None else None
        negative = ", ".join(negative) if negative is not None else None
        self.positive = positive if positive else None
        self.negative = negative if negative else None
        with torch.no_grad():
            self._set_multiplier(0.0)           # adapter off: the static encoder of the reference
            self.neu_tokens = text_tokenize(tok_a[clip_index], [""]).to(self.device)
            self.neutral = hidden(self.peft_encoder, self.neu_tokens)
            self.pos = hidden(self.peft_encoder, text_tokenize(tok_b[clip_index], [self.positive])) if self.positive else None
            self.neg = hidden(self.peft_encoder, text_tokenize(tok_b[clip_index], [self.negative])) if self.negative else None
        if self.pos is None and self.neg is None:
            raise ValueError("train_notrigger: give --positive and / or --negative")
        self.lam = config.train.lambda_similarity
        self.distance1 = self.distance2 = None
        self.last = {}
        # engine mode (LoRA on CUDA): flat parameter / fp32 gradient buffers, the whole iteration replayed as one graph
        self.engine = bool(engine) and self.device.type == "cuda" and peft_type != "dora"
        self._graph, self._capturable = None, True
        if self.engine:
            self.network.flatten()
            self._neg_lr = torch.zeros((), device=self.device, dtype=torch.float32)
            self._obj = torch.zeros((), device=self.device, dtype=torch.float32)

    def _set_multiplier(self, m: float) -> None:
        for lora in self.network.unet_loras:
            lora.multiplier = m

    def _objective(self, i: int, on_device: bool = False) -> torch.Tensor:
        """Forward passes and the scalar that is back-propagated (:293-432).  `on_device`: no host read of a tensor
        (the curriculum's `min(0.95, t)` becomes a clamp with the same value and gradient), so the body can be captured
        in a CUDA graph."""
        pos, neg, neutral = self.pos, self.neg, self.neutral
        preg = nreg = None
        with self.network:
            if pos is not None:
                self._set_multiplier(1.0)
                e_p = self._hidden(self.peft_encoder, self.neu_tokens)
                if i == 0:
                    self.distance1 = torch.norm(pos - e_p.detach()).mean() / SPLIT
                ploss = ((pos - e_p) ** 2).mean() if neg is None else fixed_distance_loss(e_p, pos, self.distance1).mean()
                if neg is not None:
                    preg = direction_regulariser(e_p, pos, neg, neutral)
            if neg is not None:
                self._set_multiplier(-1.0)
                e_n = self._hidden(self.peft_encoder, self.neu_tokens)
                if i == 0:
                    self.distance2 = torch.norm(neg - e_n.detach()).mean() / SPLIT
                nloss = ((neg - e_n) ** 2).mean() if pos is None else fixed_distance_loss(e_n, neg, self.distance2).mean()
                if pos is not None:
                    nreg = direction_regulariser(e_n, neg, pos, neutral)
        if pos is None:
            return nloss
        if neg is None:
            return ploss
        # remaining fractions of the initial distances (:393-404), then the curriculum (:413-432)
        nperc = torch.norm(neg - e_n).mean() / (self.distance2 * SPLIT)
        pperc = torch.norm(pos - e_p).mean() / (self.distance1 * SPLIT)
        similarity = self.lam * preg + self.lam * nreg
        if on_device:
            gap = torch.abs(pperc - nperc)
            boost = 1 + 4 * (1 - torch.exp(-gap / 0.05))
            w_p = (pperc * boost) / (pperc * boost + nperc * boost)
            w_n = (nperc * boost) / (pperc * boost + nperc * boost)
            w_r = torch.clamp((nperc + pperc) / 2, max=0.95)
            return w_p * pperc + w_n * nperc + (1.0 - w_r) * similarity
        loss, similarity, w = curriculum(pperc, nperc, similarity)
        self.last = {k: float(torch.as_tensor(v).detach()) for k, v in
                     (("w_p", w[0]), ("w_n", w[1]), ("w_r", w[2]), ("pperc", pperc), ("nperc", nperc))}
        return loss + similarity

    def _step_schedule(self, i: int) -> None:
        if i < WARMUP_STEPS:
            self.scheduler.step()
        else:
            self.lr_scheduler.step()

    def iteration(self, i: int):
        """One pass of the loop body (:293-445).  Returns the value that was back-propagated (a float; with
        `engine=True` the 0-dim device tensor it was copied to, so that nothing waits for the GPU)."""
        if self.engine:
            return self._engine_iteration(i)
        objective = self._objective(i)
        objective.backward()
        torch.nn.utils.clip_grad_value_(self.network.parameters(), clip_value=1.0)       # :437
        self.optimizer.step()
        # no zero_grad in the reference loop: gradients ACCUMULATE across iterations (:293-445 never clears them)
        self._step_schedule(i)
        return float(objective.detach())

    # ---- engine: the iteration as ONE replayed CUDA graph over the flat parameter / gradient buffers ----------------
    def _engine_body(self, i: int) -> None:
        net = self.network
        with net.packed_transposes():
            objective = self._objective(i, on_device=True)
            objective.backward()            # the fused backward launches accumulate straight into net.flat_grad (fp32)
        net.flat_grad.clamp_(-1.0, 1.0)                                                   # clip_grad_value_(1.0)
        net.flat_param.add_((net.flat_grad * self._neg_lr).to(net.flat_param.dtype))      # SGD: p -= lr * g
        self._obj.copy_(objective.detach())

    def _engine_iteration(self, i: int):
        net = self.network
        self._neg_lr.fill_(-float(self.optimizer.param_groups[0]["lr"]))      # the schedule objects stay the lr source
        if i == 0 or self._graph is None and not self._capturable:
            self._engine_body(i)            # iteration 0 fixes the reference distances (data-dependent init): eager
        elif self._graph is None:
            saved = (net.flat_param.clone(), net.flat_grad.clone())
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):          # warm-up outside capture (lazy inits, cudaFuncSetAttribute)
                    self._engine_body(i)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            net.flat_param.copy_(saved[0]); net.flat_grad.copy_(saved[1])
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self._engine_body(i)
            self._graph.replay()
        else:
            self._graph.replay()
        self._step_schedule(i)
        return self._obj

    def state_dict(self, dtype=None):
        return self.network.get_state_dict(dtype)


def train(config: RootConfig, prompts: list, device, on_step_complete=None, save_file: bool = True,
          clip_index: int = 0, peft_type: str = "lora", rank: int = 4, model: str = "SDXL", attributes=None,
          positive=None, negative=None, encoders=None, seed: int = 0, engine: bool = False):
    """train_notrigger.train (:67-486), same arguments (+ `engine`: replay the iteration as one CUDA graph).  `prompts` and `attributes` are accepted and unused, as in the
    reference (its attribute list is reset to [] at :266, so the stabilising term is always zero).  `encoders`
    (two (tokenizers, text_encoders) pairs) replaces the checkpoint load for callers that bring their own models."""
    save_path = Path(config.save.path)
    save_dtype = config_util.parse_precision(config.train.precision)        # the reference saves in the TRAIN precision (:93)
    trainer = NoTriggerTrainer(config, device, clip_index=clip_index, peft_type=peft_type, rank=rank, model=model,
                               positive=positive, negative=negative, encoders=encoders, seed=seed, engine=engine)
    last_recon = None
    for i in range(config.train.iterations):
        trainer.iteration(i)
        if i % 800 == 0 and i > 1000:                                       # early stop on a frozen reconstruction (:387-392)
            with torch.no_grad():
                recon = _reconstruction(trainer)
            if last_recon is not None and last_recon == recon:
                print("loss stopped moving. exitting early.")
                break
            last_recon = recon
        if i % config.save.per_steps == 0 and i != 0 and i != config.train.iterations - 1 and save_file:
            save_path.mkdir(parents=True, exist_ok=True)
            trainer.network.save_weights(save_path / f"{config.save.name}_{clip_index}_{i}steps.safetensors", dtype=save_dtype)
        if on_step_complete is not None:
            on_step_complete(i)
    if save_file:
        save_path.mkdir(parents=True, exist_ok=True)
        trainer.network.save_weights(save_path / f"{config.save.name}_{clip_index}_last.safetensors", dtype=save_dtype)
        return None
    return trainer.state_dict(save_dtype)


def _reconstruction(trainer: NoTriggerTrainer) -> float:
    """The `full_loss` of :380-385: mse of each side's encoding against its target, adapter applied."""
    total = 0.0
    with trainer.network:
        if trainer.pos is not None:
            trainer._set_multiplier(1.0)
            total += float(((trainer.pos - trainer._hidden(trainer.peft_encoder, trainer.neu_tokens)) ** 2).mean())
        if trainer.neg is not None:
            trainer._set_multiplier(-1.0)
            total += float(((trainer.neg - trainer._hidden(trainer.peft_encoder, trainer.neu_tokens)) ** 2).mean())
    return total


def _decorate(config: RootConfig, alpha: float, rank: int, name: Optional[str]) -> None:
    """Name / path suffixes of :494-500 and :535-541."""
    if name is not None:
        config.save.name = name
    config.network.alpha = alpha
    config.network.rank = rank
    config.save.name += f"_alpha{alpha}"
    config.save.name += f"_rank{rank}"
    config.save.name += f"_{config.network.training_method}"
    config.save.path += f"/{config.save.name}"


def train_lora(target, positive, negative, unconditional, alpha=1.0, rank=4, device=0, name=None, attributes=None,
               batch_size=1, config_file="data/config-xl.yaml", resolution=512, steps=None, on_step_complete=None,
               clip_index=0, peft_type="lora", model="SDXL"):
    """Programmatic entry (:506-545): returns the trained state dict instead of writing a file.  The reference also
    rewrites data/prompts-xl.yaml here although the trainer never reads it; that side effect is not reproduced."""
    config = config_util.load_config_from_yaml(config_file)
    if steps is not None:
        config.train.iterations = steps
    _decorate(config, alpha, rank, name)
    dev = torch.device(f"cuda:{device}") if isinstance(device, int) else torch.device(device)
    return train(config, [], dev, on_step_complete, positive=[positive], negative=[negative], save_file=False,
                 clip_index=clip_index, peft_type=peft_type, rank=rank, model=model, attributes=attributes)


def build_parser() -> argparse.ArgumentParser:
    """The reference's flags (:547-631), same names, defaults and required-ness."""
    ap = argparse.ArgumentParser()
    ap.add_argument("--config_file", required=True, help="Config file for training.")
    ap.add_argument("--alpha", type=float, required=True, help="LoRA weight.")
    ap.add_argument("--rank", type=int, required=False, default=4, help="Rank of LoRA.")
    ap.add_argument("--device", type=int, required=False, default=0, help="Device to train on.")
    ap.add_argument("--name", type=str, required=False, default=None)
    ap.add_argument("--attributes", type=str, required=False, default=None,
                    help="attributes to disentangle (comma separated string)")
    ap.add_argument("--peft_type", type=str, required=False, default="dora", help="dora (default) or lora")
    ap.add_argument("--positive", type=str, nargs="+", required=False, default=None, help="positive term(s)")
    ap.add_argument("--model", type=str, required=False, default="SDXL", help="model SDXL or SD3-Medium")
    ap.add_argument("--negative", type=str, required=False, default=None, help="negative term(s)")
    ap.add_argument("--clip_index", type=int, required=True, help="0 based clip index (sdxl has two)")
    ap.add_argument("--engine", action="store_true",
                    help="(not in the reference) replay each iteration as one CUDA graph over flat parameter buffers")
    return ap


def main(args) -> None:
    config = config_util.load_config_from_yaml(args.config_file)
    _decorate(config, args.alpha, args.rank, args.name)
    device = torch.device(f"cuda:{args.device}")
    # the reference passes `--negative` (a plain string) straight to ", ".join (:279): its characters get joined by
    # ", " — reproduced by handing the string over unchanged
    train(config, [], device, on_step_complete=None, positive=args.positive, negative=args.negative,
          clip_index=args.clip_index, peft_type=args.peft_type, rank=args.rank, model=args.model,
          engine=getattr(args, "engine", False))


if __name__ == "__main__":
    main(build_parser().parse_args())
