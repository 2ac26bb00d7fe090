"""One implementation of the text-slider training iteration, shared by the SD-1.x and SD-XL CLIs.

Follows the reference loops statement by statement where results depend on it
(conceptmod/textsliders/train_lora.py:155-321, train_lora_xl.py:165-375): prompt-pair draw, `timesteps_to`
draw, partial denoise with the adapter on, three adapter-off predictions, grad prediction, guidance loss,
backward, optimizer + LR-scheduler step, periodic save.  What differs is how it is issued: every sampler /
objective call below lands in libslider_b200 kernels, there is no per-iteration `flush()` and the loss is
only read back to the host when it is displayed.
"""
from __future__ import annotations

import ast
from pathlib import Path
from typing import Callable, Optional

import torch

from . import config_util, model_util, train_util
from .lora import DEFAULT_TARGET_REPLACE, UNET_TARGET_REPLACE_MODULE_CONV, LoRANetwork
from .prompt_util import PromptEmbedsCache, PromptEmbedsPair, PromptEmbedsXL, PromptSettings


def _encode(xl: bool, tokenizers, text_encoders, prompt: str):
    if not xl:
        return model_util.encode_prompts(tokenizers, text_encoders, [prompt])
    enc = text_encoders[0]
    return PromptEmbedsXL(enc.encode([prompt]), enc.encode_pooled([prompt]))


class TextSliderTrainer:
    """State of one training run; `iteration(i)` is one pass of the reference loop body."""

    def __init__(self, config: config_util.RootConfig, prompts: list[PromptSettings], device, xl: bool = False,
                 rank: Optional[int] = None, seed: Optional[int] = None, peft_type: str = "lora"):
        self.config, self.xl, self.device = config, xl, torch.device(device)
        self.weight_dtype = config_util.parse_precision(config.train.precision)
        self.save_weight_dtype = config_util.parse_precision(config.train.precision)   # save.precision is unused upstream
        self.metadata = {"prompts": ",".join(p.json() for p in prompts), "config": config.json()}
        if seed is not None:
            torch.manual_seed(seed)
        name = config.pretrained_model.name_or_path
        if xl:
            tokenizers, text_encoders, unet, sched = model_util.load_models_xl(name, scheduler_name=config.train.noise_scheduler)
            for te in text_encoders:
                te.to(self.device)
        else:
            tokenizers, text_encoders, unet, sched = model_util.load_models(
                name, scheduler_name=config.train.noise_scheduler, v2=config.pretrained_model.v2,
                v_pred=config.pretrained_model.v_pred)
            text_encoders.to(self.device)
        unet.to(self.device, dtype=self.weight_dtype)
        if config.other.use_xformers:
            unet.enable_xformers_memory_efficient_attention()
        unet.requires_grad_(False)
        unet.eval()
        self.unet, self.scheduler = unet, sched
        if xl:
            # train_lora_xl.py:92-105: attention-only targets, alpha default, hard-coded optimizer
            if peft_type == "dora":                      # train_lora_xl.py:87-90
                from .dora import DoRANetwork as peft_class
            else:
                peft_class = LoRANetwork
            self.network = peft_class(unet, rank=rank if rank is not None else 4, multiplier=1.0, delimiter="_",
                                      target_replace=["Attention"], prefix="lora_unet",
                                      train_method=config.network.training_method).to(self.device, dtype=self.weight_dtype)
            self.optimizer = torch.optim.AdamW(self.network.parameters(), lr=1e-4, weight_decay=1e-6)
            self.lr_scheduler = torch.optim.lr_scheduler.CosineAnnealingLR(self.optimizer, T_max=50, eta_min=1e-6)
            self.max_grad_norm = 0.2
            self.denoise_guidance = config.train.cfg
            self.predict_guidance = config.train.cfg
        else:
            modules = list(DEFAULT_TARGET_REPLACE)
            if config.network.type == "c3lier":          # train_lora.py:44-46
                modules += UNET_TARGET_REPLACE_MODULE_CONV
            self.network = LoRANetwork(unet, rank=config.network.rank, multiplier=1.0, alpha=config.network.alpha,
                                       target_replace=modules,
                                       train_method=config.network.training_method).to(self.device, dtype=self.weight_dtype)
            opt_cls = train_util.get_optimizer(config.train.optimizer)
            kwargs = {}
            if config.train.optimizer_args:              # "k=v k=v" (train_lora.py:82-87)
                for arg in config.train.optimizer_args.split(" "):
                    key, value = arg.split("=")
                    kwargs[key] = ast.literal_eval(value)
            self.optimizer = opt_cls(self.network.prepare_optimizer_params(), lr=config.train.lr, **kwargs)
            self.lr_scheduler = train_util.get_lr_scheduler(config.train.lr_scheduler, self.optimizer,
                                                            max_iterations=config.train.iterations,
                                                            lr_min=config.train.lr / 100)
            self.max_grad_norm = None
            self.denoise_guidance = 3          # train_lora.py:206
            self.predict_guidance = 1          # train_lora.py:226-272
        self.criteria = torch.nn.MSELoss()
        cache = PromptEmbedsCache()
        self.prompt_pairs: list[PromptEmbedsPair] = []
        with torch.no_grad():
            for settings in prompts:
                names = [settings.target, settings.positive, settings.neutral, settings.unconditional]
                if xl:
                    names.append(settings.negative if settings.negative is not None else settings.unconditional)
                for prompt in names:
                    if cache[prompt] is None:
                        cache[prompt] = _encode(xl, tokenizers, text_encoders, prompt)
                args = [self.criteria, cache[settings.target], cache[settings.positive], cache[settings.unconditional],
                        cache[settings.neutral]]
                if xl:
                    args.append(cache[names[-1]])
                self.prompt_pairs.append(PromptEmbedsPair(*args, settings))
        del tokenizers, text_encoders
        self.last_loss = None

    # ---- helpers --------------------------------------------------------------------------------
    def _cat(self, u, c, n):
        return train_util.concat_embeddings(u, c, n).to(self.device, dtype=self.weight_dtype)

    def _predict(self, pair, t, latents, cond, add_time_ids, guidance):
        b = pair.batch_size
        if not self.xl:
            return train_util.predict_noise(self.unet, self.scheduler, t, latents,
                                            self._cat(pair.unconditional, cond, b), guidance_scale=guidance)
        return train_util.predict_noise_xl(
            self.unet, self.scheduler, t, latents,
            text_embeddings=self._cat(pair.unconditional.text_embeds, cond.text_embeds, b),
            add_text_embeddings=self._cat(pair.unconditional.pooled_embeds, cond.pooled_embeds, b),
            add_time_ids=self._cat(add_time_ids, add_time_ids, b), guidance_scale=guidance)

    # ---- one reference iteration ------------------------------------------------------------------
    def iteration(self, i: int) -> torch.Tensor:
        cfg, sched, net = self.config, self.scheduler, self.network
        n_den = cfg.train.max_denoising_steps
        with torch.no_grad():
            sched.set_timesteps(n_den, device=self.device)
            self.optimizer.zero_grad()
            pair = self.prompt_pairs[torch.randint(0, len(self.prompt_pairs), (1,)).item()]
            timesteps_to = torch.randint(1, n_den, (1,)).item()
            height, width = pair.resolution, pair.resolution
            if pair.dynamic_resolution:
                height, width = train_util.get_random_resolution_in_bucket(pair.resolution)
            latents = train_util.get_initial_latents(sched, pair.batch_size, height, width, 1).to(
                self.device, dtype=self.weight_dtype)
            add_time_ids = None
            if self.xl:
                add_time_ids = train_util.get_add_time_ids(height, width, dynamic_crops=pair.dynamic_crops,
                                                           dtype=self.weight_dtype).to(self.device)
            with net:       # adapter on: partially denoised latents
                if not self.xl:
                    denoised = train_util.diffusion(
                        self.unet, sched, latents, self._cat(pair.unconditional, pair.target, pair.batch_size),
                        start_timesteps=0, total_timesteps=timesteps_to, guidance_scale=self.denoise_guidance)
                else:
                    b = pair.batch_size
                    denoised = train_util.diffusion_xl(
                        self.unet, sched, latents,
                        text_embeddings=self._cat(pair.unconditional.text_embeds, pair.target.text_embeds, b),
                        add_text_embeddings=self._cat(pair.unconditional.pooled_embeds, pair.target.pooled_embeds, b),
                        add_time_ids=self._cat(add_time_ids, add_time_ids, b),
                        start_timesteps=0, total_timesteps=timesteps_to, guidance_scale=self.denoise_guidance)
            sched.set_timesteps(1000)
            t_cur = sched.timesteps[int(timesteps_to * 1000 / n_den)]
            g = self.predict_guidance      # adapter off outside `with network:` (multiplier 0)
            positive = self._predict(pair, t_cur, denoised, pair.positive, add_time_ids, g)
            neutral = self._predict(pair, t_cur, denoised, pair.neutral, add_time_ids, g)
            q_emb = pair.negative if self.xl else pair.unconditional
            q_lat = self._predict(pair, t_cur, denoised, q_emb, add_time_ids, g)
        with net:
            target = self._predict(pair, t_cur, denoised, pair.target, add_time_ids, g)
        kw = {"negative_latents": q_lat} if self.xl else {"unconditional_latents": q_lat}
        loss = pair.loss(target_latents=target, positive_latents=positive, neutral_latents=neutral, **kw)
        loss.backward()
        if self.max_grad_norm is not None:
            torch.nn.utils.clip_grad_norm_(self.network.parameters(), max_norm=self.max_grad_norm)
        self.optimizer.step()
        self.lr_scheduler.step()
        self.last_loss = loss.detach()
        return self.last_loss

    # ---- the loop -----------------------------------------------------------------------------------
    def run(self, on_step_complete: Optional[Callable[[int], None]] = None, save_file: bool = True,
            log_every: int = 10, suffix: str = ".pt"):
        cfg = self.config
        save_path = Path(cfg.save.path)
        try:
            from tqdm import tqdm
            pbar = tqdm(range(cfg.train.iterations))
        except Exception:       # pragma: no cover
            pbar = range(cfg.train.iterations)
        for i in pbar:
            loss = self.iteration(i)
            if hasattr(pbar, "set_description") and (i % log_every == 0 or i == cfg.train.iterations - 1):
                pbar.set_description(f"Loss*1k: {loss.item() * 1000:.4f}")     # the only host sync, every log_every
            if i % cfg.save.per_steps == 0 and i != 0 and i != cfg.train.iterations - 1 and save_file:
                save_path.mkdir(parents=True, exist_ok=True)
                self.network.save_weights(save_path / f"{cfg.save.name}_{i}steps{suffix}", dtype=self.save_weight_dtype)
            if on_step_complete is not None:
                on_step_complete(i)
        if save_file:
            save_path.mkdir(parents=True, exist_ok=True)
            self.network.save_weights(save_path / f"{cfg.save.name}_last{suffix}", dtype=self.save_weight_dtype)
            return None
        return self.network.get_state_dict(self.save_weight_dtype)
