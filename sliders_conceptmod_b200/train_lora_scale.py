"""Image-pair slider trainer — drop-in for trainscripts/imagesliders/train_lora-scale.py
(`prev_step` :31-39, `train(config, prompts, device, folder_main, folders, scales)` :41-373, CLI :421-499);
importable module name of the hyphenated script `train_lora-scale.py` next to it.

Objective (reference :284-334): at a random scale s from |scales|, two grad predictions of the SAME noisy
pair — adapter at +s on the "high" latents with the positive prompt, adapter at -s on the "low" latents with the
neutral prompt — each regressed on the noise that was added (plain epsilon MSE, no guidance term), gradients of
both accumulated before ONE optimizer step.  The reference computes both losses on the host CPU in fp32 after a
D2H copy of the predictions (:297-305); here they stay on the device (`slider_guidance_loss`, p = NULL).

VAE encoding and image files are outside the hot path (no pretrained VAE can be fetched here): a folder entry is a
`.pt` file holding the already-encoded latent `[1,4,h,w]` (scaling factor applied) with the same file name in the
`-s` and `+s` folders; `--synthetic N` draws N seeded paired latents instead (BASELINE config 4).
"""
from __future__ import annotations

import argparse
import ast
import os
import random
from pathlib import Path

import torch

from . import config_util, model_util, ops, prompt_util, train_util
from .lora import DEFAULT_TARGET_REPLACE, UNET_TARGET_REPLACE_MODULE_CONV, LoRANetwork
from .prompt_util import PromptEmbedsCache, PromptEmbedsPair


def prev_step(model_output, timestep, scheduler, sample):
    """eta = 0 DDIM update, as restated in the reference script (:31-39); arithmetic on the device in one
    fused launch (`slider_cfg_step` with the two affine coefficients)."""
    prev_timestep = timestep - scheduler.config.num_train_timesteps // scheduler.num_inference_steps
    a_t = float(scheduler.alphas_cumprod[int(timestep)])
    a_p = float(scheduler.alphas_cumprod[int(prev_timestep)]) if prev_timestep >= 0 else float(scheduler.final_alpha_cumprod)
    c_x = (a_p / a_t) ** 0.5
    c_e = (1 - a_p) ** 0.5 - (a_p * (1 - a_t) / a_t) ** 0.5
    return ops.cfg_step(sample.contiguous(), model_output.contiguous(), model_output.contiguous(), None,
                        guidance=1.0, c_x=c_x, c_eps=c_e)


@torch.no_grad()
def get_noisy_latents(init_latents: torch.Tensor, generator, scheduler, total_timesteps: int, device, dtype):
    """`get_noisy_image` (trainscripts/imagesliders/train_util.py:200-235) from the VAE output onwards:
    noise = randn(generator); noisy = scheduler.add_noise(latents, noise, timesteps[t:t+1])."""
    noise = torch.randn(init_latents.shape, generator=generator).to(device, dtype)
    lat = init_latents.to(device, dtype)
    timestep = scheduler.timesteps[total_timesteps:total_timesteps + 1]
    return scheduler.add_noise(lat, noise, timestep), noise


class PairedLatents:
    """Source of (low, high) latent pairs per |scale|."""

    def __init__(self, folder_main=None, folders=None, scales=None, synthetic: int = 0, latent_hw: int = 32, seed: int = 0):
        self.scales = [float(s) for s in scales]
        self.synthetic = synthetic
        self.folder_main, self.folders = folder_main, list(folders) if folders is not None else None
        self.latent_hw, self.seed = latent_hw, seed

    def unique_scales(self):
        return sorted({abs(s) for s in self.scales})

    def sample(self, scale: float):
        if self.synthetic:
            idx = random.randint(0, self.synthetic - 1)
            g = torch.Generator().manual_seed(self.seed * 100003 + idx)
            base = torch.randn(1, 4, self.latent_hw, self.latent_hw, generator=g) * 0.18215
            delta = torch.randn(1, 4, self.latent_hw, self.latent_hw, generator=g) * 0.18215 * 0.1
            return base - scale * delta, base + scale * delta
        lo = self.folders[self.scales.index(-scale)]
        hi = self.folders[self.scales.index(scale)]
        names = sorted(n for n in os.listdir(f"{self.folder_main}/{lo}/") if n.endswith(".pt"))
        name = names[random.randint(0, len(names) - 1)]
        return torch.load(f"{self.folder_main}/{lo}/{name}"), torch.load(f"{self.folder_main}/{hi}/{name}")


class ImageSliderTrainer:
    """State of one image-slider run; `iteration(i)` is one pass of the reference loop body (:196-340)."""

    def __init__(self, config, prompts, device, data: PairedLatents, seed=None, engine: bool = False):
        self.config, self.device, self.data = config, torch.device(device), data
        self.use_engine, self.engine, self._staged_pair = engine, None, None
        self.weight_dtype = config_util.parse_precision(config.train.precision)
        self.save_weight_dtype = config_util.parse_precision(config.train.precision)
        if seed is not None:
            torch.manual_seed(seed)
            random.seed(seed)
        tokenizer, text_encoder, unet, sched = model_util.load_models(
            config.pretrained_model.name_or_path, scheduler_name=config.train.noise_scheduler,
            v2=config.pretrained_model.v2, v_pred=config.pretrained_model.v_pred)
        text_encoder.to(self.device)
        unet.to(self.device, dtype=self.weight_dtype)
        unet.requires_grad_(False)
        unet.eval()
        self.unet, self.scheduler = unet, sched
        modules = list(DEFAULT_TARGET_REPLACE)
        if config.network.type == "c3lier":
            modules += UNET_TARGET_REPLACE_MODULE_CONV
        self.network = LoRANetwork(unet, rank=config.network.rank, multiplier=1.0, alpha=config.network.alpha,
                                   target_replace=modules,
                                   train_method=config.network.training_method).to(self.device, dtype=self.weight_dtype)
        kwargs = {}
        if config.train.optimizer_args:
            for arg in config.train.optimizer_args.split(" "):
                key, value = arg.split("=")
                kwargs[key] = ast.literal_eval(value)
        self.optimizer = train_util.get_optimizer(config.train.optimizer)(self.network.prepare_optimizer_params(),
                                                                          lr=config.train.lr, **kwargs)
        self.lr_scheduler = train_util.get_lr_scheduler(config.train.lr_scheduler, self.optimizer,
                                                        max_iterations=config.train.iterations,
                                                        lr_min=config.train.lr / 100)
        criteria = torch.nn.MSELoss()
        cache = PromptEmbedsCache()
        self.prompt_pairs = []
        with torch.no_grad():
            for settings in prompts:
                for prompt in [settings.target, settings.positive, settings.neutral, settings.unconditional]:
                    if cache[prompt] is None:
                        cache[prompt] = model_util.encode_prompts(tokenizer, text_encoder, [prompt])
                self.prompt_pairs.append(PromptEmbedsPair(criteria, cache[settings.target], cache[settings.positive],
                                                          cache[settings.unconditional], cache[settings.neutral],
                                                          settings))
        self.last = None          # draws of the latest iteration (tests replay them through the oracle)

    def iteration(self, i: int):
        cfg, sched, net, dev, dt = self.config, self.scheduler, self.network, self.device, self.weight_dtype
        n_den = cfg.train.max_denoising_steps
        cat = lambda u, c, n: train_util.concat_embeddings(u, c, n).to(dev, dtype=dt)
        with torch.no_grad():
            sched.set_timesteps(n_den, device=dev)
            self.optimizer.zero_grad()
            pair = self.prompt_pairs[torch.randint(0, len(self.prompt_pairs), (1,)).item()]
            timesteps_to = torch.randint(1, n_den - 1, (1,)).item()
            scale_to_look = abs(random.choice(self.data.unique_scales()))
            lat_low, lat_high = self.data.sample(scale_to_look)
            noise_seed = random.randint(0, 2 * 15)
            if self.use_engine:
                return self._engine_iteration(pair, timesteps_to, scale_to_look, lat_low, lat_high, noise_seed)
            denoised_low, low_noise = get_noisy_latents(lat_low, torch.manual_seed(noise_seed), sched, timesteps_to,
                                                        dev, dt)
            denoised_high, high_noise = get_noisy_latents(lat_high, torch.manual_seed(noise_seed), sched, timesteps_to,
                                                          dev, dt)
            sched.set_timesteps(1000)
            t_cur = sched.timesteps[int(timesteps_to * 1000 / n_den)]
            # (the reference also evaluates two adapter-off predictions here, :254-278, and never uses them)
        self.last = dict(pair=pair, timesteps_to=timesteps_to, scale=scale_to_look, lat_low=lat_low, lat_high=lat_high,
                         noise_seed=noise_seed)
        net.set_lora_slider(scale=scale_to_look)
        with net:
            target_high = train_util.predict_noise(self.unet, sched, t_cur, denoised_high,
                                                   cat(pair.unconditional, pair.positive, pair.batch_size), guidance_scale=1)
        loss_high = prompt_util.plain_mse(target_high, high_noise)
        loss_high.backward()
        net.set_lora_slider(scale=-scale_to_look)
        with net:
            target_low = train_util.predict_noise(self.unet, sched, t_cur, denoised_low,
                                                  cat(pair.unconditional, pair.neutral, pair.batch_size), guidance_scale=1)
        loss_low = prompt_util.plain_mse(target_low, low_noise)
        loss_low.backward()              # NO zero_grad in between: the two sides accumulate (:327-334)
        self.optimizer.step()
        self.lr_scheduler.step()
        return loss_high.detach(), loss_low.detach()

    def _engine_iteration(self, pair, timesteps_to, scale, lat_low, lat_high, noise_seed):
        """Same draws, issued through engine.SliderEngine.image_slider_step: add_noise, the two grad predictions,
        both losses and backward passes replay as CUDA graphs, gradients land in the flat buffer, one fused AdamW."""
        from .engine import EngineConfig, SliderEngine
        cfg = self.config
        if cfg.train.optimizer.lower() != "adamw":
            raise NotImplementedError("engine=True implements AdamW (the reference default)")
        if self.engine is None:
            ecfg = EngineConfig(max_denoising_steps=cfg.train.max_denoising_steps, lr=cfg.train.lr, use_graphs=True)
            for k, v in self.optimizer.defaults.items():
                if k == "betas":
                    ecfg.betas = tuple(v)
                elif k == "eps":
                    ecfg.eps = float(v)
                elif k == "weight_decay":
                    ecfg.weight_decay = float(v)
            h, w = lat_low.shape[-2] * 8, lat_low.shape[-1] * 8
            self.engine = SliderEngine(self.unet, self.network, self.scheduler, ecfg, pair.batch_size, h, w,
                                       tuple(pair.target.shape[1:]), self.weight_dtype, self.device)
        if self._staged_pair is not pair:
            self.engine.load_image_pair_prompts(pair.positive, pair.neutral, pair.unconditional)
            self._staged_pair = pair
        # shared by both sides (:231); leaves the global generator where the eager loop's second draw leaves it
        noise = torch.randn(lat_low.shape, generator=torch.manual_seed(noise_seed))
        self.last = dict(pair=pair, timesteps_to=timesteps_to, scale=scale, lat_low=lat_low, lat_high=lat_high,
                         noise_seed=noise_seed)
        self.engine.lr = self.optimizer.param_groups[0]["lr"]
        with torch.enable_grad():
            loss_high, loss_low = self.engine.image_slider_step(lat_low, lat_high, noise, timesteps_to, scale)
        self.optimizer.step()               # no .grad attached (flat buffer): only keeps the LR-schedule bookkeeping
        self.lr_scheduler.step()
        return loss_high.clone(), loss_low.clone()

    def run(self, save_file: bool = True):
        cfg = self.config
        save_path = Path(cfg.save.path)
        losses = []
        for i in range(cfg.train.iterations):
            losses.append(self.iteration(i))
            if save_file and i % cfg.save.per_steps == 0 and i != 0 and i != cfg.train.iterations - 1:
                save_path.mkdir(parents=True, exist_ok=True)
                self.network.save_weights(save_path / f"{cfg.save.name}_{i}steps.pt", dtype=self.save_weight_dtype)
        if save_file:
            save_path.mkdir(parents=True, exist_ok=True)
            self.network.save_weights(save_path / f"{cfg.save.name}_last.pt", dtype=self.save_weight_dtype)
        return losses


def train(config, prompts, device, folder_main=None, folders=None, scales=None, synthetic: int = 0, seed=None,
          save_file: bool = True, engine: bool = False):
    if scales is None:
        raise TypeError("train() needs scales (and folders or synthetic=N)")       # upstream crashes here
    data = PairedLatents(folder_main, folders, scales, synthetic=synthetic, seed=seed or 0,
                         latent_hw=prompts[0].resolution // 8 if prompts else 32)
    trainer = ImageSliderTrainer(config, prompts, device, data, seed=seed, engine=engine)
    losses = trainer.run(save_file=save_file)
    return trainer.network, losses


def main(args):
    config = config_util.load_config_from_yaml(args.config_file)
    if args.name is not None:
        config.save.name = args.name
    attributes = [a.strip() for a in args.attributes.split(",")] if args.attributes is not None else []
    config.network.alpha = args.alpha
    config.network.rank = args.rank
    config.save.name += f"_alpha{args.alpha}"
    config.save.name += f"_rank{config.network.rank}"
    config.save.name += f"_{config.network.training_method}"
    config.save.path += f"/{config.save.name}"
    prompts = prompt_util.load_prompts_from_yaml(config.prompts_file, attributes)
    device = torch.device(f"cuda:{args.device}")
    folders = [f.strip() for f in args.folders.split(",")]
    scales = [int(s.strip()) for s in args.scales.split(",")]
    if len(scales) != len(folders):
        raise Exception("the number of folders need to match the number of scales")
    train(config, prompts, device, folder_main=args.folder_main, folders=folders, scales=scales, synthetic=args.synthetic,
          engine=args.engine)


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser()
    parser.add_argument("--config_file", required=False, default="data/config.yaml", help="Config file for training.")
    parser.add_argument("--alpha", type=float, required=True, help="LoRA weight.")
    parser.add_argument("--rank", type=int, required=False, default=4, help="Rank of LoRA.")
    parser.add_argument("--device", type=int, required=False, default=0, help="Device to train on.")
    parser.add_argument("--name", type=str, required=False, default=None, help="Name of the slider.")
    parser.add_argument("--attributes", type=str, required=False, default=None,
                        help="attributes to disentangle (comma separated string)")
    parser.add_argument("--folder_main", type=str, required=False, default=None,
                        help="The folder to check (holds one sub-folder of encoded latents per scale)")
    parser.add_argument("--stylecheck", type=str, required=False, default=None, help="(accepted for CLI parity; unused)")
    parser.add_argument("--folders", type=str, required=False, default="verylow, low, high, veryhigh",
                        help="folders with different attribute-scaled images")
    parser.add_argument("--scales", type=str, required=False, default="-2, -1,1, 2",
                        help="scales for different attribute-scaled images")
    parser.add_argument("--engine", action="store_true",
                        help="B200 extension: CUDA-graph engine + flat AdamW for the iteration body")
    parser.add_argument("--synthetic", type=int, default=0, help="draw N seeded synthetic latent pairs instead of folders")
    return parser


if __name__ == "__main__":
    main(build_parser().parse_args())
