"""SliderEngine — one slider training iteration (reference conceptmod/textsliders/train_lora.py:155-309)
as three replayable CUDA graphs over static buffers.

  G_denoise : adapter ON,  CFG UNet forward (batch 2b) + fused CFG/scheduler step   (train_lora.py:193-207)
  G_predict : adapter OFF, the three no-grad predictions as one batch-6b forward   (train_lora.py:216-252)
  G_train   : adapter ON,  grad forward + guidance loss + backward into the flat
              LoRA-gradient buffer                                                 (train_lora.py:261-298)
followed by (multi-GPU) one all-reduce of the flat gradient buffer and one fused AdamW launch.

The arithmetic per sample is the reference's; what changes is how it is issued: no per-iteration
host synchronisation, timestep and scheduler coefficients live in device buffers so the same graph
serves every denoising step, and the three adapter-off predictions share one launch sequence.
"""
from __future__ import annotations

import os

from dataclasses import dataclass

import torch
import torch.distributed as dist

from . import ops, train_util
from .lora import LoRANetwork
from .prompt_util import guidance_mse


@dataclass
class EngineConfig:
    max_denoising_steps: int = 50
    denoise_guidance: float = 3.0       # train_lora.py:206
    lr: float = 2e-4
    betas: tuple = (0.9, 0.999)
    eps: float = 1e-8
    weight_decay: float = 1e-2          # torch.optim.AdamW default
    max_grad_norm: float | None = None  # train_lora_xl.py:349 uses 0.2
    use_graphs: bool = True
    # predictions on a parallel graph branch beside the grad forward (SLIDER_DISABLE=ptrain: A/B switch)
    overlap_predict: bool = "ptrain" not in os.environ.get("SLIDER_DISABLE", "")


class SliderEngine:
    def __init__(self, unet, network: LoRANetwork, scheduler, cfg: EngineConfig, batch_size: int, height: int,
                 width: int, embed_shape: tuple, dtype: torch.dtype, device, xl_cond: dict | None = None):
        self.unet, self.network, self.sched, self.cfg = unet, network, scheduler, cfg
        self.device, self.dtype = torch.device(device), dtype
        self.b = batch_size
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        if network.flat_param is None:
            network.flatten()
        n = network.flat_param.numel()
        self.m = torch.zeros(n, device=self.device, dtype=torch.float32)
        self.v = torch.zeros(n, device=self.device, dtype=torch.float32)
        self.opt_step = 0
        self.lr = cfg.lr
        lat_shape = (batch_size, 4, height // 8, width // 8)
        mk = lambda *s, dt=dtype: torch.zeros(*s, device=self.device, dtype=dt)
        self.lat = mk(*lat_shape)
        self.emb_den = mk(2 * batch_size, *embed_shape)
        self.emb_pred = mk(6 * batch_size, *embed_shape)
        self.emb_tgt = mk(2 * batch_size, *embed_shape)
        # per-step scheduler state: ONE 12-float device row {t, -, -, -, g, c_x, c_eps, c_noise, _, input scale, 0, 0} so
        # that a denoising step costs one small device copy (row of the table -> here) before its graph replay
        self.step_row = torch.zeros(12, device=self.device, dtype=torch.float32)
        self.step_row[8:12] = 1.0
        self.t_dev = self.step_row[0:1]
        self.coef = self.step_row[4:8]
        self.in_scale = self.step_row[8:12]                                      # {_, c_x, 0, 0} for scale_model_input
        self.positive = mk(*lat_shape)
        self.neutral = mk(*lat_shape)
        self.uncond = mk(*lat_shape)
        self.loss = torch.zeros((), device=self.device, dtype=torch.float32)
        self.xl = None
        if xl_cond is not None:   # {"text_embeds": [1,1280] pair rows..., "time_ids": [1,6]} filled per step
            self.xl = {"den": {k: mk(2 * batch_size, v) for k, v in xl_cond.items()},
                       "pred": {k: mk(6 * batch_size, v) for k, v in xl_cond.items()},
                       "tgt": {k: mk(2 * batch_size, v) for k, v in xl_cond.items()}}
        self.guidance = 1.0
        self.sign = 1
        self._graphs: dict[str, torch.cuda.CUDAGraph] = {}
        self._graph_launches: dict[str, int] = {}
        self.kernel_launches = 0        # libslider_b200 kernels issued (directly or by graph replay)
        self._tables = {}
        self._needs_scale = hasattr(scheduler, "_sigmas")        # Euler-ancestral divides the model input
        if hasattr(unet, "enable_context_cache"):      # cross-attention K/V once per prompt pair (see unet.py)
            unet.enable_context_cache([self.emb_den, self.emb_pred, self.emb_tgt])
        self._stochastic = any(self.sched.step_coefficients(t)[2] != 0.0 for t in self._table(cfg.max_denoising_steps)[3][:2])

    # ---- scheduler tables on the device (one row per denoising step) ---------------------------------
    def _table(self, n_steps: int):
        tab = self._tables.get(n_steps)
        if tab is None:
            self.sched.set_timesteps(n_steps)
            ts = [float(t) for t in self.sched.timesteps]
            rows, scales = [], []
            for t in ts:
                c_x, c_e, c_n = self.sched.step_coefficients(t)
                rows.append([self.cfg.denoise_guidance, c_x, c_e, c_n])
                scales.append([0.0, self.sched.input_scale(t), 0.0, 0.0])
            t_t = torch.tensor(ts, device=self.device, dtype=torch.float32)
            c_t = torch.tensor(rows, device=self.device, dtype=torch.float32)
            s_t = torch.tensor(scales, device=self.device, dtype=torch.float32)
            packed = torch.zeros(len(ts), 12, device=self.device, dtype=torch.float32)       # rows of `step_row`
            packed[:, 0] = t_t; packed[:, 4:8] = c_t; packed[:, 8:12] = s_t
            tab = (t_t, c_t, s_t, ts, packed)
            self._tables[n_steps] = tab
        return tab

    # ---- graph bodies ------------------------------------------------------------------------------
    def _unet(self, x, emb, cond):
        kw = {} if cond is None else {"added_cond_kwargs": cond}
        return self.unet(x, self.t_dev, encoder_hidden_states=emb, **kw).sample

    def _model_input(self, reps: int):
        x = torch.cat([self.lat] * reps)
        if self._needs_scale:
            x = ops.cfg_step(x, None, None, None, dev_coef=self.in_scale)
        return x

    def _body_denoise(self):
        with torch.no_grad(), self.network:
            eps = self._unet(self._model_input(2), self.emb_den, self.xl and self.xl["den"])
            b = self.b
            noise = None
            if self._stochastic:
                noise = torch.randn_like(self.lat)
            ops.cfg_step(self.lat, eps[:b], eps[b:], noise, dev_coef=self.coef, out=self.lat)

    def _body_predict(self):
        with torch.no_grad():
            eps = self._unet(self._model_input(6), self.emb_pred, self.xl and self.xl["pred"])
            b = self.b
            for k, out in enumerate((self.positive, self.neutral, self.uncond)):
                ops.cfg_step(None, eps[k * b:(k + 1) * b], eps[(3 + k) * b:(4 + k) * b], guidance=1.0, out=out)

    def _body_train(self):
        self.network.flat_grad.zero_()
        with self.network.packed_transposes():
            with self.network:
                eps = self._unet(self._model_input(2), self.emb_tgt, self.xl and self.xl["tgt"])
                target = train_util._CFGCombineFn.apply(eps, 1.0)
            loss = guidance_mse(target, self.positive, self.neutral, self.uncond, self.guidance, self.sign)
            loss.backward()
        self.loss.copy_(loss.detach())

    def _body_predict_train(self):
        """The three adapter-off predictions (one batch-6 forward) do not depend on the grad forward: they run on a
        second stream — a parallel branch of the captured graph, with its own scratch lane — and join before the
        loss needs them (train_lora.py:216-289 evaluates them one after the other)."""
        main = torch.cuda.current_stream()
        if not hasattr(self, "_predict_stream"):
            self._predict_stream = torch.cuda.Stream(device=self.device)
        side = self._predict_stream
        self.network.flat_grad.zero_()
        side.wait_stream(main)
        with torch.cuda.stream(side), ops.workspace_lane(1):
            self._body_predict()
        with self.network.packed_transposes():
            with self.network:
                eps = self._unet(self._model_input(2), self.emb_tgt, self.xl and self.xl["tgt"])
                target = train_util._CFGCombineFn.apply(eps, 1.0)
            main.wait_stream(side)
            loss = guidance_mse(target, self.positive, self.neutral, self.uncond, self.guidance, self.sign)
            loss.backward()
        self.loss.copy_(loss.detach())

    def _run(self, name: str, body):
        if not self.cfg.use_graphs:
            n0 = ops.launch_count()
            body()
            self.kernel_launches += ops.launch_count() - n0
            return
        g = self._graphs.get(name)
        if g is None:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):      # warm-up outside capture (lazy inits, cudaFuncSetAttribute, autotune)
                    body()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            n0 = ops.launch_count()
            with torch.cuda.graph(g):
                body()
            self._graph_launches[name] = ops.launch_count() - n0
            self._graphs[name] = g
            return              # capture does not execute; the caller replays
        g.replay()
        self.kernel_launches += self._graph_launches[name]

    def _exec(self, name, body, restore=()):
        first = self.cfg.use_graphs and name not in self._graphs
        saved = [t.clone() for t in restore] if first else []
        self._run(name, body)
        if first:       # warm-up ran the body for real: put the in-place state back, then replay once
            for t, s in zip(restore, saved):
                t.copy_(s)
            self._graphs[name].replay()
            self.kernel_launches += self._graph_launches[name]

    # ---- one reference iteration ---------------------------------------------------------------------
    def load_pair(self, target, positive, neutral, unconditional, guidance_scale: float, action: str, xl=None):
        """Stage one prompt pair's embeddings ([1,77,D] each) into the static buffers."""
        b = self.b
        rep = lambda e: e.to(self.device, self.dtype).repeat_interleave(b, dim=0)
        u, t, p, n = rep(unconditional), rep(target), rep(positive), rep(neutral)
        self.emb_den.copy_(torch.cat([u, t]))
        self.emb_tgt.copy_(torch.cat([u, t]))
        self.emb_pred.copy_(torch.cat([u, u, u, p, n, u]))
        if hasattr(self.unet, "refresh_context_cache"):
            self.unet.refresh_context_cache()
        if (guidance_scale, action) != (self.guidance, "enhance" if self.sign > 0 else "erase"):
            self.guidance, self.sign = float(guidance_scale), (1 if action == "enhance" else -1)
            self._graphs.pop("train", None)     # host scalars are baked into the captured launch
            self._graphs.pop("predict_train", None)
        if self.xl is not None and xl is not None:
            for key in self.xl["den"]:
                cu, ct, cp, cn = (rep(xl[key][r]) for r in ("unconditional", "target", "positive", "neutral"))
                self.xl["den"][key].copy_(torch.cat([cu, ct]))
                self.xl["tgt"][key].copy_(torch.cat([cu, ct]))
                self.xl["pred"][key].copy_(torch.cat([cu, cu, cu, cp, cn, cu]))

    def train_step(self, latents: torch.Tensor, timesteps_to: int, optimizer: bool = True) -> torch.Tensor:
        """latents: initial noise [b,4,h/8,w/8] (host or device).  Returns the device loss (no sync).
        `optimizer=False` stops after the backward: the sample's gradient is left in `network.flat_grad` (gradient
        accumulation over several samples before one `optimizer_step()`; the data-parallel equivalence check)."""
        n_den = self.cfg.max_denoising_steps
        packed = self._table(n_den)[4]
        self.lat.copy_(latents.to(self.dtype), non_blocking=True)
        for i in range(timesteps_to):                                   # train_lora.py:195-207
            self.step_row.copy_(packed[i])
            self._exec("denoise", self._body_denoise, restore=(self.lat,))
        t1000, _, s1000, ts = self._table(1000)[:4]                    # train_lora.py:209-213
        idx = int(timesteps_to * 1000 / n_den)
        self.t_dev.copy_(t1000[idx:idx + 1]); self.in_scale.copy_(s1000[idx])
        if self.cfg.use_graphs and self.cfg.overlap_predict:
            self._exec("predict_train", self._body_predict_train)
        else:
            self._exec("predict", self._body_predict)
            self._exec("train", self._body_train)
        if optimizer:
            self.optimizer_step()
        return self.loss

    # ---- image-pair slider iteration (trainscripts/imagesliders/train_lora-scale.py:196-340) ------------------
    def load_image_pair_prompts(self, positive, neutral, unconditional):
        """Stage the two CFG embedding batches of the image-slider objective: [uncond, positive] for the "high"
        side, [uncond, neutral] for the "low" side."""
        b = self.b
        rep = lambda e: e.to(self.device, self.dtype).repeat_interleave(b, dim=0)
        u, p, n = rep(unconditional), rep(positive), rep(neutral)
        self.emb_den.copy_(torch.cat([u, p]))          # high side
        self.emb_tgt.copy_(torch.cat([u, n]))          # low side
        if hasattr(self.unet, "refresh_context_cache"):
            self.unet.refresh_context_cache()

    def _body_image_side(self, emb, scale: float):
        """noisy = add_noise(side latents, noise, t) -> grad prediction with the adapter at `scale` -> MSE against
        the noise that was added -> backward into the flat gradient buffer (accumulating)."""
        noisy = ops.cfg_step(self.lat, None, None, self.noise, dev_coef=self.coef)
        self.network.set_lora_slider(scale)
        from .prompt_util import plain_mse
        with self.network.packed_transposes():
            with self.network:
                x = torch.cat([noisy] * 2)
                if self._needs_scale:
                    x = ops.cfg_step(x, None, None, None, dev_coef=self.in_scale)
                eps = self._unet(x, emb, None)
                target = train_util._CFGCombineFn.apply(eps, 1.0)
            loss = plain_mse(target, self.noise)
            loss.backward()
        return loss.detach()

    def image_slider_step(self, lat_low: torch.Tensor, lat_high: torch.Tensor, noise: torch.Tensor, timesteps_to: int,
                          scale: float):
        """One optimizer step of the image-pair objective: both sides share `noise` and the timestep; the adapter
        runs at +scale on the high latents / positive prompt and at -scale on the low latents / neutral prompt; the
        two backward passes accumulate before ONE AdamW launch.  Each (side, scale) body is a replayed CUDA graph
        (the scale is a host scalar of the captured launches).  Returns (loss_high, loss_low) device scalars."""
        if not hasattr(self, "noise"):
            self.noise = torch.zeros_like(self.lat)
            self.loss_pair = torch.zeros(2, device=self.device, dtype=torch.float32)
        n_den = self.cfg.max_denoising_steps
        t_add = self._table(n_den)[3][timesteps_to]
        if self._needs_scale:
            raise NotImplementedError("image_slider_step: add_noise of the sigma-parameterised schedulers is not tabled")
        a = float(self.sched._ac[int(t_add)])
        t1000, _, s1000, _ = self._table(1000)[:4]
        idx = int(timesteps_to * 1000 / n_den)
        self.t_dev.copy_(t1000[idx:idx + 1]); self.in_scale.copy_(s1000[idx])
        self.coef.copy_(torch.tensor([0.0, a ** 0.5, 0.0, (1.0 - a) ** 0.5]), non_blocking=True)
        self.noise.copy_(noise.to(self.dtype), non_blocking=True)
        self.network.flat_grad.zero_()
        for k, (lat, emb, sc) in enumerate(((lat_high, self.emb_den, float(scale)), (lat_low, self.emb_tgt, -float(scale)))):
            self.lat.copy_(lat.to(self.dtype), non_blocking=True)
            name = f"image_{k}_{sc:+.4f}"
            self._exec(name, lambda emb=emb, sc=sc, k=k: self.loss_pair[k:k + 1].copy_(self._body_image_side(emb, sc).reshape(1)),
                       restore=(self.network.flat_grad,))      # the warm-up passes of a first capture must not accumulate
        self.optimizer_step()
        return self.loss_pair[0], self.loss_pair[1]

    def optimizer_step(self):
        g = self.network.flat_grad
        pre = 1.0
        if self.world > 1:
            dist.all_reduce(g, op=dist.ReduceOp.SUM)
            pre = 1.0 / self.world
        scale = None
        if self.cfg.max_grad_norm is not None:
            scale = ops.clip_coef(g, self.cfg.max_grad_norm, pre)
        elif pre != 1.0:
            scale = torch.full((), pre, device=self.device, dtype=torch.float32)
        self.opt_step += 1
        self.kernel_launches += 1 + (1 if self.cfg.max_grad_norm is not None else 0)
        ops.adamw_flat(self.network.flat_param, g, self.m, self.v, self.lr, self.cfg.betas[0], self.cfg.betas[1],
                       self.cfg.eps, self.cfg.weight_decay, self.opt_step, scale)
