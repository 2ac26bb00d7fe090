"""GPU parity of the drop-in trainers and the inference sweep against the CPU oracle loops
(oracle/slider_oracle.py): SD-1.x text slider (train_lora.py), SD-XL text slider with Euler-ancestral sampling and
gradient clipping (train_lora_xl.py), image-pair slider with two-sided gradient accumulation
(train_lora-scale.py), the `--engine` CUDA-graph loop, and generate_images' per-step adapter switch.

Tolerances (BASELINE.json north_star): loss rel <= 1e-3 in fp32; LoRA gradients and post-step weights
cosine >= 0.999."""
import random
import sys
from pathlib import Path

import pytest
import torch
import yaml

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def _config(tmp_path, model="tiny-sd", scheduler="ddim", iterations=3, precision="float32", negative=False, cfg=1.0):
    from sliders_conceptmod_b200 import config_util, prompt_util
    prompts = [{"target": "person", "positive": "old person", "unconditional": "young person", "neutral": "person",
                "action": "enhance", "guidance_scale": 4, "resolution": 128, "batch_size": 1,
                "dynamic_resolution": False},
               {"target": "dog", "positive": "big dog", "unconditional": "small dog", "neutral": "dog",
                "action": "erase", "guidance_scale": 2, "resolution": 128, "batch_size": 1,
                "dynamic_resolution": False}]
    if negative:
        for p in prompts:
            p["negative"] = "bad " + p["target"]
    (tmp_path / "prompts.yaml").write_text(yaml.safe_dump(prompts))
    c = {"prompts_file": str(tmp_path / "prompts.yaml"),
         "pretrained_model": {"name_or_path": model, "v2": False, "v_pred": False},
         "network": {"type": "c3lier", "rank": 4, "alpha": 1.0, "training_method": "noxattn"},
         "train": {"precision": precision, "noise_scheduler": scheduler, "iterations": iterations, "lr": 2e-4,
                   "optimizer": "AdamW", "lr_scheduler": "constant", "max_denoising_steps": 10, "cfg": cfg},
         "save": {"name": "temp", "path": str(tmp_path / "models"), "per_steps": 500, "precision": precision},
         "logging": {"use_wandb": False, "verbose": False}, "other": {"use_xformers": False}}
    (tmp_path / "config.yaml").write_text(yaml.safe_dump(c))
    config = config_util.load_config_from_yaml(str(tmp_path / "config.yaml"))
    return config, prompt_util.load_prompts_from_yaml(config.prompts_file, [])


def _randomize_up(network, std=0.05, seed=9):
    """lora_up starts at zero (lora.py:84); give it values so that dA and the adapter output are exercised."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for m in network.unet_loras:
            m.lora_up.weight.copy_((torch.randn(m.lora_up.weight.shape, generator=g) * std).to(m.lora_up.weight))


def _oracle_copy(unet, network, **net_kw):
    """CPU fp32 copies of (unet, network) wired to the oracle's adapter arithmetic."""
    from oracle import slider_oracle
    from sliders_conceptmod_b200.lora import LoRANetwork
    from sliders_conceptmod_b200.unet import UNet2DConditionModel
    unet_c = UNet2DConditionModel(unet.config).eval().requires_grad_(False)
    unet_c.load_state_dict({k: v.detach().float().cpu() for k, v in unet.state_dict().items()})
    net_c = LoRANetwork(unet_c, **net_kw)
    net_c.load_state_dict({k: v.detach().float().cpu() for k, v in network.state_dict().items()})
    slider_oracle.attach(net_c)
    return unet_c, net_c


def _flat(network, what):
    ps = dict(network.named_parameters())
    pick = (lambda p: p.grad) if what == "grad" else (lambda p: p.data)
    return torch.cat([pick(ps[k]).detach().float().flatten().cpu() for k in sorted(ps)])


def _cpu(e):
    return e.detach().float().cpu()


def test_sd1_trainer_iterations_match_oracle(tmp_path):
    from helpers import cosine
    from oracle import slider_oracle
    from sliders_conceptmod_b200 import train_util, workload
    from sliders_conceptmod_b200.slider_loop import TextSliderTrainer
    config, prompts = _config(tmp_path)
    tr = TextSliderTrainer(config, prompts, "cuda", seed=1)
    _randomize_up(tr.network)
    unet_c, net_c = _oracle_copy(tr.unet, tr.network, rank=4, multiplier=1.0, alpha=1.0,
                                 target_replace=workload.target_modules(True), train_method="noxattn")
    opt_c = torch.optim.AdamW(net_c.prepare_optimizer_params(), lr=config.train.lr)
    osch = slider_oracle.OracleScheduler("ddim")
    for it in range(3):
        torch.manual_seed(100 + it)
        loss_g = tr.iteration(it)
        # replay the host draws of the iteration (train_lora.py:163-193) for the oracle
        torch.manual_seed(100 + it)
        pair = tr.prompt_pairs[torch.randint(0, len(tr.prompt_pairs), (1,)).item()]
        timesteps_to = torch.randint(1, 10, (1,)).item()
        osch.set_timesteps(10)
        latents = train_util.get_initial_latents(osch, 1, 128, 128, 1)
        emb = {k: _cpu(getattr(pair, k)) for k in ("target", "positive", "neutral", "unconditional")}
        opt_c.zero_grad()
        loss_c, _ = slider_oracle.train_iteration(unet_c, net_c, osch, emb, latents, timesteps_to, pair.guidance_scale,
                                                  pair.action, max_denoising_steps=10)
        rel = abs(loss_g.item() - loss_c.item()) / abs(loss_c.item())
        cg = cosine(_flat(tr.network, "grad"), _flat(net_c, "grad"))
        opt_c.step()
        cw = cosine(_flat(tr.network, "data"), _flat(net_c, "data"))
        print(f"iter {it}: t_to {timesteps_to} {pair.action} loss {loss_g.item():.6e} vs {loss_c.item():.6e} "
              f"rel {rel:.2e} grad cos {cg:.6f} weight cos {cw:.6f}")
        assert rel <= 1e-3 and cg >= 0.999 and cw >= 0.999


def test_xl_trainer_euler_a_clip_matches_oracle(tmp_path):
    from helpers import cosine
    from oracle import slider_oracle
    from sliders_conceptmod_b200 import train_util
    from sliders_conceptmod_b200.slider_loop import TextSliderTrainer
    config, prompts = _config(tmp_path, model="tiny-xl", scheduler="euler_a", negative=True, cfg=2.0)
    tr = TextSliderTrainer(config, prompts, "cuda", xl=True, rank=4, seed=1)
    _randomize_up(tr.network)
    unet_c, net_c = _oracle_copy(tr.unet, tr.network, rank=4, multiplier=1.0, delimiter="_",
                                 target_replace=["Attention"], prefix="lora_unet", train_method="noxattn")
    opt_c = torch.optim.AdamW(net_c.parameters(), lr=1e-4, weight_decay=1e-6)        # train_lora_xl.py:104
    osch = slider_oracle.OracleScheduler("euler_a")
    for it in range(2):
        torch.manual_seed(200 + it)
        torch.cuda.manual_seed(300 + it)
        loss_g = tr.iteration(it)
        torch.manual_seed(200 + it)
        pair = tr.prompt_pairs[torch.randint(0, len(tr.prompt_pairs), (1,)).item()]
        timesteps_to = torch.randint(1, 10, (1,)).item()
        osch.set_timesteps(10)
        latents = train_util.get_initial_latents(osch, 1, 128, 128, 1)
        # the ancestral noise is drawn on the device, one tensor per denoising step, in order
        torch.cuda.manual_seed(300 + it)
        noises = [torch.randn(latents.shape, device="cuda", dtype=torch.float32).cpu() for _ in range(timesteps_to)]
        ids = train_util.get_add_time_ids(128, 128, dynamic_crops=False)
        text = {k: _cpu(getattr(pair, k).text_embeds) for k in ("target", "positive", "neutral", "unconditional", "negative")}
        pooled = {k: _cpu(getattr(pair, k).pooled_embeds) for k in text}
        opt_c.zero_grad()
        loss_c, _ = slider_oracle.train_iteration(
            unet_c, net_c, osch, text, latents, timesteps_to, pair.guidance_scale, pair.action, max_denoising_steps=10,
            denoise_guidance=2.0, predict_guidance=2.0, noises=noises, pooled=pooled, time_ids=ids)
        rel = abs(loss_g.item() - loss_c.item()) / abs(loss_c.item())
        torch.nn.utils.clip_grad_norm_(net_c.parameters(), max_norm=0.2)               # train_lora_xl.py:349
        cg = cosine(_flat(tr.network, "grad"), _flat(net_c, "grad"))
        gn = (_flat(tr.network, "grad").norm().item(), _flat(net_c, "grad").norm().item())
        opt_c.step()
        cw = cosine(_flat(tr.network, "data"), _flat(net_c, "data"))
        print(f"iter {it}: t_to {timesteps_to} loss {loss_g.item():.6e} vs {loss_c.item():.6e} rel {rel:.2e} "
              f"grad cos {cg:.6f} clipped norms {gn[0]:.4f}/{gn[1]:.4f} weight cos {cw:.6f}")
        assert rel <= 1e-3 and cg >= 0.999 and cw >= 0.999
        assert abs(gn[0] - gn[1]) <= 1e-2 * gn[1]


def test_image_slider_two_sided_accumulation_matches_oracle(tmp_path):
    from helpers import cosine
    from oracle import slider_oracle
    from sliders_conceptmod_b200 import workload
    from sliders_conceptmod_b200.train_lora_scale import ImageSliderTrainer, PairedLatents
    config, prompts = _config(tmp_path)
    data = PairedLatents(scales=[-2, -1, 1, 2], synthetic=4, latent_hw=16, seed=3)
    tr = ImageSliderTrainer(config, prompts, "cuda", data, seed=3)
    _randomize_up(tr.network)
    unet_c, net_c = _oracle_copy(tr.unet, tr.network, rank=4, multiplier=1.0, alpha=1.0,
                                 target_replace=workload.target_modules(True), train_method="noxattn")
    opt_c = torch.optim.AdamW(net_c.prepare_optimizer_params(), lr=config.train.lr)
    osch = slider_oracle.OracleScheduler("ddim")
    for it in range(3):
        lh_g, ll_g = tr.iteration(it)
        d = tr.last
        noise = torch.randn(d["lat_low"].shape, generator=torch.Generator().manual_seed(d["noise_seed"]))
        pair = d["pair"]
        emb = {k: _cpu(getattr(pair, k)) for k in ("positive", "neutral", "unconditional")}
        opt_c.zero_grad()
        lh_c, ll_c = slider_oracle.image_slider_iteration(unet_c, net_c, osch, emb, d["lat_low"], d["lat_high"], noise,
                                                          d["timesteps_to"], d["scale"], max_denoising_steps=10)
        rh = abs(lh_g.item() - lh_c.item()) / abs(lh_c.item())
        rl = abs(ll_g.item() - ll_c.item()) / abs(ll_c.item())
        cg = cosine(_flat(tr.network, "grad"), _flat(net_c, "grad"))
        opt_c.step()
        cw = cosine(_flat(tr.network, "data"), _flat(net_c, "data"))
        print(f"iter {it}: scale {d['scale']} t_to {d['timesteps_to']} high {lh_g.item():.6e}/{lh_c.item():.6e} "
              f"low {ll_g.item():.6e}/{ll_c.item():.6e} grad cos {cg:.6f} weight cos {cw:.6f}")
        assert rh <= 1e-3 and rl <= 1e-3 and cg >= 0.999 and cw >= 0.999


def test_image_slider_train_entry_runs_and_saves(tmp_path):
    from sliders_conceptmod_b200 import train_lora_scale
    config, prompts = _config(tmp_path, iterations=2, precision="bfloat16")
    network, losses = train_lora_scale.train(config, prompts, "cuda", scales=[-1, 1], synthetic=2, seed=5)
    assert len(losses) == 2 and all(torch.isfinite(l).item() for pair in losses for l in pair)
    sd = torch.load(Path(config.save.path) / "temp_last.pt")
    assert set(sd) == set(network.state_dict()) and all(v.dtype == torch.bfloat16 for v in sd.values())


@pytest.mark.parametrize("scheduler,steps", [("ddim", 8), ("euler_a", 6), ("lms", 8)])
def test_inference_sweep_matches_oracle(scheduler, steps):
    from helpers import build_pair
    from oracle import slider_oracle
    from sliders_conceptmod_b200 import generate_images, model_util
    from sliders_conceptmod_b200.unet import UNetConfig
    cfg = UNetConfig.tiny()
    unet_c, net_c, unet_g, net_g = build_pair(cfg, "cuda", torch.float32)
    g = torch.Generator().manual_seed(4)
    emb = torch.randn(2, 77, cfg.cross_attention_dim, generator=g)          # [uncond, cond]
    lat = torch.randn(1, 4, 16, 16, generator=g)
    sched = model_util.create_noise_scheduler(scheduler)
    osch = slider_oracle.OracleScheduler(scheduler)
    sched.set_timesteps(steps); osch.set_timesteps(steps)
    lat_c = lat * float(osch.init_noise_sigma)
    lat_g = (lat * float(sched.init_noise_sigma)).cuda()
    noises = None
    gen = None
    if scheduler == "euler_a":
        gen = torch.Generator(device="cuda").manual_seed(77)
        g2 = torch.Generator(device="cuda").manual_seed(77)
        noises = [torch.randn(lat.shape, generator=g2, device="cuda").cpu() for _ in range(steps)]
    for scale in (2.0, -1.0):
        if gen is not None:
            gen.manual_seed(77)
        out_g = generate_images.denoise_with_slider(unet_g, net_g, sched, emb.cuda(), lat_g.clone(), scale, 500, 7.5, steps,
                                                    noise_generator=gen)
        out_c = slider_oracle.inference_sweep(unet_c, net_c, osch, emb, lat_c.clone(), scale, 500, 7.5, steps, noises=noises)
        err = (out_g.float().cpu() - out_c).abs().max().item()
        ref = out_c.abs().max().item()
        print(f"{scheduler} scale {scale:+.1f}: max-abs {err:.3e} of {ref:.3e}")
        assert err <= 5e-3 * ref
    # the adapter really is switched per step: scale 0 and scale 2 differ, and differ from each other only
    # through the steps with t <= start_noise
    base = generate_images.denoise_with_slider(unet_g, net_g, sched, emb.cuda(), lat_g.clone(), 0.0, 500, 7.5, steps,
                                               noise_generator=gen.manual_seed(77) if gen is not None else None)
    late = generate_images.denoise_with_slider(unet_g, net_g, sched, emb.cuda(), lat_g.clone(), 2.0, -1, 7.5, steps,
                                               noise_generator=gen.manual_seed(77) if gen is not None else None)
    assert torch.equal(base, late)          # start_noise below every t: adapter never on == scale 0


def test_sweep_deals_scales_to_ranks():
    from helpers import build_pair
    from sliders_conceptmod_b200 import generate_images, model_util
    from sliders_conceptmod_b200.unet import UNetConfig
    cfg = UNetConfig.tiny()
    _, _, unet_g, net_g = build_pair(cfg, "cuda", torch.bfloat16)
    g = torch.Generator().manual_seed(4)
    pe = torch.randn(1, 77, cfg.cross_attention_dim, generator=g).cuda().bfloat16()
    ue = torch.randn(1, 77, cfg.cross_attention_dim, generator=g).cuda().bfloat16()
    sched = model_util.create_noise_scheduler("ddim")
    scales = [-2.0, -1.0, 0.0, 1.0, 2.0]
    full = generate_images.sweep(unet_g, net_g, sched, pe, ue, scales, seed=3, height=128, width=128, start_noise=600,
                                 num_inference_steps=5)
    parts = {}
    for r in range(2):
        parts.update(generate_images.sweep(unet_g, net_g, sched, pe, ue, scales, seed=3, height=128, width=128,
                                           start_noise=600, num_inference_steps=5, rank=r, world=2))
    assert sorted(parts) == sorted(full) == scales
    for s in scales:
        assert torch.equal(parts[s], full[s])
    assert not torch.equal(full[2.0], full[-2.0])


def test_engine_loop_matches_eager_trainer(tmp_path):
    """`train_lora --engine` (CUDA graphs + flat AdamW) against the eager drop-in loop, same seeds."""
    from helpers import cosine
    from sliders_conceptmod_b200 import engine_loop
    from sliders_conceptmod_b200.slider_loop import TextSliderTrainer
    config, prompts = _config(tmp_path, iterations=4)
    prompts = prompts[:1]
    outs = []
    for use_engine in (False, True):
        tr = TextSliderTrainer(config, prompts, "cuda", seed=1)
        _randomize_up(tr.network)
        torch.manual_seed(42)
        if use_engine:
            engine_loop.run_with_engine(tr, suffix=".pt")
        else:
            tr.run(save_file=False)
        outs.append(torch.cat([v.float().flatten().cpu() for _, v in sorted(tr.network.state_dict().items())]))
    cw = cosine(outs[0], outs[1])
    rel = (outs[0] - outs[1]).norm().item() / outs[0].norm().item()
    print(f"engine vs eager after 4 iterations: weight cos {cw:.7f} rel diff {rel:.2e}")
    assert cw >= 0.999 and rel <= 1e-2
    assert (Path(config.save.path) / "temp_last.pt").exists()


def test_image_slider_engine_matches_eager_loop(tmp_path):
    """`train_lora-scale.py --engine` (graph-replayed sides + flat AdamW) against the eager drop-in loop, same seeds."""
    import random
    from helpers import cosine
    from sliders_conceptmod_b200.train_lora_scale import ImageSliderTrainer, PairedLatents
    config, prompts = _config(tmp_path)
    outs, losses = [], []
    for use_engine in (False, True):
        data = PairedLatents(scales=[-2, -1, 1, 2], synthetic=4, latent_hw=16, seed=3)
        tr = ImageSliderTrainer(config, prompts[:1], "cuda", data, seed=3, engine=use_engine)
        _randomize_up(tr.network)
        torch.manual_seed(21); random.seed(21)
        ls = [tuple(float(v) for v in tr.iteration(i)) for i in range(5)]
        losses.append(ls)
        outs.append(torch.cat([v.float().flatten().cpu() for _, v in sorted(tr.network.state_dict().items())]))
    for (h0, l0), (h1, l1) in zip(*losses):
        assert abs(h0 - h1) <= 1e-3 * abs(h0) and abs(l0 - l1) <= 1e-3 * abs(l0), (losses[0], losses[1])
    cw = cosine(outs[0], outs[1])
    rel = (outs[0] - outs[1]).norm().item() / outs[0].norm().item()
    print(f"image-slider engine vs eager after 5 iterations: weight cos {cw:.7f} rel diff {rel:.2e}")
    assert cw >= 0.999 and rel <= 1e-2
