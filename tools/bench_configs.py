"""Secondary workloads of BASELINE.json (configs 3-5) on one GPU, CUDA-event timed, one JSON line each:
  config 3  SD-XL text slider step (attention-only LoRA, euler_a, AdamW + clip 0.2), timesteps_to = 6 of 12
  config 4  image-pair slider iteration (SD-1.4 stand-in, 256 px latents 32x32, two-sided accumulated MSE)
  config 5  inference sweep (SD-XL stand-in, batch 8 -> UNet batch 16, 30 DDIM steps, CFG 5, 7 scales, start_noise 750)
Not the headline bench (bench.py); numbers go to profiles/."""
import json
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import yaml  # noqa: E402

from sliders_conceptmod_b200 import generate_images, model_util, workload  # noqa: E402
from sliders_conceptmod_b200.lora import LoRANetwork  # noqa: E402


def timed(fn, steps, warmup=2):
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(warmup + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / steps


def config3(steps=4):
    wl = workload.WORKLOADS["sdxl_text_slider_rank4_1024px"]
    engine, unet, network, pairs = workload.build(wl, "cuda", torch.bfloat16, use_graphs=True)
    workload.load_pair(engine, wl, pairs[0])
    lat = torch.randn(steps + 4, wl.batch_size, 4, wl.resolution // 8, wl.resolution // 8, device="cuda")
    sec = timed(lambda i: engine.train_step(lat[i], 6), steps)
    print(json.dumps({"config": 3, "workload": wl.name, "metric": "slider_train_steps_per_sec", "value": round(1 / sec, 3),
                      "ms_per_step": round(sec * 1e3, 2), "timesteps_to": 6, "lora_modules": len(network.unet_loras),
                      "lora_params": int(network.flat_param.numel()), "loss": float(engine.loss)}), flush=True)
    del engine, unet, network
    torch.cuda.empty_cache()


def config4(steps=6):
    from sliders_conceptmod_b200 import config_util, prompt_util
    from sliders_conceptmod_b200.train_lora_scale import ImageSliderTrainer, PairedLatents
    with tempfile.TemporaryDirectory() as tmp:
        prompts = [{"target": "person", "positive": "person, smiling", "unconditional": "", "neutral": "person",
                    "action": "enhance", "guidance_scale": 4, "resolution": 256, "batch_size": 1}]
        Path(tmp, "p.yaml").write_text(yaml.safe_dump(prompts))
        cfg = {"prompts_file": str(Path(tmp, "p.yaml")), "pretrained_model": {"name_or_path": "sd14-standin"},
               "network": {"type": "c3lier", "rank": 4, "alpha": 1.0, "training_method": "noxattn"},
               "train": {"precision": "bfloat16", "noise_scheduler": "ddim", "iterations": steps, "lr": 2e-4,
                         "optimizer": "AdamW", "lr_scheduler": "constant", "max_denoising_steps": 50},
               "save": {"name": "t", "path": tmp, "per_steps": 1000, "precision": "bfloat16"},
               "logging": {"use_wandb": False, "verbose": False}, "other": {"use_xformers": False}}
        Path(tmp, "c.yaml").write_text(yaml.safe_dump(cfg))
        config = config_util.load_config_from_yaml(str(Path(tmp, "c.yaml")))
        ps = prompt_util.load_prompts_from_yaml(config.prompts_file, [])
        data = PairedLatents(scales=[-1, 1], synthetic=8, latent_hw=32, seed=1)
        for engine in (False, True):
            data = PairedLatents(scales=[-1, 1], synthetic=8, latent_hw=32, seed=1)
            tr = ImageSliderTrainer(config, ps, "cuda", data, seed=1, engine=engine)
            sec = timed(lambda i: tr.iteration(i), steps, warmup=3)
            print(json.dumps({"config": 4, "workload": "sd14_image_slider_256px", "metric": "image_slider_iterations_per_sec",
                              "value": round(1 / sec, 3), "ms_per_iteration": round(sec * 1e3, 2),
                              "note": ("--engine: two graph-replayed grad forward/backward sides + flat AdamW" if engine else
                                       "eager drop-in loop (2 grad forwards + 2 backwards + torch AdamW per iteration)")}), flush=True)
            del tr
    torch.cuda.empty_cache()


def config5(n_scales=7, batch=8, steps=30):
    _, encs, unet, sched = model_util.load_models_xl("sdxl-standin", scheduler_name="ddim")
    unet.to("cuda", torch.bfloat16).requires_grad_(False).eval()
    net = LoRANetwork(unet, rank=4, multiplier=1.0, alpha=1.0, target_replace=workload.target_modules(False),
                      train_method="noxattn").to("cuda", torch.bfloat16)
    g = torch.Generator().manual_seed(0)
    for m in net.unet_loras:
        m.lora_up.weight.data.copy_((torch.randn(m.lora_up.weight.shape, generator=g) * 0.02).to(torch.bfloat16))
    net.flatten()
    enc = encs[0]
    prompts = [f"image of a person {i}" for i in range(batch)]
    pe, ue = enc.encode(prompts).cuda().bfloat16(), enc.encode([""] * batch).cuda().bfloat16()
    from sliders_conceptmod_b200.train_util import get_add_time_ids
    ids = get_add_time_ids(1024, 1024).cuda().bfloat16().repeat(2 * batch, 1)
    pooled = torch.cat([enc.encode_pooled([""] * batch), enc.encode_pooled(prompts)]).cuda().bfloat16()
    added = {"text_embeds": pooled, "time_ids": ids}
    scales = [-3.0, -2.0, -1.0, 0.0, 1.0, 2.0, 3.0][:n_scales]
    run = lambda: generate_images.sweep(unet, net, sched, pe, ue, scales, seed=0, height=1024, width=1024, start_noise=750,
                                        guidance_scale=5.0, num_inference_steps=steps, added_cond=added)
    generate_images.sweep(unet, net, sched, pe, ue, scales[:1], seed=0, height=1024, width=1024, start_noise=750,
                          guidance_scale=5.0, num_inference_steps=3, added_cond=added)      # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = run()
    e1.record()
    torch.cuda.synchronize()
    sec = e0.elapsed_time(e1) * 1e-3
    print(json.dumps({"config": 5, "workload": "sdxl_inference_sweep_1024px", "metric": "unet_steps_per_sec",
                      "value": round(len(scales) * steps / sec, 2), "latents_per_sec": round(len(scales) * batch / sec, 3),
                      "unet_batch": 2 * batch, "ddim_steps": steps, "scales": scales, "seconds": round(sec, 2),
                      "host_seconds": round(time.perf_counter() - t0, 2),
                      "finite": all(bool(torch.isfinite(v).all()) for v in out.values())}), flush=True)


if __name__ == "__main__":
    which = [int(a) for a in sys.argv[1:]] or [3, 4, 5]
    if 3 in which:
        config3()
    if 4 in which:
        config4()
    if 5 in which:
        config5()
