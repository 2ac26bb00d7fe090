"""Dev: how far is a bf16 run of the ORACLE (reference arithmetic on CPU) from the fp32 oracle, and the CUDA path from both?"""
import sys, copy
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent)); sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
import torch
from helpers import build_pair
from oracle import slider_oracle
from sliders_conceptmod_b200 import model_util
from sliders_conceptmod_b200.engine import EngineConfig, SliderEngine
from sliders_conceptmod_b200.unet import UNetConfig

cfg = UNetConfig.tiny()
dtype = torch.bfloat16
for emb_scale in (1.0, 4.0):
    unet_c, net_c, unet_g, net_g = build_pair(cfg, "cuda", dtype)
    g = torch.Generator().manual_seed(11)
    emb = {k: (torch.randn(1, 77, cfg.cross_attention_dim, generator=g) * emb_scale).to(dtype).float() for k in ("target", "positive", "neutral", "unconditional")}
    lat = torch.randn(1, 4, 16, 16, generator=g).to(dtype).float()
    loss32, i32 = slider_oracle.train_iteration(unet_c, net_c, slider_oracle.OracleScheduler("ddim"), emb, lat, 3, 4.0, "enhance", max_denoising_steps=10)
    # oracle in bf16 on CPU
    unet_b, net_b, _, _ = build_pair(cfg, "cpu", torch.float32)
    unet_b.to(torch.bfloat16); net_b.to(torch.bfloat16)
    embb = {k: v.to(dtype) for k, v in emb.items()}
    loss16, i16 = slider_oracle.train_iteration(unet_b, net_b, slider_oracle.OracleScheduler("ddim"), embb, lat.to(dtype), 3, 4.0, "enhance", max_denoising_steps=10)
    eng = SliderEngine(unet_g, net_g, model_util.create_noise_scheduler("ddim"), EngineConfig(max_denoising_steps=10, use_graphs=False), 1, 128, 128, (77, cfg.cross_attention_dim), dtype, "cuda")
    eng.load_pair(emb["target"], emb["positive"], emb["neutral"], emb["unconditional"], 4.0, "enhance")
    lg = eng.train_step(lat, 3).item()
    print(f"emb_scale {emb_scale}: loss fp32-oracle {loss32.item():.5e}  bf16-oracle {loss16.item():.5e}  cuda-bf16 {lg:.5e}")
    for k in ("positive", "neutral", "unconditional"):
        ref = i32[k]; sc = ref.abs().max().item()
        print(f"   {k}: scale {sc:.3f} bf16-oracle err {(i16[k].float()-ref).abs().max().item():.3e} cuda err {(getattr(eng, 'uncond' if k=='unconditional' else k).float().cpu()-ref).abs().max().item():.3e}")
    print("   (pos-unc) magnitude", (i32['positive']-i32['unconditional']).abs().max().item())
