"""Shared builders for parity tests: the same seeded tiny UNet + LoRANetwork twice — one copy on CPU wired
to the oracle arithmetic, one on the GPU running the product kernels."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

from oracle import slider_oracle  # noqa: E402
from sliders_conceptmod_b200 import workload  # noqa: E402
from sliders_conceptmod_b200.lora import LoRANetwork  # noqa: E402
from sliders_conceptmod_b200.unet import UNet2DConditionModel, UNetConfig  # noqa: E402


def build_pair(cfg: UNetConfig, device, dtype, rank=4, alpha=1.0, c3lier=True, method="noxattn", b_std=0.05, seed=0):
    """Returns (unet_cpu, net_cpu [oracle arithmetic, fp32], unet_gpu, net_gpu [product kernels, dtype])."""
    torch.manual_seed(seed)
    unet_cpu = UNet2DConditionModel(cfg).eval().requires_grad_(False)
    net_cpu = LoRANetwork(unet_cpu, rank=rank, alpha=alpha, target_replace=workload.target_modules(c3lier),
                          train_method=method)
    g = torch.Generator().manual_seed(seed + 1)
    for m in net_cpu.unet_loras:
        m.lora_up.weight.data.copy_(torch.randn(m.lora_up.weight.shape, generator=g) * b_std)
    unet_gpu = UNet2DConditionModel(cfg).eval().requires_grad_(False)
    unet_gpu.load_state_dict(unet_cpu.state_dict())
    unet_gpu.to(device, dtype)
    net_gpu = LoRANetwork(unet_gpu, rank=rank, alpha=alpha, target_replace=workload.target_modules(c3lier),
                          train_method=method)
    net_gpu.load_state_dict(net_cpu.state_dict())
    net_gpu.to(device, dtype)
    if dtype != torch.float32:      # the oracle sees exactly the rounded parameter values
        unet_cpu.load_state_dict({k: v.to(dtype).float() for k, v in unet_cpu.state_dict().items()})
        net_cpu.load_state_dict({k: v.to(dtype).float() for k, v in net_cpu.state_dict().items()})
    slider_oracle.attach(net_cpu)
    return unet_cpu, net_cpu, unet_gpu, net_gpu


def cosine(a, b):
    a, b = a.double().flatten().cpu(), b.double().flatten().cpu()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))
