"""Library SDPA forward + backward at the training shapes of the SD-1.4 UNet (batch 2, 8 heads), per backend.
Attention is not a libslider kernel; this only picks which library kernel the grad-recording forward asks for."""
import torch
import torch.nn.functional as F
from torch.nn.attention import SDPBackend, sdpa_kernel

dev, dt = "cuda", torch.bfloat16


def timed(fn, reps=10):
    """Sum of the CUDA kernel durations per call (torch profiler): the eager call itself is host-launch bound."""
    from torch.profiler import profile, ProfilerActivity
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
    tot = 0.0
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            tot += ev.device_time_total if hasattr(ev, "device_time_total") else ev.cuda_time_total
    return tot / reps


for (L, S, d) in [(4096, 4096, 40), (1024, 1024, 80), (256, 256, 160), (64, 64, 160), (4096, 77, 40), (1024, 77, 80), (256, 77, 160)]:
    q = torch.randn(2, 8, L, d, device=dev, dtype=dt, requires_grad=True)
    k = torch.randn(2, 8, S, d, device=dev, dtype=dt, requires_grad=True)
    v = torch.randn(2, 8, S, d, device=dev, dtype=dt, requires_grad=True)
    go = torch.randn(2, 8, L, d, device=dev, dtype=dt)
    row = []
    for name, be in [("default", None), ("cudnn", SDPBackend.CUDNN_ATTENTION), ("flash", SDPBackend.FLASH_ATTENTION),
                     ("efficient", SDPBackend.EFFICIENT_ATTENTION)]:
        def run():
            if be is None:
                o = F.scaled_dot_product_attention(q, k, v)
            else:
                with sdpa_kernel(be):
                    o = F.scaled_dot_product_attention(q, k, v)
            o.backward(go)
            q.grad = k.grad = v.grad = None
        try:
            row.append(f"{name} {timed(run):7.1f}")
        except RuntimeError as e:
            row.append(f"{name}    n/a")
    print(f"L={L:5d} S={S:5d} d={d:4d}  fwd+bwd us: " + " | ".join(row))
