"""In-situ kernel timing (torch.profiler / CUPTI, warm L2, CUDA-graph replay) of one phase of the slider step.
  python tools/torch_profile.py [denoise|predict|train] [--workload NAME]
Unlike the ncu launch list (cold caches, serialised), these are the durations inside the real replay."""
import sys, collections, re
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from torch.profiler import profile, ProfilerActivity
from sliders_conceptmod_b200 import workload

phase = next((a for a in sys.argv[1:] if a in ("denoise", "predict", "train", "predict_train")), "denoise")
name = sys.argv[sys.argv.index("--workload") + 1] if "--workload" in sys.argv else "sd14_text_slider_rank4_512px"
wl = workload.WORKLOADS[name]
engine, unet, network, pairs = workload.build(wl, "cuda", torch.bfloat16, use_graphs=True, lora_b_std=0.02)
workload.load_pair(engine, wl, pairs[0])
lat = torch.randn(wl.batch_size, 4, wl.resolution // 8, wl.resolution // 8)
for _ in range(2):
    engine.train_step(lat, 2)
torch.cuda.synchronize()
g = engine._graphs[phase]
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
reps = 5
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    g.replay()
e1.record()
torch.cuda.synchronize()
wall = e0.elapsed_time(e1) / reps
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(reps):
        g.replay()
    torch.cuda.synchronize()
tot, cnt = collections.Counter(), collections.Counter()
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        n = re.sub(r"^void ", "", ev.name)[:90]
        tot[n] += ev.device_time_total if hasattr(ev, "device_time_total") else ev.cuda_time_total
        cnt[n] += 1
T = sum(tot.values())
print(f"phase {phase}: graph replay {wall:.3f} ms; sum of kernel durations {T / reps / 1e3:.3f} ms; launches {sum(cnt.values()) // reps}")
for n, v in tot.most_common(30):
    print(f"{v / reps:9.1f} us {100 * v / T:5.1f}% x{cnt[n] // reps:4d} {n}")
