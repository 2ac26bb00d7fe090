"""GPU idle gaps inside real train steps (torch profiler timeline): total busy time vs step time and the largest gaps
between consecutive kernels, with the kernel that follows each gap."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from torch.profiler import profile, ProfilerActivity
from sliders_conceptmod_b200 import workload

wl = workload.WORKLOADS["sd14_text_slider_rank4_512px"]
engine, unet, network, pairs = workload.build(wl, "cuda", torch.bfloat16, use_graphs=True, lora_b_std=0.02)
workload.load_pair(engine, wl, pairs[0])
lat = torch.randn(wl.batch_size, 4, wl.resolution // 8, wl.resolution // 8).pin_memory()
for _ in range(3):
    engine.train_step(lat, 25)
torch.cuda.synchronize()
steps = 3
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    engine.train_step(lat, 25)
e1.record()
torch.cuda.synchronize()
print("step %.3f ms (events, no profiler)" % (e0.elapsed_time(e1) / steps))
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(steps):
        engine.train_step(lat, 25)
    torch.cuda.synchronize()
evs = sorted(((e.time_range.start, e.time_range.end, e.name) for e in prof.events()
              if e.device_type == torch.autograd.DeviceType.CUDA), key=lambda t: t[0])
busy_end = evs[0][0]
busy = 0.0
gaps = []
for s, e, n in evs:
    if s > busy_end:
        gaps.append((s - busy_end, n))
        busy += 0
    busy += max(0.0, e - max(s, busy_end))
    busy_end = max(busy_end, e)
span = evs[-1][1] - evs[0][0]
print("profiled span %.3f ms per step, GPU busy %.3f ms per step, idle %.3f ms per step in %d gaps" %
      (span / steps / 1e3, busy / steps / 1e3, (span - busy) / steps / 1e3, len(gaps) // steps))
big = sorted(gaps, reverse=True)[:12]
for g, n in big:
    print("  gap %8.1f us before %s" % (g, n[:90]))
import collections
c = collections.Counter()
for g, n in gaps:
    c[n[:60]] += g
print("idle time by the kernel that follows the gap (us per step):")
for n, g in c.most_common(8):
    print("  %8.1f  %s" % (g / steps, n))
