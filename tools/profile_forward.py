"""Profiling target: one adapter-on CFG UNet forward (+ fused CFG/DDIM step) and one grad forward/backward of the
SD-1.4 workload, eager (no CUDA graphs), bracketed by cudaProfilerStart/Stop.
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv \
      python tools/profile_forward.py [--train]
"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import workload

wl = workload.WORKLOADS[sys.argv[sys.argv.index("--workload") + 1] if "--workload" in sys.argv else "sd14_text_slider_rank4_512px"]
engine, unet, network, pairs = workload.build(wl, "cuda", torch.bfloat16, use_graphs=False, lora_b_std=0.02)
workload.load_pair(engine, wl, pairs[0])
engine.lat.normal_()
t_tab, c_tab, s_tab = engine._table(wl.max_denoising_steps)[:3]
engine.t_dev.copy_(t_tab[0:1]); engine.coef.copy_(c_tab[0]); engine.in_scale.copy_(s_tab[0])
for _ in range(2):
    engine._body_denoise()
    if "--train" in sys.argv:
        engine._body_predict(); engine._body_train()
torch.cuda.synchronize()
torch.cuda.profiler.start()
engine._body_denoise()
if "--train" in sys.argv:
    engine._body_predict()
    engine._body_train()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done")
