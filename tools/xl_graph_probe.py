import sys, time, torch
sys.path.insert(0, "/root/repo")
from sliders_conceptmod_b200 import model_util, workload
from sliders_conceptmod_b200.lora import LoRANetwork
from sliders_conceptmod_b200.train_util import get_add_time_ids
batch = 8
_, encs, unet, sched = model_util.load_models_xl("sdxl-standin", scheduler_name="ddim")
unet.to("cuda", torch.bfloat16).requires_grad_(False).eval()
net = LoRANetwork(unet, rank=4, multiplier=1.0, alpha=1.0, target_replace=workload.target_modules(False), train_method="noxattn").to("cuda", torch.bfloat16)
net.flatten()
enc = encs[0]
emb = torch.cat([enc.encode([""] * batch), enc.encode(["p"] * batch)]).cuda().bfloat16()
ids = get_add_time_ids(1024, 1024).cuda().bfloat16().repeat(2 * batch, 1)
pooled = torch.cat([enc.encode_pooled([""] * batch), enc.encode_pooled(["p"] * batch)]).cuda().bfloat16()
added = {"text_embeds": pooled, "time_ids": ids}
x = torch.randn(2 * batch, 4, 128, 128, device="cuda", dtype=torch.bfloat16)
t = torch.tensor([500.0], device="cuda")
net.set_lora_slider(2.0)
def fwd():
    with torch.no_grad(), net:
        return unet(x, t, encoder_hidden_states=emb, added_cond_kwargs=added).sample
for _ in range(2): fwd()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): fwd()
e1.record(); torch.cuda.synchronize()
print("eager ms/forward", e0.elapsed_time(e1) / 3)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    y = fwd()
g.replay(); torch.cuda.synchronize()
e0.record()
for _ in range(3): g.replay()
e1.record(); torch.cuda.synchronize()
print("graph ms/forward", e0.elapsed_time(e1) / 3)
import collections, re
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    g.replay(); torch.cuda.synchronize()
tot, cnt = collections.Counter(), collections.Counter()
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        n = re.sub(r"^void ", "", ev.name)[:100]
        tot[n] += ev.device_time_total; cnt[n] += 1
T = sum(tot.values())
print("sum kernel ms", T / 1e3)
for n, v in tot.most_common(16):
    print(f"{v/1e3:8.2f} ms {100*v/T:5.1f}% x{cnt[n]:4d} {n}")
