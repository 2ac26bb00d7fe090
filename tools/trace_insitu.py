"""Per-CTA clock64() timeline of every lora_gemm launch INSIDE one eager adapter-on CFG UNet forward (cold weights,
instruction cache shared with the other kernels of the forward) next to the same call repeated back to back (warm).
Dev tool."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops, workload, _lib

wl = workload.WORKLOADS["sd14_text_slider_rank4_512px"]
engine, unet, network, pairs = workload.build(wl, torch.device("cuda", 0), torch.bfloat16, seed=0, use_graphs=False,
                                              lora_b_std=0.0)
workload.load_pair(engine, wl, pairs[0])
engine._body_denoise()
torch.cuda.synchronize()
trace = torch.zeros(148 * 16, dtype=torch.int64, device="cuda")
real = ops.lora_forward
rows = []


def stamp():
    t = trace.view(148, 16).cpu()
    t = t[t[:, 0] > 0]
    d = (t[:, 1:8] - t[:, :1]).double()
    return [int(d[:, i].median().item()) for i in range(7)], t.shape[0]


def spy(x, w, bias, a, b, s, ksize=1, want_h=False, residual=None, stride=1):
    torch.cuda.synchronize()
    trace.zero_()
    _lib.lib().slider_debug_gemm_trace(trace.data_ptr())
    out = real(x, w, bias, a, b, s, ksize=ksize, want_h=want_h, residual=None if residual is None else residual, stride=stride)
    torch.cuda.synchronize()
    cold, n = stamp()
    res2 = None if residual is None else residual.clone()
    for _ in range(2):
        trace.zero_()
        real(x, w, bias, a, b, s, ksize=ksize, want_h=want_h, residual=res2, stride=stride)
        torch.cuda.synchronize()
    warm, _ = stamp()
    _lib.lib().slider_debug_gemm_trace(None)
    if n:
        rows.append((tuple(x.shape), tuple(w.shape), residual is not None, n, cold, warm))
    return out


ops.lora_forward = spy
try:
    engine._body_denoise()
finally:
    ops.lora_forward = real
torch.cuda.synchronize()
names = ["setup", "tma1", "full1", "mma_end", "acc", "epi", "exit"]
print("x w res ctas | in-situ: " + " ".join(names) + " | warm repeat: " + " ".join(names))
tot_c = tot_w = 0
for (xs, ws, res, n, cold, warm) in rows:
    tot_c += cold[6]; tot_w += warm[6]
    print(xs, ws, int(res), n, "|", " ".join(f"{v:6d}" for v in cold), "|", " ".join(f"{v:6d}" for v in warm))
print(f"sum of exit cycles: in-situ {tot_c}  warm {tot_w}  ({len(rows)} tensor-core launches)")
