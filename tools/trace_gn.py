"""Phase timeline of the single-launch GroupNorm (gn_group_kernel): globaltimer stamps per CTA (slider_debug_gn_trace).
Dev tool: prints, per shape, the kernel span and the median / max duration of every phase over the CTAs."""
import ctypes
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from sliders_conceptmod_b200 import _lib, ops

h = _lib.lib()
h.slider_debug_gn_trace.argtypes = [ctypes.c_void_p]
h.slider_debug_gn_group.argtypes = [ctypes.c_int, ctypes.c_int]
NAMES = ["params", "pdl_wait", "load+stats", "reduce", "cluster/sync", "dsmem+stat", "apply+store", "cluster_wait"]
kb = int(sys.argv[1]) if len(sys.argv) > 1 else 24
h.slider_debug_gn_group(1, kb)
for shape in [(2, 64, 64, 320), (2, 64, 64, 960), (2, 32, 32, 640), (2, 16, 16, 1280), (2, 8, 8, 2560)]:
    B, H, W, C = shape
    xs = [torch.randn(B, H, W, C, device="cuda").bfloat16() for _ in range(3)]
    w = torch.ones(C, device="cuda").bfloat16(); b = torch.zeros(C, device="cuda").bfloat16()
    buf = torch.zeros(4096 * 8, dtype=torch.int64, device="cuda")
    for x in xs:
        ops.groupnorm_nhwc(x, 32, w, b, 1e-5, True)
    torch.cuda.synchronize()
    h.slider_debug_gn_trace(buf.data_ptr())
    ops.groupnorm_nhwc(xs[0], 32, w, b, 1e-5, True)
    torch.cuda.synchronize()
    h.slider_debug_gn_trace(None)
    t = buf.view(-1, 8).cpu()
    t = t[t[:, 0] > 0].double()
    if t.numel() == 0:
        print(shape, "takes the two-kernel form (no single-launch plan): no trace")
        continue
    span = (t[:, 7].max() - t[:, 0].min()).item() / 1e3
    start_spread = (t[:, 0].max() - t[:, 0].min()).item() / 1e3
    d = (t[:, 1:] - t[:, :-1]) / 1e3
    print(shape, f"kb {kb} ctas {t.shape[0]} span {span:.2f} us, start spread {start_spread:.2f} us")
    print("   " + "  ".join(f"{n} {d[:, i].median().item():.2f}/{d[:, i].max().item():.2f}" for i, n in enumerate(NAMES[1:])))
