"""GroupNorm forms at the shapes whose slab needs one CTA per SM (> 100 KB per CTA at k = 8): us per GroupNorm in a graph."""
import ctypes, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sliders_conceptmod_b200 import ops, _lib
h = _lib.lib()
h.slider_debug_gn_group.argtypes = [ctypes.c_int, ctypes.c_int]
def bench(shape):
    B, H, W, C = shape
    xs = [torch.randn(B, H, W, C, device="cuda").bfloat16() for _ in range(2)]
    w = torch.ones(C, device="cuda").bfloat16(); b = torch.zeros(C, device="cuda").bfloat16()
    for x in xs: ops.groupnorm_nhwc(x, 32, w, b, 1e-5, True)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(3):
            for x in xs: ops.groupnorm_nhwc(x, 32, w, b, 1e-5, True)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): g.replay()
    e1.record(); torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) * 1e3 / 30, 2)
for shape in [(2, 64, 64, 960), (6, 64, 64, 960), (2, 128, 128, 320), (16, 128, 128, 320), (2, 128, 128, 640), (16, 64, 64, 640)]:
    res = {}
    for name, on in (("two_kernel", 0), ("group", 1)):
        h.slider_debug_gn_group(on, 24)
        res[name] = bench(shape)
    h.slider_debug_gn_group(1, 24)
    print(shape, res, flush=True)
