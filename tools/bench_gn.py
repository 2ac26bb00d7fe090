import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch, ctypes
from sliders_conceptmod_b200 import ops, _lib
h = _lib.lib()
h.slider_debug_gn_variant.argtypes = [ctypes.c_int]
def bench(shape, variant):
    h.slider_debug_gn_variant(variant)
    B, H, W, C = shape
    xs = [torch.randn(B, H, W, C, device="cuda").bfloat16() for _ in range(4)]
    w = torch.ones(C, device="cuda").bfloat16(); b = torch.zeros(C, device="cuda").bfloat16()
    for x in xs: ops.groupnorm_nhwc(x, 32, w, b, 1e-5, True)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(5):
            for x in xs: ops.groupnorm_nhwc(x, 32, w, b, 1e-5, True)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (5 * 20)
for shape in [(2, 64, 64, 320), (2, 32, 32, 640), (2, 16, 16, 1280), (2, 8, 8, 2560), (6, 64, 64, 320)]:
    print(shape, {v: round(bench(shape, v), 2) for v in (0, 1, 2, 3, 4)}, "us per GN (stats+apply)")
