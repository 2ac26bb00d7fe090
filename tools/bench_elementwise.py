"""HBM-roofline sweep of the bandwidth-bound kernels of the slider path:
guidance loss (+grad), CFG + scheduler step, flat AdamW, clip coefficient, and the low-rank backward
(rowdot + colreduce).  Prints one JSON line per (kernel, size): algorithmic bytes / CUDA-event time vs
the measured HBM peak (MEASURED_PEAKS.json) and the nominal 8 TB/s.  Native sizes (16 384 / 65 536
elements) are reported as latency.  Inputs rotate through > 126 MB of distinct buffers so no launch
re-reads L2-resident data.
"""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from sliders_conceptmod_b200 import ops  # noqa: E402

PEAK = 6545.9
p = ROOT / "MEASURED_PEAKS.json"
if p.exists():
    PEAK = float(json.loads(p.read_text())["hbm_gbs"])


def time_rotating(make_call, n_sets, iters=40):
    """make_call(i) launches the kernel on buffer set i % n_sets.  Returns seconds per launch."""
    for i in range(min(n_sets, 4) + 2):
        make_call(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        make_call(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / iters


def sets_for(bytes_per_set):
    return max(2, min(64, int(400e6 // max(bytes_per_set, 1)) + 1))


def report(kernel, dtype, n, bytes_alg, sec):
    gbs = bytes_alg / sec / 1e9
    print(json.dumps({"kernel": kernel, "dtype": str(dtype).replace("torch.", ""), "n_elem": n,
                      "us": round(sec * 1e6, 2), "algorithmic_bytes": bytes_alg, "GBps": round(gbs, 1),
                      "frac_of_measured_hbm": round(gbs / PEAK, 3), "frac_of_8TBps": round(gbs / 8000.0, 3)}),
          flush=True)


def main():
    dev = "cuda"
    sizes = [1 << 14, 1 << 16, 1 << 20, 1 << 22, 1 << 24, 1 << 26, 1 << 27]
    for dtype in (torch.bfloat16, torch.float32):
        es = 2 if dtype == torch.bfloat16 else 4
        for n in sizes:
            ns = sets_for(5 * n * es)
            bufs = [[torch.randn(n, device=dev).to(dtype) for _ in range(4)] for _ in range(ns)]
            sec = time_rotating(lambda i: ops.guidance_loss(bufs[i % ns][0], bufs[i % ns][1], bufs[i % ns][2],
                                                            bufs[i % ns][3], 4.0, 1), ns)
            report("guidance_loss_fwd_bwd", dtype, n, 5 * n * es + 4, sec)
            outs = [torch.empty(n, device=dev, dtype=dtype) for _ in range(ns)]
            sec = time_rotating(lambda i: ops.cfg_step(bufs[i % ns][0], bufs[i % ns][1], bufs[i % ns][2], None,
                                                       guidance=3.0, c_x=1.01, c_eps=-0.02, out=outs[i % ns]), ns)
            report("cfg_ddim_step", dtype, n, 4 * n * es, sec)
            sec = time_rotating(lambda i: ops.cfg_step(bufs[i % ns][0], bufs[i % ns][1], bufs[i % ns][2],
                                                       bufs[i % ns][3], guidance=3.0, c_x=1.01, c_eps=-0.02,
                                                       c_noise=0.1, out=outs[i % ns]), ns)
            report("cfg_ddpm_step_noise", dtype, n, 5 * n * es, sec)
            del bufs, outs
    # flat AdamW at the real LoRA sizes and a large buffer
    for n in (2906880, 4320000, 1 << 26):
        for dtype in (torch.bfloat16, torch.float32):
            es = 2 if dtype == torch.bfloat16 else 4
            ns = sets_for((2 * es + 20) * n)
            P = [torch.randn(n, device=dev).to(dtype) for _ in range(ns)]
            G = [torch.randn(n, device=dev) for _ in range(ns)]
            Mo = [torch.zeros(n, device=dev) for _ in range(ns)]
            Vo = [torch.zeros(n, device=dev) for _ in range(ns)]
            sec = time_rotating(lambda i: ops.adamw_flat(P[i % ns], G[i % ns], Mo[i % ns], Vo[i % ns], 1e-4, 0.9,
                                                         0.999, 1e-8, 0.01, 3), ns)
            report("adamw_flat", dtype, n, (2 * es + 20) * n, sec)
            sec = time_rotating(lambda i: ops.clip_coef(G[i % ns], 0.2), ns)
            report("clip_coef", torch.float32, n, 4 * n, sec)
            del P, G, Mo, Vo
    # low-rank backward at the SD-1.4 / SD-XL layer shapes: bytes = gy read twice + x + h + dh
    r = 4
    for (B, H, W, C, N, k) in [(1, 1, 8192, 320, 320, 1), (1, 1, 2048, 1280, 1280, 1), (1, 1, 65536, 640, 640, 1),
                               (2, 64, 64, 320, 320, 3), (2, 16, 16, 1280, 1280, 3), (16, 64, 64, 640, 640, 3)]:
        dtype = torch.bfloat16
        M = B * H * W
        ns = sets_for((M * C + M * N) * 2)
        shape = (M, C) if k == 1 and B == 1 and H == 1 else (B, H, W, C)
        X = [torch.randn(*shape, device=dev).to(dtype) for _ in range(ns)]
        GY = [torch.randn(*shape[:-1], N, device=dev).to(dtype) for _ in range(ns)]
        Hh = [torch.randn(M, r, device=dev) for _ in range(ns)]
        b = (torch.randn(N, r, device=dev) * 0.02).to(dtype)
        dA = torch.zeros(r, k * k, C, device=dev)
        dB = torch.zeros(N, r, device=dev)
        # captured in a CUDA graph (as the engine runs it): host-side descriptor encoding is not in the timing
        ns = min(ns, 8)
        for i in range(ns):
            ops.lora_grad(GY[i], X[i], Hh[i], b, 0.25, k, dA, dB)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            for i in range(ns):
                ops.lora_grad(GY[i], X[i], Hh[i], b, 0.25, k, dA, dB)
        sec = time_rotating(lambda i: graph.replay(), 1, iters=10) / ns
        bytes_alg = (2 * M * N + M * C) * 2 + M * (r * 4 + 32)
        print(json.dumps({"kernel": "lora_grad(pack_bt + dh GEMM + colreduce dB + colreduce dA, 2 memsets)", "shape": [B, H, W, C, N, k],
                          "us": round(sec * 1e6, 2), "algorithmic_bytes": bytes_alg,
                          "GBps": round(bytes_alg / sec / 1e9, 1),
                          "frac_of_measured_hbm": round(bytes_alg / sec / 1e9 / PEAK, 3)}), flush=True)
        del X, GY, Hh


if __name__ == "__main__":
    main()
