"""Per-signature breakdown of the wrapped-layer launches of ONE adapter-on CFG UNet forward (the set bench.py's
roofline replays): every distinct (kind, shape) is captured as a CUDA graph of `--reps` back-to-back launches and
timed with CUDA events, so the per-launch time is the in-graph one (launch gap + prologue + main loop + epilogue
[+ split-K finalize]).  Dev tool: prints one line per signature sorted by its share of the forward."""
import argparse
import json
import sys
from collections import OrderedDict
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from sliders_conceptmod_b200 import ops, workload


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="sd14_text_slider_rank4_512px")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--json", default=None)
    ap.add_argument("--cold", action="store_true", help="rotate over >= 300 MB of weights per signature: every launch streams its weights from HBM, as in the real forward")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    wl = workload.WORKLOADS[args.workload]
    engine, unet, network, pairs = workload.build(wl, dev, torch.bfloat16, seed=0, use_graphs=False, lora_b_std=0.0)
    workload.load_pair(engine, wl, pairs[0])
    calls = []
    real, real_grouped = ops.lora_forward, ops.lora_forward_grouped

    def spy(x, w, bias, a, b, s, ksize=1, want_h=False, residual=None, stride=1):
        calls.append(("single", tuple(x.shape), tuple(w.shape), bias is not None, None if a is None else a.shape[0], s,
                      ksize * 10 + stride, residual is not None))
        return real(x, w, bias, a, b, s, ksize=ksize, want_h=want_h, residual=residual, stride=stride)

    def spy_grouped(x, w_cat, a_cat, b_cat, groups, s):
        calls.append(("grouped", tuple(x.shape), tuple(w_cat.shape), False, a_cat.shape[0], s, groups, False))
        return real_grouped(x, w_cat, a_cat, b_cat, groups, s)

    real_multi = ops.lora_forward_small_multi
    multi = []

    def spy_multi(x, plan, scales):
        multi.append((x.clone(), plan, tuple(scales)))
        return real_multi(x, plan, scales)

    ops.lora_forward, ops.lora_forward_grouped, ops.lora_forward_small_multi = spy, spy_grouped, spy_multi
    try:
        engine._body_denoise()
    finally:
        ops.lora_forward, ops.lora_forward_grouped, ops.lora_forward_small_multi = real, real_grouped, real_multi
    torch.cuda.synchronize()
    sigs = OrderedDict()
    for c in calls:
        sigs[c] = sigs.get(c, 0) + 1
    dt = torch.bfloat16
    rows = []
    for (x, plan, scales) in multi:
        def run_m():
            for _ in range(args.reps):
                real_multi(x, plan, scales)
        run_m(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            run_m()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (3 * args.reps)
        M, K = x.shape
        fl = sum(2.0 * M * n * K + 2.0 * M * plan.rank * (K + n) for n in plan.n)
        rows.append({"kind": "multi", "x": list(x.shape), "w": [sum(plan.n), 1, K], "k": 1, "res": False,
                     "rank": plan.rank, "count": 1, "us": round(us, 2), "total_us": round(us, 1),
                     "gflop": round(fl / 1e9, 2), "tflops": round(fl / us / 1e6, 1)})
    for (kind, xs, ws, has_b, r, s, k, has_res), count in sigs.items():
        stride = 1
        if kind == "single":
            k, stride = k // 10, k % 10
        M = 1
        for i, d in enumerate(xs[:-1]):
            M *= ((d - 1) // 2 + 1) if (stride == 2 and i in (1, 2)) else d
        N, T, C = ws
        # several operand sets so that back-to-back launches do not re-hit the same weights in L2
        nset = max(1, min(8, int(256e6 / max(1, N * T * C * 2))))
        if args.cold:
            nset = max(2, min(args.reps, -(-int(300e6) // (N * T * C * 2))))
        sets = []
        for _ in range(nset):
            x = torch.randn(xs, device=dev, dtype=dt)
            w = torch.randn(ws, device=dev, dtype=dt) * 0.02
            if kind == "grouped":
                a = torch.randn(r, 1, C, device=dev, dtype=dt) * 0.02
                b = torch.randn(N, r // k, device=dev, dtype=dt) * 0.02
                sets.append((x, w, a, b))
            else:
                a = torch.randn(r, T, C, device=dev, dtype=dt) * 0.02 if r else None
                b = torch.randn(N, r, device=dev, dtype=dt) * 0.02 if r else None
                bias = torch.zeros(N, device=dev, dtype=dt) if has_b else None
                res = torch.zeros(*xs[:-1], N, device=dev, dtype=dt) if has_res else None     # stride-2 layers never fuse a residual
                sets.append((x, w, bias, a, b, res))
        if kind == "grouped":
            flops = 2.0 * M * N * C + 2.0 * M * r * C + 2.0 * M * (r // k) * N
        else:
            flops = 2.0 * M * N * T * C + (2.0 * M * r * (T * C + N) if r else 0.0)

        def run():
            for i in range(args.reps):
                o = sets[i % nset]
                if kind == "grouped":
                    real_grouped(o[0], o[1], o[2], o[3], k, s)
                else:
                    real(o[0], o[1], o[2], o[3], o[4], s, ksize=k, residual=o[5], stride=stride)
        run(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            run()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (3 * args.reps)
        rows.append({"kind": kind, "x": list(xs), "w": list(ws), "k": k, "stride": stride, "res": has_res, "rank": r, "count": count,
                     "us": round(us, 2), "total_us": round(us * count, 1), "gflop": round(flops / 1e9, 2),
                     "tflops": round(flops / us / 1e6, 1)})
    tot = sum(r["total_us"] for r in rows)
    totf = sum(r["gflop"] * r["count"] for r in rows)
    rows.sort(key=lambda r: -r["total_us"])
    for r in rows:
        r["share"] = round(r["total_us"] / tot, 3)
        print(json.dumps(r))
    print(json.dumps({"total_us": round(tot, 1), "total_gflop": round(totf, 1), "tflops": round(totf / tot * 1e-3, 1),
                      "launches": len(calls) + len(multi)}))
    if args.json:
        Path(args.json).write_text(json.dumps(rows, indent=1))


if __name__ == "__main__":
    main()
