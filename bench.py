#!/usr/bin/env python
"""Headline benchmark: slider train steps/sec (SD-1.4 rank-4) — BASELINE.json `metric`.

  python bench.py --gpus N --steps K --warmup W              (N>1: launched by torch.distributed.run)
  python bench.py --impl reference --gpus N --steps K --warmup W

One "step" = one reference training iteration (conceptmod/textsliders/train_lora.py:155-309) on the
configuration BASELINE.json quotes the metric on (configs[1]: SD-1.4-shaped random-init UNet, 512 px,
rank 4, alpha 1, noxattn + c3lier, 'male, female' attributes, bf16, DDIM): 25 adapter-on CFG UNet
forwards (the mean of randint(1,50), fixed so every step does the same work), 3 adapter-off
predictions, 1 grad forward, guidance loss, backward into the LoRA factors, AdamW step.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

WORKLOAD = "sd14_text_slider_rank4_512px"
TIMESTEPS_TO = 25
METRIC = "slider_train_steps_per_sec"
UNIT = "steps/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--timesteps-to", type=int, default=TIMESTEPS_TO, help="denoising steps per iteration (headline: 25)")
    ap.add_argument("--no-graphs", action="store_true")
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=240.0)
    ap.add_argument("--skip-eager-baseline", action="store_true")
    ap.add_argument("--profile-step", action="store_true",
                    help="after the timed regions run ONE more step between cudaProfilerStart/Stop, for "
                         "`ncu --profile-from-start off` (the launch list of this same command; tools/gpu_launch_list_bench.sh)")
    return ap.parse_args()


def peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        d["source"] = "measured"
        return d
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "200"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------- roofline of the dominant kernel
def fused_gemm_roofline(engine, pk: dict, reps: int = 5) -> dict:
    """Dominant kernel = lora_gemm_kernel (fused base+LoRA implicit GEMM).  Records the wrapped-layer calls of
    ONE adapter-on CFG UNet forward, then replays exactly those launches back to back on the current stream,
    timed with CUDA events: achieved = sum(algorithmic FLOPs) / sum(launch durations)."""
    from sliders_conceptmod_b200 import ops
    calls = []
    real, real_grouped = ops.lora_forward, ops.lora_forward_grouped

    def spy(x, w, bias, a, b, s, ksize=1, want_h=False, residual=None, stride=1):
        calls.append(("single", tuple(x.shape), tuple(w.shape), bias is not None, None if a is None else a.shape[0], s,
                      (ksize, stride), x.dtype, residual is not None))
        return real(x, w, bias, a, b, s, ksize=ksize, want_h=want_h, residual=residual, stride=stride)

    def spy_grouped(x, w_cat, a_cat, b_cat, groups, s):
        calls.append(("grouped", tuple(x.shape), tuple(w_cat.shape), False, a_cat.shape[0], s, groups, x.dtype, False))
        return real_grouped(x, w_cat, a_cat, b_cat, groups, s)

    real_multi = ops.lora_forward_small_multi

    def spy_multi(x, plan, scales):
        calls.append(("multi", tuple(x.shape), plan, True, plan.rank, tuple(scales), 1, x.dtype, False))
        return real_multi(x, plan, scales)

    ops.lora_forward, ops.lora_forward_grouped, ops.lora_forward_small_multi = spy, spy_grouped, spy_multi
    try:
        engine._body_denoise()
    finally:
        ops.lora_forward, ops.lora_forward_grouped, ops.lora_forward_small_multi = real, real_grouped, real_multi
    torch.cuda.synchronize()
    flops, alg_bytes, args = 0.0, 0.0, []
    for (kind, xs, ws, has_b, r, s, k, dt, has_res) in calls:
        if kind == "multi":         # every time_emb_proj of the forward in one weight-streaming launch
            plan = ws
            es = 2 if dt == torch.bfloat16 else 4
            M, K = xs
            for n in plan.n:
                flops += 2.0 * M * n * K + 2.0 * M * r * (K + n)
                alg_bytes += es * (float(n) * K + M * n)
            alg_bytes += es * float(M) * K
            args.append(("multi", torch.randn(xs, device="cuda", dtype=dt), plan, s))
            continue
        es = 2 if dt == torch.bfloat16 else 4
        stride = 1
        if kind == "single":
            k, stride = k
        M = 1                     # OUTPUT pixels (a stride-2 layer produces a quarter of its input's)
        for i, d in enumerate(xs[:-1]):
            M *= ((d - 1) // 2 + 1) if (stride == 2 and len(xs) == 4 and i in (1, 2)) else d
        alg_bytes += es * (float(torch.Size(xs).numel()) + float(torch.Size(ws).numel()) +
                           float(M) * ws[0] * (2 if has_res else 1))
        N, T, C = ws
        x = torch.randn(xs, device="cuda", dtype=dt)
        w = torch.randn(ws, device="cuda", dtype=dt) * 0.02
        if kind == "grouped":
            groups, rg = k, r // k
            flops += 2.0 * M * N * C + 2.0 * M * r * C + 2.0 * M * rg * N
            a = torch.randn(r, 1, C, device="cuda", dtype=dt) * 0.02
            b = torch.randn(N, rg, device="cuda", dtype=dt) * 0.02
            args.append(("grouped", x, w, a, b, groups, s))
            continue
        flops += 2.0 * M * N * T * C + (2.0 * M * r * (T * C + N) if r else 0.0)
        a = torch.randn(r, T, C, device="cuda", dtype=dt) * 0.02 if r else None
        b = torch.randn(N, r, device="cuda", dtype=dt) * 0.02 if r else None
        bias = torch.zeros(N, device="cuda", dtype=dt) if has_b else None
        res = torch.zeros(*xs[:-1], N, device="cuda", dtype=dt) if has_res else None
        args.append(("single", x, w, bias, a, b, s, k, res, stride))

    def timed(subset):
        def run():
            for c in subset:
                if c[0] == "grouped":
                    real_grouped(*c[1:])
                elif c[0] == "multi":
                    real_multi(c[1], c[2], c[3])
                else:
                    real(c[1], c[2], c[3], c[4], c[5], c[6], ksize=c[7], residual=c[8], stride=c[9])
        # capture the launch sequence in a CUDA graph so host-side descriptor encoding is not in the timing
        run(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            run()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / reps
    sec = timed(args)
    achieved = flops / sec / 1e12
    # the same figure over the launches earlier rounds counted (adapter launches + the plain 1x1 convolutions): the frozen
    # un-wrapped nn.Linear layers (attn2.to_q / to_out, second feed-forward Linear; 2-D activations, no adapter) joined
    # the fused kernel in r02 and are small launches that pull the average down while making the forward faster
    def frozen_linear(c):
        return c[0] == "single" and c[4] is None and c[1].dim() == 2
    sub = [c for c in args if not frozen_linear(c)]
    sub_info = None
    if 0 < len(sub) < len(args):
        fl_lin = sum(2.0 * c[1].shape[0] * c[2].shape[0] * c[2].shape[2] for c in args if frozen_linear(c))
        sec_sub = timed(sub)
        sub_info = {"launches": len(sub), "flops": flops - fl_lin, "achieved": round((flops - fl_lin) / sec_sub / 1e12, 1),
                    "avg_launch_us": round(sec_sub / len(sub) * 1e6, 2)}
    peak = float(pk.get("bf16_tflops_sustained", pk["bf16_tflops"]))
    burst = float(pk["bf16_tflops"])
    traffic, tc_util = measured_traffic()
    return {"bound": "tensor", "kernel": "lora_gemm_kernel", "achieved": round(achieved, 1), "peak": peak,
            "peak_source": pk["source"] + ":bf16_tflops_sustained", "unit": "TFLOP/s", "frac": round(achieved / peak, 4),
            "frac_burst": round(achieved / burst, 4), "peak_burst": burst,
            "tc_util_pct": tc_util, "tc_util_source": "profiles/gemm_traffic.json: duration-weighted ncu "
            "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active over the tensor-core launches of the "
            "same forward (ncu --set full capture, committed; not measured inside this run)",
            "traffic": traffic, "launches": len(calls), "avg_launch_us": round(sec / len(calls) * 1e6, 2),
            "flops_per_forward": flops, "algorithmic_bytes_per_launch": round(alg_bytes / len(calls)),
            "without_frozen_linear": None if sub_info is None else dict(sub_info, frac=round(sub_info["achieved"] / peak, 4))}


def measured_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per lora_gemm_kernel launch, from the committed `ncu --set full`
    capture of the same forward (tools/profile_forward.py -> tools/ncu_traffic.py -> profiles/gemm_traffic.json)."""
    f = ROOT / "profiles" / "gemm_traffic.json"
    if not f.exists():
        return None, None
    try:
        d = json.loads(f.read_text())
        ks = [v for n, v in d["kernels"].items() if "lora_gemm_kernel" in n]
        traffic = round(sum(k["dram_read_bytes"] + k["dram_write_bytes"] for k in ks) / sum(k["launches"] for k in ks))
        pl = [p for p in d.get("per_launch", []) if "lora_gemm_kernel" in p.get("kernel", "") and "tc_active" in p]
        tc = round(sum(p["tc_active"] * p["dur"] for p in pl) / sum(p["dur"] for p in pl), 1) if pl else None
        return traffic, tc
    except Exception:
        return None, None


def backward_roofline(engine, pk: dict, reps: int = 5) -> dict:
    """The LoRA backward launches of one grad step (input-gradient GEMMs + the low-rank gradient kernels), replayed
    back to back inside a CUDA graph: algorithmic FLOPs 2MNK (dx) + 4Mr(K+N) (dh, dx +=, dA, dB) per wrapped layer."""
    from sliders_conceptmod_b200 import ops
    calls = []
    real_ig, real_g, real_f = ops.lora_input_grad, ops.lora_grad, ops.lora_backward_1x1

    def spy_f(gy, x, h, wt, a, b, s, dA, dB, accumulate=False, bt=None):
        calls.append(("fused", tuple(gy.shape), tuple(wt.shape), 1, b.shape[1], gy.dtype, {"train": dA is not None}))
        return real_f(gy, x, h, wt, a, b, s, dA, dB, accumulate=accumulate, bt=bt)

    def spy_ig(gy, wt, dh, at, cin, ksize=1, rank=0, **kw):
        calls.append(("dx", tuple(gy.shape), tuple(wt.shape), ksize, rank if dh is not None else 0, gy.dtype, dict(kw)))
        return real_ig(gy, wt, dh, at, cin, ksize, rank, **kw)

    def spy_g(gy, x, h, b, s, ksize, dA, dB, accumulate=False, **kw):
        calls.append(("grad", tuple(gy.shape), tuple(x.shape), ksize, b.shape[1], gy.dtype, dict(kw)))
        return real_g(gy, x, h, b, s, ksize, dA, dB, accumulate=accumulate, **kw)

    ops.lora_input_grad, ops.lora_grad, ops.lora_backward_1x1 = spy_ig, spy_g, spy_f
    try:
        engine._body_train()
    finally:
        ops.lora_input_grad, ops.lora_grad, ops.lora_backward_1x1 = real_ig, real_g, real_f
    torch.cuda.synchronize()
    flops, args = 0.0, []
    for kind, gs, other, k, r, dt, kw in calls:
        M = 1
        for d in gs[:-1]:
            M *= d
        N = gs[-1]
        if kind == "fused":       # 1x1: dx + g in one GEMM launch, dA + dB in one reduction launch
            cin = other[0]
            flops += 2.0 * M * N * cin + 2.0 * M * r * (N + cin) + (2.0 * M * r * (N + cin) if kw["train"] else 0.0)
            gy = torch.randn(gs, device="cuda", dtype=dt)
            x = torch.randn(*gs[:-1], cin, device="cuda", dtype=dt)
            h = torch.randn(M, r, device="cuda", dtype=torch.float32)
            wt = torch.randn(other, device="cuda", dtype=dt) * 0.02
            a = torch.randn(r, 1, cin, device="cuda", dtype=dt) * 0.02
            b = torch.randn(N, r, device="cuda", dtype=dt) * 0.02
            dA = torch.zeros(r, 1, cin, device="cuda", dtype=torch.float32) if kw["train"] else None
            dB = torch.zeros(N, r, device="cuda", dtype=torch.float32) if kw["train"] else None
            args.append(("fused", gy, x, h, wt, a, b, dA, dB, b.t().contiguous()))
            continue
        if kind == "dx":
            cin, taps, _ = other
            stride = kw.get("stride", 1)
            flops += 2.0 * M * N * taps * cin + (2.0 * M * r * taps * cin if r else 0.0)
            gy = torch.randn(gs, device="cuda", dtype=dt)
            wt = torch.randn(other, device="cuda", dtype=dt) * 0.02
            dh = torch.randn(M, 16, device="cuda", dtype=dt) if r else None
            at = torch.randn(cin, taps, 16, device="cuda", dtype=dt) * 0.02 if r else None
            args.append(("dx", gy, wt, dh, at, cin, k, r, kw))
        else:
            cin = other[-1]
            taps = k * k
            flops += 2.0 * M * r * N * 2 + 2.0 * M * r * taps * cin
            gy = torch.randn(gs, device="cuda", dtype=dt)
            x = torch.randn(other, device="cuda", dtype=dt)
            Mx = x.numel() // cin
            h = torch.randn(M, r, device="cuda", dtype=torch.float32)
            b = torch.randn(N, r, device="cuda", dtype=dt) * 0.02
            dA = torch.zeros(r, taps, cin, device="cuda", dtype=torch.float32)
            dB = torch.zeros(N, r, device="cuda", dtype=torch.float32)
            args.append(("grad", gy, x, h, b, k, dA, dB, kw))

    def run():
        for c in args:
            if c[0] == "fused":
                real_f(c[1], c[2], c[3], c[4], c[5], c[6], 0.25, c[7], c[8], accumulate=True, bt=c[9])
            elif c[0] == "dx":
                real_ig(c[1], c[2], c[3], c[4], c[5], c[6], c[7], **c[8])
            else:
                real_g(c[1], c[2], c[3], c[4], 0.25, c[5], c[6], c[7], accumulate=True, **c[8])
    n0 = ops.launch_count()
    run(); torch.cuda.synchronize()
    launches = ops.launch_count() - n0
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        run()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    sec = e0.elapsed_time(e1) * 1e-3 / reps
    achieved = flops / sec / 1e12
    peak = float(pk.get("bf16_tflops_sustained", pk["bf16_tflops"]))
    return {"bound": "tensor", "kernels": "lora_gemm_kernel (1x1: dx + gy.B fused; 3x3: dx extended-K, dh) + colreduce_kernel (dA, dB)", "achieved": round(achieved, 1),
            "peak": peak, "unit": "TFLOP/s", "frac": round(achieved / peak, 4), "frac_burst": round(achieved / float(pk["bf16_tflops"]), 4),
            "entry_calls": len(calls), "launches": int(launches), "ms": round(sec * 1e3, 3), "flops_per_backward": flops}


def phase_breakdown(engine, reps: int = 5) -> dict:
    """Device time of each captured phase of a step (CUDA events around graph replays): one adapter-on CFG denoise
    forward (x timesteps_to per step), the batched adapter-off predictions, the grad forward+loss+backward, AdamW."""
    out = {}
    saved = engine.lat.clone()
    for name in ("denoise", "predict", "train", "predict_train"):
        g = engine._graphs.get(name)
        if g is None:
            continue
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        out[name] = round(e0.elapsed_time(e1) / reps, 3)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        engine.optimizer_step()
    e1.record()
    torch.cuda.synchronize()
    out["optimizer"] = round(e0.elapsed_time(e1) / reps, 3)
    engine.lat.copy_(saved)
    return out


def config_dict(wl, n_modules: int, n_params: int, gpus: int) -> dict:
    """`config` of the JSON line: identical for both arms (the reference arm runs OUR arm's configuration)."""
    return {"workload": wl.name, "timesteps_to": TIMESTEPS_TO, "rank": wl.rank, "alpha": wl.alpha,
            "resolution": wl.resolution, "batch_size": wl.batch_size, "lora_modules": n_modules,
            "lora_params": n_params, "scheduler": wl.scheduler,
            "parallelism": f"dp{gpus} (prompt pairs sharded, one flat-grad all-reduce)",
            "l2": "working set (1.7 GB bf16 UNet weights) exceeds the 126 MB L2; no explicit flush"}


# --------------------------------------------------------------------------------- CPU arm (oracle port)
class OracleRun:
    """The reference's iteration (conceptmod/textsliders/train_lora.py:155-309) restated with plain PyTorch ops
    (oracle/slider_oracle.py) over the same random-init SD-1.4-shaped UNet.  On the host cores in fp32 it is the
    CPU baseline / `--impl reference` arm; on the GPU in bf16 (`device="cuda"`, cuBLAS / cuDNN / SDPA eager, no
    libslider kernel) it is the "reference arithmetic through PyTorch eager on the same B200" baseline of SURVEY 8d.
    `src`: (unet, network) whose parameter values are copied, so that losses are comparable with the CUDA engine."""

    def __init__(self, wl_name: str, device="cpu", dtype=torch.float32, src=None, threads: int | None = None):
        from oracle import slider_oracle
        from sliders_conceptmod_b200 import workload
        from sliders_conceptmod_b200.lora import LoRANetwork
        from sliders_conceptmod_b200.unet import UNet2DConditionModel
        if threads:
            torch.set_num_threads(threads)
        self.so, self.device, self.dtype = slider_oracle, torch.device(device), dtype
        self.wl = wl = workload.WORKLOADS[wl_name]
        cfg = workload.unet_config(wl.family)
        torch.manual_seed(0)
        with torch.device(self.device):
            unet = UNet2DConditionModel(cfg)
        unet.to(dtype=dtype).eval().requires_grad_(False)
        if src is not None:
            unet.load_state_dict(src[0].state_dict())
        net = LoRANetwork(unet, rank=wl.rank, alpha=wl.alpha, target_replace=workload.target_modules(wl.c3lier),
                          train_method=wl.training_method)
        if src is not None:
            net.load_state_dict({k: v.detach().to("cpu") for k, v in src[1].state_dict().items()})
        net.to(self.device, dtype=dtype)
        slider_oracle.attach(net)
        self.unet, self.net, self.cfg = unet, net, cfg
        self.sched = slider_oracle.OracleScheduler(wl.scheduler)
        self.opt = torch.optim.AdamW(net.prepare_optimizer_params(), lr=wl.lr)
        self.pairs = workload.prompt_pairs(wl, cfg.cross_attention_dim)

    def iteration(self, pair: dict, latents: torch.Tensor, timesteps_to: int, round_to=None) -> tuple[float, float]:
        """One full reference iteration incl. AdamW.  Returns (loss, seconds)."""
        from sliders_conceptmod_b200 import unet as unet_mod
        rd = (lambda x: x.to(round_to).to(self.dtype)) if round_to is not None else (lambda x: x.to(self.dtype))
        emb = {k: rd(pair[k]).to(self.device) for k in ("target", "positive", "neutral", "unconditional")}
        lat = rd(latents).to(self.device)
        if self.device.type == "cuda":
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        self.opt.zero_grad()
        with unet_mod.torch_reference_mode():
            loss, _ = self.so.train_iteration(self.unet, self.net, self.sched, emb, lat, timesteps_to, self.wl.guidance_scale,
                                              self.wl.action, max_denoising_steps=self.wl.max_denoising_steps)
        self.opt.step()
        val = float(loss.item())          # train_lora.py:292 reads the loss back every iteration
        return val, time.perf_counter() - t0


def run_reference_arm(args):
    """`--impl reference`: the oracle port on all host cores, REAL iterations (25 adapter-on CFG forwards + 3 adapter-off
    predictions + grad forward / loss / backward + AdamW each), as many as fit `--cpu-budget-s`; `steps` in the
    printed line is the number actually timed."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    run = OracleRun(args.workload, "cpu", torch.float32, threads=threads)
    wl = run.wl
    gen = torch.Generator().manual_seed(1234)
    shape = (wl.batch_size, 4, wl.resolution // 8, wl.resolution // 8)
    t_start = time.perf_counter()
    n_warm = 0
    if args.warmup > 0:                      # one real iteration warms the allocator / thread pool
        run.iteration(run.pairs[0], torch.randn(shape, generator=gen), TIMESTEPS_TO)
        n_warm = 1
    times = []
    while len(times) < args.steps:
        _, sec = run.iteration(run.pairs[len(times) % len(run.pairs)], torch.randn(shape, generator=gen), TIMESTEPS_TO)
        times.append(sec)
        spent = time.perf_counter() - t_start
        if spent + 1.2 * sec > args.cpu_budget_s:
            break
    sec = sum(times) / len(times)
    val = 1.0 / sec
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": len(times), "steps_requested": args.steps, "warmup": n_warm, "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(wl, len(run.net.unet_loras), sum(p.numel() for p in run.net.parameters()), args.gpus),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{len(times)} full iterations of the oracle port (oracle/slider_oracle.py: "
                                       f"{TIMESTEPS_TO} adapter-on CFG forwards + 3 predictions + grad forward/loss/backward + "
                                       f"AdamW), fp32, bounded by --cpu-budget-s {args.cpu_budget_s:.0f}"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch.distributed as dist
    from sliders_conceptmod_b200 import workload
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    wl = workload.WORKLOADS[args.workload]
    dtype = torch.bfloat16
    engine, unet, network, pairs = workload.build(wl, dev, dtype, seed=0, use_graphs=not args.no_graphs,
                                                  lora_b_std=0.0)
    pair = pairs[rank % len(pairs)]              # prompt pairs / attributes sharded over ranks
    workload.load_pair(engine, wl, pair)
    lat_shape = (wl.batch_size, 4, wl.resolution // 8, wl.resolution // 8)
    gen = torch.Generator().manual_seed(1234 + rank)
    n_lat = args.steps + args.warmup + 2
    host_lat = torch.randn((n_lat,) + lat_shape, generator=gen).pin_memory()     # e2e: per-step H2D source
    dev_lat = host_lat.to(dev)                                                    # value: inputs resident in HBM
    host_loss = torch.zeros(1).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item() * 1e-3

    resident = lambda i: engine.train_step(dev_lat[args.warmup + i], TIMESTEPS_TO)

    def end_to_end(i):       # host latents in, host loss out, every step
        loss = engine.train_step(host_lat[args.warmup + i], TIMESTEPS_TO)
        host_loss.copy_(loss.reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    loss0 = None
    init_state = {k: v.detach().clone() for k, v in network.state_dict().items()} if rank == 0 else None
    for i in range(max(args.warmup, 3)):
        l = engine.train_step(dev_lat[i % n_lat], TIMESTEPS_TO)
        if i == 0:
            loss0 = float(l.item())      # step 0: weights at their initial values (checked against the oracle below)
    barrier()
    launches0 = engine.kernel_launches
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    sec = timed(resident, args.steps)
    launches = engine.kernel_launches - launches0
    sec_e2e = timed(end_to_end, args.steps)
    clk = clocks.stop() if rank == 0 else None
    final_loss = float(host_loss.item())
    # secondary: the reference's own draw timesteps_to = randint(1, max_denoising_steps)
    # (train_lora.py:168-170), seeded per rank
    draws = [int(torch.randint(1, wl.max_denoising_steps, (1,), generator=gen)) for _ in range(args.steps)]
    sec_rand = timed(lambda i: engine.train_step(dev_lat[args.warmup + i], draws[i]), args.steps)
    if args.profile_step and rank == 0:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        engine.train_step(dev_lat[args.warmup], TIMESTEPS_TO)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pk = peaks()
    roof = fused_gemm_roofline(engine, pk)
    phases = phase_breakdown(engine) if (not args.no_graphs and world == 1) else None   # optimizer_step all-reduces
    value = world * args.steps / sec
    value_e2e = world * args.steps / sec_e2e
    lat_bytes = host_lat[0].numel() * host_lat.element_size()
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": config_dict(wl, len(network.unet_loras), int(network.flat_param.numel()), world),
            "cuda_graphs": not args.no_graphs,
            "e2e": {"value": value_e2e, "unit": UNIT, "h2d_bytes_per_step": lat_bytes, "d2h_bytes_per_step": 4,
                    "ms_per_step": sec_e2e / args.steps * 1e3},
            "gpu_launches": int(launches), "clocks": clk, "roofline": roof, "final_loss": final_loss,
            "phases_ms": phases,
            "secondary": {"metric": "slider_train_steps_per_sec_random_timesteps", "value": world * args.steps / sec_rand,
                          "unit": UNIT, "timesteps_to": "randint(1, %d) seeded 1234+rank" % wl.max_denoising_steps,
                          "mean_timesteps_to_rank0": sum(draws) / len(draws)}}
    if world == 1 and wl.family == "sd14":
        try:
            line["roofline_bwd"] = backward_roofline(engine, pk)
        except Exception as e:      # measurement extra: never lose the headline line to it
            line["roofline_bwd"] = {"error": str(e)[:200]}
    if not args.skip_cpu_baseline and world == 1:
        # ONE real iteration of the oracle port on the host cores, on the SAME parameter values and the SAME step-0
        # inputs as the CUDA engine: it is both the bounded CPU-baseline sample and the check of the benched loss.
        threads = os.cpu_count() or 1
        src_net = type("S", (), {"state_dict": staticmethod(lambda: init_state)})
        cpu = OracleRun(args.workload, "cpu", torch.float32, src=(unet, src_net), threads=threads)
        loss_c, s = cpu.iteration(pair, host_lat[0], TIMESTEPS_TO, round_to=dtype)
        rel = abs(loss0 - loss_c) / max(abs(loss_c), 1e-30)
        line["cpu_baseline"] = {"value": 1.0 / s, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"1 full iteration of the oracle port (fp32, {TIMESTEPS_TO} adapter-on CFG forwards + 3 "
                                          f"predictions + grad forward/loss/backward + AdamW) on the bench's step-0 inputs"}
        line["loss_check"] = {"step": 0, "oracle_fp32": loss_c, "cuda_bf16": loss0, "rel": rel,
                              "note": "lora_up = 0 at init (lora.py:124) and target == neutral prompt, so the step-0 loss is "
                                      "g^2 * mean((positive - unconditional)^2): a difference of two bf16 UNet evaluations"}
        del cpu
    if not args.skip_eager_baseline and world == 1:
        # the reference's arithmetic (plain torch ops: cuBLAS / cuDNN / SDPA, no libslider kernel, no CUDA graph) on the
        # same GPU in bf16 -- SURVEY 8d's "real bar"
        src_net = type("S", (), {"state_dict": staticmethod(lambda: init_state)})
        eager = OracleRun(args.workload, dev, dtype, src=(unet, src_net))
        eager.iteration(pair, host_lat[0], TIMESTEPS_TO)
        ts = [eager.iteration(pair, host_lat[1 + k], TIMESTEPS_TO)[1] for k in range(3)]
        s = sum(ts) / len(ts)
        line["gpu_eager_baseline"] = {"value": 1.0 / s, "unit": UNIT, "ms_per_step": s * 1e3, "ours_over_eager": value * s,
                                      "kind": "oracle arithmetic = LoRAModule.forward as 3 torch ops per layer "
                                              "(lora.py:134-138) + torch AdamW, PyTorch eager bf16 on the same GPU, 3 iterations"}
        del eager
        torch.cuda.empty_cache()
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    chk = line.get("loss_check")
    if chk is not None and not (chk["rel"] <= 0.5):
        raise SystemExit(f"bench: step-0 loss {chk['cuda_bf16']} is not the oracle's {chk['oracle_fp32']}")


def main():
    global TIMESTEPS_TO
    args = parse_args()
    TIMESTEPS_TO = args.timesteps_to
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
