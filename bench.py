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
    ap.add_argument("--no-graphs", action="store_true")
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget-s", type=float, default=200.0)
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

    def spy(x, w, bias, a, b, s, ksize=1, want_h=False, residual=None):
        calls.append(("single", tuple(x.shape), tuple(w.shape), bias is not None, None if a is None else a.shape[0], s,
                      ksize, x.dtype, residual is not None))
        return real(x, w, bias, a, b, s, ksize=ksize, want_h=want_h, residual=residual)

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
        alg_bytes += es * (float(torch.Size(xs).numel()) + float(torch.Size(ws).numel()) +
                           float(torch.Size(xs[:-1]).numel()) * ws[0] * (2 if has_res else 1))
        M = 1
        for d in xs[:-1]:
            M *= d
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
        args.append(("single", x, w, bias, a, b, s, k, res))

    def run():
        for c in args:
            if c[0] == "grouped":
                real_grouped(*c[1:])
            elif c[0] == "multi":
                real_multi(c[1], c[2], c[3])
            else:
                real(c[1], c[2], c[3], c[4], c[5], c[6], ksize=c[7], residual=c[8])
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
    sec = e0.elapsed_time(e1) * 1e-3 / reps
    achieved = flops / sec / 1e12
    peak = float(pk.get("bf16_tflops_sustained", pk["bf16_tflops"]))
    return {"bound": "tensor", "kernel": "lora_gemm_kernel", "achieved": round(achieved, 1), "peak": peak,
            "peak_source": pk["source"] + ":bf16_tflops_sustained", "unit": "TFLOP/s", "frac": round(achieved / peak, 4),
            "traffic": measured_traffic(), "launches": len(calls), "avg_launch_us": round(sec / len(calls) * 1e6, 2),
            "flops_per_forward": flops, "algorithmic_bytes_per_launch": round(alg_bytes / len(calls))}


def measured_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per lora_gemm_kernel launch, from the committed `ncu --set full`
    capture of the same forward (tools/profile_forward.py -> tools/ncu_traffic.py -> profiles/gemm_traffic.json)."""
    f = ROOT / "profiles" / "gemm_traffic.json"
    if not f.exists():
        return None
    try:
        ks = [v for n, v in json.loads(f.read_text())["kernels"].items() if "lora_gemm_kernel" in n]
        return round(sum(k["dram_read_bytes"] + k["dram_write_bytes"] for k in ks) / sum(k["launches"] for k in ks))
    except Exception:
        return None


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


# --------------------------------------------------------------------------------- CPU arm (oracle port)
def cpu_reference_sample(wl_name: str, threads: int):
    """Builds the reference's CPU path for the workload: the oracle arithmetic (oracle/slider_oracle.py) over the
    same random-init SD-1.4-shaped UNet in fp32 on the host cores.  Returns a closure timing ONE bounded sample:
    one adapter-on no-grad CFG forward (t_f) and one grad forward + loss + backward (t_fb)."""
    from oracle import slider_oracle
    from sliders_conceptmod_b200 import workload
    from sliders_conceptmod_b200.lora import LoRANetwork
    from sliders_conceptmod_b200.unet import UNet2DConditionModel
    torch.set_num_threads(threads)
    wl = workload.WORKLOADS[wl_name]
    cfg = workload.unet_config(wl.family)
    torch.manual_seed(0)
    unet = UNet2DConditionModel(cfg).eval().requires_grad_(False)
    net = LoRANetwork(unet, rank=wl.rank, alpha=wl.alpha, target_replace=workload.target_modules(wl.c3lier),
                      train_method=wl.training_method)
    slider_oracle.attach(net)
    sched = slider_oracle.OracleScheduler("ddim")
    sched.set_timesteps(1000)
    pair = workload.prompt_pairs(wl, cfg.cross_attention_dim)[0]
    lat = torch.randn(wl.batch_size, 4, wl.resolution // 8, wl.resolution // 8)
    emb = torch.cat([pair["unconditional"], pair["target"]])
    t = sched.timesteps[500]

    def sample():
        t0 = time.perf_counter()
        with torch.no_grad(), net:
            slider_oracle.predict_noise(unet, sched, t, lat, emb, 3.0)
        t1 = time.perf_counter()
        with net:
            tgt = slider_oracle.predict_noise(unet, sched, t, lat, emb, 1.0)
        loss = slider_oracle.guidance_loss(tgt, lat, lat, lat, wl.guidance_scale, wl.action)
        loss.backward()
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1

    return sample


def step_seconds_from_sample(t_f: float, t_fb: float) -> float:
    # train_lora.py:155-309: TIMESTEPS_TO denoise forwards + 3 predictions (no grad) + 1 grad forward/backward
    return (TIMESTEPS_TO + 3) * t_f + t_fb


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = cpu_reference_sample(args.workload, threads)
    t_start = time.perf_counter()
    for _ in range(args.warmup):
        sample()
        if time.perf_counter() - t_start > args.cpu_budget_s * 0.3:
            break
    times = []
    t_meas = time.perf_counter()
    for _ in range(args.steps):
        times.append(step_seconds_from_sample(*sample()))
        if time.perf_counter() - t_meas > args.cpu_budget_s * 0.7:
            break
    sec = sum(times) / len(times)
    val = 1.0 / sec
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "timesteps_to": TIMESTEPS_TO, "rank": 4, "resolution": 512},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"per step: 1 no-grad CFG UNet forward + 1 grad forward/loss/backward at full size, "
                                       f"step time = {TIMESTEPS_TO + 3}*t_fwd + t_fwd_bwd; {len(times)} samples measured"},
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

    for i in range(max(args.warmup, 3)):
        engine.train_step(dev_lat[i % n_lat], TIMESTEPS_TO)
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
            "config": {"workload": args.workload, "timesteps_to": TIMESTEPS_TO, "rank": wl.rank, "alpha": wl.alpha,
                       "resolution": wl.resolution, "batch_size": wl.batch_size, "lora_modules": len(network.unet_loras),
                       "lora_params": int(network.flat_param.numel()), "scheduler": wl.scheduler,
                       "parallelism": f"dp{world} (prompt pairs sharded, one flat-grad all-reduce)",
                       "cuda_graphs": not args.no_graphs,
                       "l2": "working set (1.7 GB bf16 UNet weights) exceeds the 126 MB L2; no explicit flush"},
            "e2e": {"value": value_e2e, "unit": UNIT, "h2d_bytes_per_step": lat_bytes, "d2h_bytes_per_step": 4,
                    "ms_per_step": sec_e2e / args.steps * 1e3},
            "gpu_launches": int(launches), "clocks": clk, "roofline": roof, "final_loss": final_loss,
            "phases_ms": phases}
    if not args.skip_cpu_baseline and world == 1:
        threads = os.cpu_count() or 1
        sample = cpu_reference_sample(args.workload, threads)
        t0 = time.perf_counter()
        vals = []
        while len(vals) < 2 and time.perf_counter() - t0 < 60.0:
            vals.append(step_seconds_from_sample(*sample()))
        s = min(vals)
        line["cpu_baseline"] = {"value": 1.0 / s, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"oracle port, fp32, {len(vals)} samples of (1 no-grad CFG forward + 1 grad "
                                          f"forward/backward) at full size; step = {TIMESTEPS_TO + 3}*t_fwd + t_fwd_bwd"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
