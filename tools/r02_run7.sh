#!/bin/bash
# secondary workloads (BASELINE configs 3-5) with a clock record; SD-XL bench line with roofline
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > gpurun_out/r02_clocks_configs.csv &
SMI=$!
timeout 1500 python bench.py --workload sdxl_text_slider_rank4_1024px --steps 3 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/bench_r02_xl.json 2> gpurun_out/bench_r02_xl.err
echo "xl bench rc=$?"; cut -c1-400 gpurun_out/bench_r02_xl.json; tail -3 gpurun_out/bench_r02_xl.err
timeout 1500 python tools/bench_configs.py > gpurun_out/r02_bench_configs.jsonl 2> gpurun_out/r02_bench_configs.err
echo "configs rc=$?"; cat gpurun_out/r02_bench_configs.jsonl; tail -3 gpurun_out/r02_bench_configs.err
kill $SMI
python - <<'PY'
import csv, statistics
rows = list(csv.reader(open('gpurun_out/r02_clocks_configs.csv')))[1:]
sm = [float(r[1].split()[0]) for r in rows if len(r) > 8]
pw = [float(r[3].split()[0]) for r in rows if len(r) > 8]
bad = [r for r in rows if len(r) > 8 and any('Active' == x.strip() for x in r[5:8])]
cap = sum(1 for r in rows if len(r) > 8 and r[8].strip() == 'Active')
print(f"clocks: {len(sm)} samples, sm median {statistics.median(sm)} MHz, min {min(sm)}, max {max(sm)}, power max {max(pw)} W, slowdown samples {len(bad)}, sw_power_cap samples {cap}")
PY
