#!/bin/bash
# secondary workloads in the current state: SD-XL text slider bench line, configs 3 / 4 / 5
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python bench.py --workload sdxl_text_slider_rank4_1024px --timesteps-to 6 --steps 3 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/bench_r02_xl_v2.json 2> gpurun_out/bench_r02_xl_v2.err
echo "xl rc=$?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r02_xl_v2.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['phases_ms'], d['roofline']['frac'], d['roofline']['avg_launch_us'])
PY
timeout 1500 python tools/bench_configs.py > gpurun_out/r02_bench_configs_v2.jsonl 2> gpurun_out/r02_bench_configs_v2.err
echo "configs rc=$?"; cut -c1-400 gpurun_out/r02_bench_configs_v2.jsonl
