#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02_tall3.log 2>&1; echo "tall rc=$?"; tail -4 gpurun_out/r02_tall3.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02_v4.json 2> gpurun_out/bench_r02_v4.err
echo "bench rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r02_v4.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['phases_ms'], d['roofline']['frac'], d['roofline']['avg_launch_us'], d['roofline_bwd']['ms'])
PY
timeout 1500 python bench.py --workload sdxl_text_slider_rank4_1024px --timesteps-to 6 --steps 3 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/bench_r02_xl_v1.json 2> gpurun_out/bench_r02_xl_v1.err
echo "xl rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r02_xl_v1.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['achieved'], d['roofline']['avg_launch_us'], d['clocks'])
PY
timeout 1200 python tools/bench_configs.py 2>/dev/null | tail -2
