#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02_tall4.log 2>&1; echo "tall rc=$?"; tail -6 gpurun_out/r02_tall4.log
timeout 300 python tools/trace_gemm.py 1,1,8192,320,320,1 2>&1 | grep -v "^ *-\|chunk0"
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02_v7.json 2>gpurun_out/bench_r02_v7.err
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r02_v7.json').read().strip().splitlines()[-1])
print(round(d['value'],3), round(d['e2e']['value'],3), d['phases_ms'], d['roofline']['frac'], d['roofline']['avg_launch_us'], d['loss_check'])
PY
