#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm_raw.py tests/test_gpu_golden.py tests/test_gpu_bwd_raw.py tests/test_gpu_dora.py tests/test_gpu_fullsize.py -x -q -m gpu > gpurun_out/r02_t16.log 2>&1; echo "t16 rc=$?"; tail -5 gpurun_out/r02_t16.log
timeout 300 python tools/trace_gemm.py 1,1,8192,320,320,1 2>&1 | grep -v "^ *-\|chunk0"
for d in 0 1 0 1; do
SLIDER_DIRECT_STORE=$d timeout 600 python bench.py --steps 5 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/bench_direct_$d.json 2>/dev/null
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_direct_$d.json').read().strip().splitlines()[-1])
print('direct=$d', round(d['value'],3), d['phases_ms'], d['roofline']['frac'], d['roofline']['avg_launch_us'])
PY
done
