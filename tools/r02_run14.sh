#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm_raw.py tests/test_gpu_golden.py tests/test_gpu_bwd_raw.py tests/test_gpu_dora.py -x -q -m gpu > gpurun_out/r02_t14.log 2>&1; echo "t14 rc=$?"; tail -12 gpurun_out/r02_t14.log
timeout 300 python tools/trace_gemm.py 1,1,8192,320,320,1 1,1,2048,640,640,1 2>&1 | grep -v "^ *-"
timeout 600 python bench.py --steps 5 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/bench_r02_v6.json 2>/dev/null
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r02_v6.json').read().strip().splitlines()[-1])
print(round(d['value'],3), d['phases_ms'], d['roofline']['frac'], d['roofline']['avg_launch_us'], d['loss_check'])
PY
