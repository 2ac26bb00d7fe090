#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm_raw.py tests/test_gpu_golden.py tests/test_gpu_transformer_ew.py tests/test_gpu_bwd_raw.py -x -q -m gpu > gpurun_out/r02_t13.log 2>&1; echo "t13 rc=$?"; tail -3 gpurun_out/r02_t13.log
timeout 300 python tools/trace_gemm.py 1,1,8192,320,320,1 1,1,2048,640,640,1 2>&1 | grep -v "^ *-"
timeout 600 python bench.py --steps 5 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/bench_r02_v5.json 2>/dev/null
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r02_v5.json').read().strip().splitlines()[-1])
print(round(d['value'],3), d['phases_ms'], d['roofline']['frac'], d['roofline']['avg_launch_us'])
PY
timeout 600 python tools/bench_layerset.py --json gpurun_out/layerset_v3.json > gpurun_out/layerset_v3.txt 2>&1; tail -1 gpurun_out/layerset_v3.txt
