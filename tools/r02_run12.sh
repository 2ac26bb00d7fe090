#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
for m in 0 -1 0 -1; do
  SLIDER_PAIR=$m timeout 600 python bench.py --steps 5 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/bench_pair_$m.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_pair_$m.json').read().strip().splitlines()[-1])
print('pair=$m', round(d['value'],3), d['phases_ms'], d['roofline']['avg_launch_us'])
PY
done
timeout 300 python tools/trace_gemm.py 1,1,8192,320,320,1 1,1,8192,320,960,1 2,64,64,320,320,3 1,1,2048,640,640,1 2>&1 | grep -v "^ *-"
