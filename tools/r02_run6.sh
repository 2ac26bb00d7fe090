#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_transformer_ew.py tests/test_gpu_fullsize.py tests/test_gpu_network.py -x -q -m gpu > gpurun_out/r02_t6.log 2>&1; echo "t6 rc=$?"; tail -4 gpurun_out/r02_t6.log
for dis in "" ffgeglu; do
  SLIDER_DISABLE=$dis timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02_geglu_$dis.json 2> gpurun_out/bench_r02_geglu_$dis.err
  echo "bench disable=[$dis] rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r02_geglu_$dis.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['phases_ms'], d['loss_check']['rel'])
PY
done
