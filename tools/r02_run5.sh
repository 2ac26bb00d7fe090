#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02_tall2.log 2>&1; echo "tall rc=$?"; tail -4 gpurun_out/r02_tall2.log
for ch in 1 5; do
  SLIDER_DENOISE_CHUNK=$ch timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02_chunk$ch.json 2> gpurun_out/bench_r02_chunk$ch.err
  echo "bench chunk=$ch rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r02_chunk$ch.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['phases_ms'], d['roofline_bwd']['ms'], d['secondary']['value'])
PY
done
timeout 600 python tools/torch_profile.py predict_train > gpurun_out/r02_torch_profile_predict_train_v2.txt 2>&1; grep -v Warn gpurun_out/r02_torch_profile_predict_train_v2.txt | head -30 | cut -c1-140
