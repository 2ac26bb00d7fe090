#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_bwd_raw.py tests/test_gpu_dora.py tests/test_gpu_golden.py tests/test_gpu_trainers.py -m gpu -x -q > gpurun_out/r30_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r30_tests.log
tail -4 gpurun_out/r30_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r30_bench.json 2> gpurun_out/r30_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r30_bench.json').read().strip().splitlines()[-1])
print(round(d['value'],3), round(d['ms_per_step'],2), d['phases_ms'], d['roofline_bwd']['ms'], d['roofline_bwd']['frac'])
PY
SLIDER_PDL=0 timeout 900 python tools/torch_profile.py predict_train 2>&1 | grep -v Warn | grep "colreduce\|graph replay"
