#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_groupnorm.py tests/test_gpu_trainers.py tests/test_gpu_network.py tests/test_gpu_fullsize.py tests/test_gpu_golden.py -m gpu -x -q > gpurun_out/r27_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r27_tests.log
tail -4 gpurun_out/r27_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r27_bench_fused.json 2> gpurun_out/r27_bench_fused.err
SLIDER_GN_BWD_GROUP=0 timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r27_bench_two.json 2> gpurun_out/r27_bench_two.err
python - <<'PY'
import json
for n in ('fused','two'):
    try:
        d=json.loads(open(f'gpurun_out/r27_bench_{n}.json').read().strip().splitlines()[-1])
        print(n, round(d['value'],3), round(d['ms_per_step'],2), d['phases_ms'])
    except Exception as e: print(n, 'failed', e)
PY
