#!/bin/bash
# r02 run 25: tensor-core colreduce -> parity, determinism, bench A/B
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_bwd_raw.py tests/test_gpu_dora.py tests/test_gpu_golden.py tests/test_gpu_network.py tests/test_gpu_trainers.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/r25_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r25_tests.log
tail -15 gpurun_out/r25_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r25_bench_tc.json 2> gpurun_out/r25_bench_tc.err
SLIDER_COLREDUCE_TC=0 timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r25_bench_simt.json 2> gpurun_out/r25_bench_simt.err
python - <<'PY'
import json
for n in ('tc','simt'):
    try:
        d=json.loads(open(f'gpurun_out/r25_bench_{n}.json').read().strip().splitlines()[-1])
        print(n, round(d['value'],3), round(d['ms_per_step'],2), d['phases_ms'], d['roofline_bwd']['ms'], d['roofline_bwd']['frac'])
    except Exception as e: print(n, 'failed', e)
PY
