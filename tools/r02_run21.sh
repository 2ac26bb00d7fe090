#!/bin/bash
# r02 run 21: branch-free erf (GEGLU epilogue), concat / upsample rewrite -> parity, then bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_transformer_ew.py tests/test_gpu_fullsize.py tests/test_gpu_trainers.py tests/test_gpu_network.py -m gpu -x -q > gpurun_out/r21_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r21_tests.log
tail -3 gpurun_out/r21_tests.log
timeout 600 python bench.py --steps 6 --warmup 3 > gpurun_out/r21_bench.json 2> gpurun_out/r21_bench.err
echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r21_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d.get('roofline'), {k:v for k,v in d.items() if 'ms' in k or 'denoise' in k})
PY
