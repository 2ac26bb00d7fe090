#!/bin/bash
# r02 run 20: concat / upsample rewrite (no 64-bit division) -> parity, then bench A/B
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_transformer_ew.py tests/test_gpu_fullsize.py tests/test_gpu_trainers.py tests/test_gpu_unet.py -m gpu -x -q > gpurun_out/r20_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r20_tests.log
tail -3 gpurun_out/r20_tests.log
timeout 600 python bench.py --steps 6 --warmup 3 > gpurun_out/r20_bench.json 2> gpurun_out/r20_bench.err
echo "bench exit $?"
cat gpurun_out/r20_bench.json | cut -c1-600
