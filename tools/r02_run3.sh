#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dora.py -x -q -m gpu > gpurun_out/r02_dora.log 2>&1; echo "dora rc=$?"; tail -25 gpurun_out/r02_dora.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02_v2.json 2> gpurun_out/bench_r02_v2.err
echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_r02_v2.json
