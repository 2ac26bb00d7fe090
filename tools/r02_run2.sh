#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
for sk in 0 1 2 3; do SLIDER_DBG_SKIP=$sk timeout 300 python tools/l2_bound_probe.py; done > gpurun_out/l2_probe.txt 2>&1
cat gpurun_out/l2_probe.txt
timeout 600 python tools/bench_layerset.py --json gpurun_out/layerset_v2.json > gpurun_out/layerset_v2.txt 2>&1; tail -1 gpurun_out/layerset_v2.txt
timeout 300 python -m pytest tests/test_gpu_gemm_raw.py -x -q -m gpu 2>&1 | tail -2
