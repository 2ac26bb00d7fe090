#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python tools/trace_insitu.py > gpurun_out/r23_trace_insitu.txt 2>&1
tail -3 gpurun_out/r23_trace_insitu.txt
timeout 300 python tools/bench_geglu.py > gpurun_out/r23_geglu.txt 2>&1; cat gpurun_out/r23_geglu.txt | tail -5
timeout 600 python -m pytest tests/test_gpu_transformer_ew.py -m gpu -x -q 2>&1 | tail -3
