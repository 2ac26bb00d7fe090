#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fp16.py -x -q -m gpu > gpurun_out/r02_fp16.log 2>&1; echo "fp16 rc=$?"; tail -25 gpurun_out/r02_fp16.log
