#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_gemm_raw.py -x -q -m gpu -k "pair" > gpurun_out/r02_pair.log 2>&1; echo "pair rc=$?"; tail -15 gpurun_out/r02_pair.log
