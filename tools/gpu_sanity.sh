#!/bin/bash
# shortest end-of-round check of the rebuilt library: smoke() + the two newest kernel families' parity tests
cd "${GRAFT_REPO_ROOT:-/root/repo}" || exit 1
mkdir -p gpurun_out
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/sanity_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/sanity_smoke.log | cut -c1-200
timeout 60 python -m pytest tests/test_gpu_attention.py tests/test_gpu_groupnorm.py -m gpu -x -q > gpurun_out/sanity_tests.log 2>&1; echo "tests rc=$?"; tail -1 gpurun_out/sanity_tests.log
