#!/bin/bash
# full GPU suite + smoke in the current state
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/ -x -q -m gpu > gpurun_out/r31_gpu_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r31_gpu_tests.log
tail -5 gpurun_out/r31_gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r31_smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/r31_smoke.log
