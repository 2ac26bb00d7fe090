#!/bin/bash
# round-2 GPU pass 1: new kernel paths first, then the whole suite, the layer-set A/B of the tile widths, the bench
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm_raw.py tests/test_gpu_bwd_raw.py tests/test_gpu_golden.py -x -q -m gpu > gpurun_out/r02_t1.log 2>&1
echo "t1 rc=$?"; tail -5 gpurun_out/r02_t1.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02_tall.log 2>&1
echo "tall rc=$?"; tail -5 gpurun_out/r02_tall.log
for thr in 0 -1; do
  SLIDER_BN80_TILES=$thr timeout 600 python tools/bench_layerset.py --json gpurun_out/layerset_bn80_$thr.json > gpurun_out/layerset_bn80_$thr.txt 2>&1
  tail -1 gpurun_out/layerset_bn80_$thr.txt
done
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02_v1.json 2> gpurun_out/bench_r02_v1.err
echo "bench rc=$?"; cut -c1-600 gpurun_out/bench_r02_v1.json
