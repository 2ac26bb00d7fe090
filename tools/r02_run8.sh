#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 1500 python bench.py --workload sdxl_text_slider_rank4_1024px --timesteps-to 6 --steps 3 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/bench_r02_xl.json 2> gpurun_out/bench_r02_xl.err
echo "xl bench rc=$?"; cut -c1-1500 gpurun_out/bench_r02_xl.json; tail -3 gpurun_out/bench_r02_xl.err
timeout 900 python tools/bench_layerset.py --workload sdxl_text_slider_rank4_1024px --json gpurun_out/layerset_xl.json > gpurun_out/layerset_xl.txt 2>&1; tail -1 gpurun_out/layerset_xl.txt
