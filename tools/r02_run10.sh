#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
for m in 0 1; do
  SLIDER_PAIR=$m timeout 600 python tools/bench_layerset.py --workload sdxl_text_slider_rank4_1024px --json gpurun_out/layerset_xl_pair$m.json > gpurun_out/layerset_xl_pair$m.txt 2>&1; echo "xl pair=$m"; grep -v "^{\"total" gpurun_out/layerset_xl_pair$m.txt | cut -c1-200 | head -5; tail -1 gpurun_out/layerset_xl_pair$m.txt
  SLIDER_PAIR=$m timeout 600 python tools/bench_layerset.py --json gpurun_out/layerset_sd14_pair$m.json > gpurun_out/layerset_sd14_pair$m.txt 2>&1; echo "sd14 pair=$m"; tail -1 gpurun_out/layerset_sd14_pair$m.txt
  SLIDER_PAIR=$m timeout 300 python tools/l2_bound_probe.py 2>&1 | grep "skip=0" 
done
