#!/bin/bash
# ncu --set full of the bandwidth-class kernels incl. the tensor-core reductions and the GroupNorm forward / backward
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 300 python tools/ncu_elementwise.py > gpurun_out/ncu_ew_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none -k regex:'guidance_loss|cfg_step|adamw|clip_coef|colreduce|gn_group|gn_bwd' -c 60 -o gpurun_out/r02_ew_full_v2 python tools/ncu_elementwise.py > gpurun_out/ncu_ew.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_ew.log
ncu -i gpurun_out/r02_ew_full_v2.ncu-rep --page raw --csv > gpurun_out/r02_ew_full_v2_raw.csv 2>/dev/null; wc -l gpurun_out/r02_ew_full_v2_raw.csv
python tools/ncu_ew_summary.py gpurun_out/r02_ew_full_v2_raw.csv > gpurun_out/r02_ncu_elementwise_summary_v2.txt; cat gpurun_out/r02_ncu_elementwise_summary_v2.txt | cut -c1-200
rm -f gpurun_out/r02_ew_full_v2.ncu-rep
