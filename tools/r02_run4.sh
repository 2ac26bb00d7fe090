#!/bin/bash
# profiling pass: in-situ kernel breakdown of the two graphs (torch profiler), then ncu --set full of the bandwidth kernels
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python tools/torch_profile.py denoise > gpurun_out/r02_torch_profile_denoise.txt 2>&1; tail -32 gpurun_out/r02_torch_profile_denoise.txt | cut -c1-150
timeout 600 python tools/torch_profile.py predict_train > gpurun_out/r02_torch_profile_predict_train.txt 2>&1; tail -45 gpurun_out/r02_torch_profile_predict_train.txt | cut -c1-150
timeout 300 python tools/ncu_elementwise.py > gpurun_out/ncu_ew_plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none -k regex:'guidance_loss|cfg_step|adamw|clip_coef|colreduce' -c 40 -o gpurun_out/r02_ew_full python tools/ncu_elementwise.py > gpurun_out/ncu_ew.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_ew.log
ncu -i gpurun_out/r02_ew_full.ncu-rep --page raw --csv > gpurun_out/r02_ew_full_raw.csv 2>/dev/null; wc -l gpurun_out/r02_ew_full_raw.csv
