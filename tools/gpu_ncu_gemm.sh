#!/bin/bash
# ncu --set full of every fused-GEMM launch (and split-K finalize) of one eager adapter-on CFG forward -> raw CSV
# (tools/ncu_traffic.py turns it into profiles/gemm_traffic.json: DRAM traffic and tensor-pipe activity per launch)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python tools/profile_forward.py > gpurun_out/pf_plain3.log 2>&1 && \
timeout 2400 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:'lora_gemm_kernel|splitk_finalize' -o gpurun_out/r02_gemm_full python tools/profile_forward.py > gpurun_out/pf_ncu3.log 2>&1
echo "gemm full rc=$?"; tail -2 gpurun_out/pf_ncu3.log
ncu -i gpurun_out/r02_gemm_full.ncu-rep --page raw --csv > gpurun_out/r02_gemm_full_raw.csv 2>/dev/null; wc -l gpurun_out/r02_gemm_full_raw.csv
rm -f gpurun_out/r02_gemm_full.ncu-rep
