#!/bin/bash
# final-state profiling artefacts: ncu launch list of one forward (+ predict + train), ncu --set full of the GEMM launches
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python tools/profile_forward.py > gpurun_out/pf_plain.log 2>&1 && \
timeout 1200 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_fwd.csv python tools/profile_forward.py > gpurun_out/pf_ncu.log 2>&1
echo "launch list rc=$?"; python tools/summarize_launches.py gpurun_out/r02_launches_fwd.csv > gpurun_out/r02_launches_fwd_summary.txt; head -12 gpurun_out/r02_launches_fwd_summary.txt
timeout 1200 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_train.csv python tools/profile_forward.py --train > gpurun_out/pf_ncu2.log 2>&1
echo "train list rc=$?"; python tools/summarize_launches.py gpurun_out/r02_launches_train.csv > gpurun_out/r02_launches_train_summary.txt; head -6 gpurun_out/r02_launches_train_summary.txt
timeout 2400 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:'lora_gemm_kernel|splitk_finalize' -o gpurun_out/r02_gemm_full python tools/profile_forward.py > gpurun_out/pf_ncu3.log 2>&1
echo "gemm full rc=$?"; tail -2 gpurun_out/pf_ncu3.log
ncu -i gpurun_out/r02_gemm_full.ncu-rep --page raw --csv > gpurun_out/r02_gemm_full_raw.csv 2>/dev/null; wc -l gpurun_out/r02_gemm_full_raw.csv
rm -f gpurun_out/r02_gemm_full.ncu-rep
