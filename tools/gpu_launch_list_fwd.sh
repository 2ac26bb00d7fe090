#!/bin/bash
# final-state launch list of one eager adapter-on CFG forward (ncu, serialised, cold caches)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python tools/profile_forward.py > gpurun_out/pf_plain.log 2>&1 && \
timeout 1200 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_fwd_final.csv python tools/profile_forward.py > gpurun_out/pf_ncu.log 2>&1
echo "launch list rc=$?"; python tools/summarize_launches.py gpurun_out/r02_launches_fwd_final.csv > gpurun_out/r02_launches_fwd_final_summary.txt; head -14 gpurun_out/r02_launches_fwd_final_summary.txt | cut -c1-150
