#!/bin/bash
# final-state launch list of denoise forward + predictions + grad forward / loss / backward
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
timeout 600 python tools/profile_forward.py --train > gpurun_out/pf_plain2.log 2>&1 && \
timeout 1500 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_train_final.csv python tools/profile_forward.py --train > gpurun_out/pf_ncu2.log 2>&1
echo "train list rc=$?"; python tools/summarize_launches.py gpurun_out/r02_launches_train_final.csv > gpurun_out/r02_launches_train_final_summary.txt; head -24 gpurun_out/r02_launches_train_final_summary.txt | cut -c1-150
