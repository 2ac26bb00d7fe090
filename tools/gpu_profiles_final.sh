#!/bin/bash
# final-state in-situ kernel lists (torch profiler over graph replays, dependent launch off so a duration is the
# kernel's own) + the ncu launch list of one eager forward
cd "${GRAFT_REPO_ROOT:-/root/repo}" || exit 1
mkdir -p gpurun_out
SLIDER_PDL=0 timeout 600 python tools/torch_profile.py denoise > gpurun_out/r02_torch_profile_denoise_v3_nopdl.txt 2>&1; echo "denoise rc=$?"
SLIDER_PDL=0 timeout 600 python tools/torch_profile.py predict_train > gpurun_out/r02_torch_profile_predict_train_v3_nopdl.txt 2>&1; echo "predict_train rc=$?"
timeout 600 python tools/torch_profile.py denoise 2>&1 | grep "graph replay"
grep "graph replay" gpurun_out/r02_torch_profile_*_v3_nopdl.txt
bash tools/gpu_launch_list_fwd.sh
