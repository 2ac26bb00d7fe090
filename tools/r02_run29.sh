#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
SLIDER_PDL=0 timeout 900 python tools/torch_profile.py predict_train --workload sdxl_text_slider_rank4_1024px > gpurun_out/r02_torch_profile_xl_predict_train.txt 2>&1; grep -v Warn gpurun_out/r02_torch_profile_xl_predict_train.txt | tail -32 | cut -c1-150
SLIDER_PDL=0 timeout 900 python tools/torch_profile.py predict_train > gpurun_out/r02_torch_profile_predict_train_v2.txt 2>&1; grep -v Warn gpurun_out/r02_torch_profile_predict_train_v2.txt | tail -32 | cut -c1-150
