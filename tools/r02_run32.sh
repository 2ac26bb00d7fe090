#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
SLIDER_PDL=0 timeout 900 python tools/torch_profile.py denoise > gpurun_out/r02_torch_profile_denoise_v2.txt 2>&1; grep -v Warn gpurun_out/r02_torch_profile_denoise_v2.txt | tail -31 | cut -c1-150
timeout 900 python tools/torch_profile.py denoise 2>&1 | grep "graph replay"
