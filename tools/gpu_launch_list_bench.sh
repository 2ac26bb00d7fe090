#!/bin/bash
# ncu launch list of ONE step of bench.py itself (the same command as the bench line; cold caches, serialised)
cd "${GRAFT_REPO_ROOT:-/root/repo}" || exit 1
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --skip-cpu-baseline --skip-eager-baseline --profile-step"
timeout 600 $CMD > gpurun_out/lb_plain.json 2> gpurun_out/lb_plain.err; rc=$?; echo "plain rc=$rc"; [ $rc -eq 0 ] || exit $rc
timeout 700 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
  --log-file gpurun_out/r02_launches_bench.csv $CMD > gpurun_out/lb_ncu.log 2>&1
echo "ncu rc=$?"
python tools/summarize_launches.py gpurun_out/r02_launches_bench.csv > gpurun_out/r02_launches_bench_summary.txt
head -16 gpurun_out/r02_launches_bench_summary.txt | cut -c1-150
