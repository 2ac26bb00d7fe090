#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
for en in "" plain plainres ""; do
SLIDER_ENABLE=$en timeout 600 python bench.py --steps 5 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/bench_en_$en.json 2>/dev/null
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_en_$en.json').read().strip().splitlines()[-1])
print('enable=[$en]', round(d['value'],3), d['phases_ms'])
PY
done
