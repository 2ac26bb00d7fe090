#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for m in 8 0 1000; do
SLIDER_PLAIN_MAX_GFLOP=$m timeout 900 python bench.py --workload sdxl_text_slider_rank4_1024px --timesteps-to 6 --steps 3 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/r37_xl_$m.json 2> gpurun_out/r37_xl_$m.err
done
timeout 900 python bench.py --steps 6 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/r37_sd14.json 2> gpurun_out/r37_sd14.err
python - <<'PY'
import json
for n in ('r37_xl_8','r37_xl_0','r37_xl_1000','r37_sd14'):
    try:
        d=json.loads(open(f'gpurun_out/{n}.json').read().strip().splitlines()[-1])
        r=d['roofline']
        print(n, round(d['value'],3), d['phases_ms'], 'roofline', r['frac'], r['launches'], r.get('without_frozen_linear'))
    except Exception as e: print(n, 'failed', e)
PY
