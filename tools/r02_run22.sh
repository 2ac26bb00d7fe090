#!/bin/bash
# r02 run 22: GEGLU launch timings vs cuBLAS; A/B of routing the frozen residual projections (attn2.to_out, ff.net.2) through the fused kernel
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python tools/bench_geglu.py > gpurun_out/r22_geglu.txt 2>&1; cat gpurun_out/r22_geglu.txt | tail -5
SLIDER_ENABLE=plainres timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r22_bench_plainres.json 2> gpurun_out/r22_bench_plainres.err
SLIDER_ENABLE=plain timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r22_bench_plain.json 2> gpurun_out/r22_bench_plain.err
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r22_bench_default.json 2> gpurun_out/r22_bench_default.err
python - <<'PY'
import json
for n in ('plainres','plain','default'):
    try:
        d=json.loads(open(f'gpurun_out/r22_bench_{n}.json').read().strip().splitlines()[-1])
        print(n, round(d['value'],3), round(d['ms_per_step'],2), d['phases_ms'])
    except Exception as e: print(n, 'failed', e)
PY
