#!/bin/bash
# final records: default bench line, reference arm (1 step), N=1 configs 3/4/5
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r38_bench_final.json 2> gpurun_out/r38_bench_final.err; echo "bench exit $?"
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r38_bench_reference.json 2> gpurun_out/r38_bench_reference.err; echo "ref exit $?"
timeout 1500 python tools/bench_configs.py > gpurun_out/r38_bench_configs.jsonl 2> gpurun_out/r38_bench_configs.err; echo "configs exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r38_bench_final.json').read().strip().splitlines()[-1])
print(round(d['value'],3), round(d['ms_per_step'],2), d['phases_ms'], 'e2e', round(d['e2e']['value'],3), 'roofline', d['roofline']['frac'], d['roofline']['without_frozen_linear'], 'bwd', d['roofline_bwd']['frac'], 'cpu', d['cpu_baseline']['value'], 'eager', d['gpu_eager_baseline']['value'], d['clocks'], d['gpu_launches'], d['secondary']['value'])
r=json.loads(open('gpurun_out/r38_bench_reference.json').read().strip().splitlines()[-1])
print('reference', r['value'], r['steps'], r.get('cpu_baseline'))
print(open('gpurun_out/r38_bench_configs.jsonl').read()[:1500])
PY
