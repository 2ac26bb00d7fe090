#!/bin/bash
# final state: full GPU suite, smoke, default bench line, reference arm, configs, forward launch list input
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/ -x -q -m gpu > gpurun_out/val_gpu_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/val_gpu_tests.log
tail -3 gpurun_out/val_gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/val_smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/val_smoke.log
timeout 900 python bench.py > gpurun_out/val_bench_final.json 2> gpurun_out/val_bench_final.err; echo "bench exit $?"
timeout 1500 python tools/bench_configs.py > gpurun_out/val_bench_configs.jsonl 2> gpurun_out/val_bench_configs.err; echo "configs exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/val_bench_final.json').read().strip().splitlines()[-1])
print(round(d['value'],3), round(d['ms_per_step'],2), d['phases_ms'], 'e2e', round(d['e2e']['value'],3), 'roofline', d['roofline']['frac'], d['roofline']['without_frozen_linear'], 'bwd', d['roofline_bwd']['frac'], 'cpu', d['cpu_baseline']['value'], 'eager', d['gpu_eager_baseline']['value'], d['clocks'], d['gpu_launches'], d['secondary']['value'], d['loss_check'])
print(open('gpurun_out/val_bench_configs.jsonl').read()[:1400])
PY
