#!/bin/bash
# frozen Linear layers on the fused kernel by default: full GPU suite, smoke, default bench line, SD-XL line, configs
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 2400 python -m pytest tests/ -x -q -m gpu > gpurun_out/r36_gpu_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r36_gpu_tests.log
tail -4 gpurun_out/r36_gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r36_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r36_smoke.log
timeout 900 python bench.py > gpurun_out/r36_bench.json 2> gpurun_out/r36_bench.err; echo "bench exit $?"
timeout 900 python bench.py --workload sdxl_text_slider_rank4_1024px --timesteps-to 6 --steps 3 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/r36_bench_xl.json 2> gpurun_out/r36_bench_xl.err; echo "xl exit $?"
python - <<'PY'
import json
for n in ('r36_bench','r36_bench_xl'):
    try:
        d=json.loads(open(f'gpurun_out/{n}.json').read().strip().splitlines()[-1])
        print(n, round(d['value'],3), round(d['ms_per_step'],2), d['phases_ms'], 'roofline', d['roofline']['frac'], d['roofline']['launches'], 'e2e', round(d['e2e']['value'],3), d.get('loss_check',{}).get('rel'))
    except Exception as e: print(n, 'failed', e)
PY
