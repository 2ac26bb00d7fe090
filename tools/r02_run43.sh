#!/bin/bash
# ResNet shortcut convolution on a side stream (parallel graph branch): parity, then interleaved A/B
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_network.py tests/test_gpu_trainers.py tests/test_gpu_fp16.py -m gpu -x -q 2>&1 | tail -3 | cut -c1-250
for round in 1 2; do
for mode in on off; do
  if [ "$mode" = on ]; then env=""; else env="SLIDER_DISABLE=temb_branch"; fi
  env $env timeout 600 python bench.py --steps 6 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/r48_${mode}_${round}.json 2> gpurun_out/r48_${mode}_${round}.err
done
done
python - <<'PY'
import json
for mode in ('on','off'):
    for r in (1,2):
        try:
            d=json.loads(open(f'gpurun_out/r48_{mode}_{r}.json').read().strip().splitlines()[-1])
            print(mode, r, round(d['value'],3), d['phases_ms'], d.get('final_loss'))
        except Exception as e: print(mode, r, 'failed', e)
PY
