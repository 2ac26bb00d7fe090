#!/bin/bash
# A/B (interleaved, two rounds): frozen nn.Linear (+ residual) through the fused kernel's plain path vs cuBLAS (+ add kernel)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for round in 1 2; do
  for mode in default plainres plain; do
    if [ "$mode" = default ]; then env=""; else env="SLIDER_ENABLE=$mode"; fi
    env $env timeout 600 python bench.py --steps 6 --warmup 3 --skip-cpu-baseline --skip-eager-baseline > gpurun_out/r34_${mode}_${round}.json 2> gpurun_out/r34_${mode}_${round}.err
  done
done
python - <<'PY'
import json
for mode in ('default','plainres','plain'):
    vals=[]
    for r in (1,2):
        try:
            d=json.loads(open(f'gpurun_out/r34_{mode}_{r}.json').read().strip().splitlines()[-1])
            vals.append((round(d['value'],3), d['phases_ms']['denoise'], d['phases_ms']['predict_train']))
        except Exception as e: vals.append(('failed',str(e)[:40]))
    print(mode, vals)
PY
