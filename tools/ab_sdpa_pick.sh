# A/B of the per-shape SDPA backend choice vs torch default on ONE box
run() { env "$@" timeout 300 python bench.py --skip-cpu-baseline --steps 6 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', round(d['value'],3), round(d['ms_per_step'],2), 'ms/step', d['phases_ms'], 'e2e', round(d['e2e']['value'],3))"; }
run SLIDER_DISABLE=sdpa_pick
run SLIDER_DISABLE=
run SLIDER_DISABLE=sdpa_pick
run SLIDER_DISABLE=
