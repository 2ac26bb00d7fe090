# A/B of the routing of frozen un-wrapped projections (unet.fused_linear) on ONE box
run() { env "$@" timeout 300 python bench.py --skip-cpu-baseline --steps 6 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', round(d['value'],3), round(d['ms_per_step'],2), 'ms/step', d['phases_ms'], 'e2e', round(d['e2e']['value'],3))"; }
run SLIDER_DISABLE=plainconv
run SLIDER_DISABLE=
run SLIDER_ENABLE=plainres
run SLIDER_ENABLE=plain
run SLIDER_DISABLE=plainconv
run SLIDER_DISABLE=
