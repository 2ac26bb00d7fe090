# A/B of the GroupNorm forms on ONE box: per-(image, group set) single launch (default) vs two-kernel form
run() { env "$@" timeout 300 python bench.py --skip-cpu-baseline --steps 6 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', round(d['value'],3), round(d['ms_per_step'],2), 'ms/step', d['phases_ms'], 'e2e', round(d['e2e']['value'],3))"; }
run SLIDER_GN_GROUP=0
run SLIDER_GN_GROUP=1
run SLIDER_GN_GROUP=1 SLIDER_GN_GROUP_KB=48
run SLIDER_GN_GROUP=1 SLIDER_GN_GROUP_KB=12
run SLIDER_GN_GROUP=0
run SLIDER_GN_GROUP=1
