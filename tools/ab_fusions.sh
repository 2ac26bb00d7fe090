for v in "" residual qkv "residual,qkv" ; do
  SLIDER_DISABLE=$v timeout 300 python bench.py --skip-cpu-baseline --steps 6 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('disable=[$v]', round(d['value'],3), d['phases_ms'], d['roofline']['frac'])"
done
