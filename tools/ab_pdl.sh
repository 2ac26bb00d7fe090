# A/B of programmatic dependent launch on ONE box
for v in 1 0 1 0 ; do
  SLIDER_PDL=$v timeout 300 python bench.py --skip-cpu-baseline --steps 6 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); p=d['phases_ms']; print('SLIDER_PDL=$v', round(d['value'],3), p, d['roofline']['frac'], d['final_loss'])"
done
