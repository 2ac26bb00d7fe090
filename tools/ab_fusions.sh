# A/B of the host-level fusions on ONE box (SLIDER_DISABLE switches in unet.py); phases_ms are graph-replay times
for v in ${AB_LIST:-"none residual qkv kvcache residual,qkv,kvcache"} ; do
  SLIDER_DISABLE=$v timeout 300 python bench.py --skip-cpu-baseline --steps 6 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); p=d['phases_ms']; print('disable=[$v]', round(d['value'],3), p, 'denoise/train', round(p['denoise']/p['train'],4), d['roofline']['frac'])"
done
