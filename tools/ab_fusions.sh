# A/B of the host-level fusions on ONE box (SLIDER_DISABLE switches in unet.py / engine.py); phases_ms are graph-replay times
for v in ${AB_LIST:-"none residual qkv kvcache temb_stream gn_train ptrain"} ; do
  SLIDER_DISABLE=$v timeout 300 python bench.py --skip-cpu-baseline --steps 6 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('disable=[$v]', round(d['value'],3), round(d['ms_per_step'],2), 'ms/step', d['phases_ms'], d['roofline']['frac'])"
done
