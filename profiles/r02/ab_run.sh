timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -q -m gpu -x 2>&1 | tail -2
for i in 1 2; do
timeout 300 python bench.py --config c4 --no-ref-class 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('c4', round(j['value']), round(j['ms_per_step'],2), {k:round(v,2) for k,v in j['stages_ms'].items()}, j['parity_on_sample'])"
done
