timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -q -m gpu -x 2>&1 | tail -2
timeout 300 python bench.py --config c4 --no-ref-class 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('c4', round(j['value']), round(j['ms_per_step'],2), {k:round(v,2) for k,v in j['stages_ms'].items()}, j['parity_on_sample'])"
for mg in 8 4 6 8 5; do
timeout 300 python bench.py --no-cpu --no-e2e --steps 3 --margin $mg 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('margin $mg', round(j['ms_per_step'],2), {k:round(v,2) for k,v in j['stages_ms'].items()}, j['uncertified_per_step'], j['tc']['keep'])"
done
