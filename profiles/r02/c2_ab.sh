timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -q -m gpu -x -k "shepmod or c2" 2>&1 | tail -2
timeout 600 python profiles/microbench/fuzz_shepmod_filter.py 24 5 2>&1 | tail -1
for L in build_ab/libcur.so util_b200/csrc/libb200approx.so build_ab/libcur.so util_b200/csrc/libb200approx.so; do
B200_LIB=$PWD/$L timeout 300 python bench.py --config c2 --no-cpu 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('$L c2', round(j['value']), round(j['ms_per_step'],3), {k:round(v,2) for k,v in j['stages_ms'].items()}, round(j['e2e']['value']))"
done
