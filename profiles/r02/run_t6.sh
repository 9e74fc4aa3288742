mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -x -q -m gpu -k "shepmod or c2" 2>&1 | tail -25 | tee gpurun_out/t_shepmod.log
timeout 900 python bench.py --config c2 --steps 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err || tail -5 gpurun_out/bench_c2.err
python -c "
import json
j=json.load(open('gpurun_out/bench_c2.json'))
print('c2', round(j['value']), 'q/s dev', round(j['e2e']['value']), 'q/s e2e', j['stages_ms'], 'frac', round(j['roofline']['frac'],4), j['roofline']['kernel'], j['parity_on_sample'])"
