mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tensor.py tests/test_gpu_configs.py -x -q -m gpu -s 2>&1 | grep -E "uncertified|worst|passed|failed|Error|error" | tail -30
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
timeout 600 python bench.py --config c4 --no-ref-class > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err || tail -5 gpurun_out/bench_c4.err
python -c "
import json
j=json.load(open('gpurun_out/bench_c4.json'))
print('c4', round(j['value']), j['ms_per_step'], j['stages_ms'], j['parity_on_sample'])"
timeout 600 python profiles/microbench/path_sweep.py 2>&1 | grep "k= 64" | tail -20
