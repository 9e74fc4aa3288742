mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "any_weighted or auto" 2>&1 | tail -4
for c in c1 c2 c3 c4; do
  timeout 900 python bench.py --config $c --steps 3 --ref-seconds 5 > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err || tail -5 gpurun_out/bench_$c.err
  python -c "
import json,sys
j=json.load(open('gpurun_out/bench_$c.json'))
print('$c', round(j['value']), 'q/s dev', round(j['e2e']['value']), 'q/s e2e fit_s', round(j['fit_s'],4), j['stages_ms'], 'frac', round(j['roofline']['frac'],4), j['roofline']['kernel'], 'cpu', round(j['cpu_baseline']['value']), j['parity_on_sample'], j.get('reference_class',{}).get('value'), j['roofline'].get('fp64_fma_peak_tflops'), j['roofline'].get('fp32_fma_peak_tflops'))"
done
