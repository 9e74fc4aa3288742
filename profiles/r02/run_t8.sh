mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tensor.py -x -q -m gpu 2>&1 | tail -3
python profiles/microbench/tc_stats.py 1000000 37888 64 16
python profiles/microbench/tc_stats.py 100000 100000 16 18
timeout 600 python bench.py --no-cpu --steps 3 2>>gpurun_out/b.err | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print(j['ms_per_step'], j['roofline']['frac'], j['stages_ms'], j['e2e']['value'], j['uncertified_per_step'], j['clocks'], j['parity_on_sample'])"
tail -3 gpurun_out/b.err
