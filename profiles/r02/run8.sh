mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensor.py -x -q -m gpu 2>&1 | tail -5
echo "== c5q"
timeout 600 python bench.py --config c5q --no-cpu --no-e2e --steps 3 2>>gpurun_out/b.err | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print(j['ms_per_step'], j['roofline']['frac'], j['stages_ms']['ms_stage1'], j['uncertified_per_step'])"
timeout 900 python bench.py --steps 3 --no-cpu 2>> gpurun_out/b.err | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print(j['ms_per_step'], j['roofline']['frac'], j['stages_ms'], j['e2e']['value'], j['uncertified_per_step'], j['clocks'])"
tail -3 gpurun_out/b.err
