mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensor.py -x -q -m gpu 2>&1 | tail -30 > gpurun_out/t_tensor.log
tail -12 gpurun_out/t_tensor.log
for args in "" "--tc-stats 1" "--tc-mode 1"; do
  echo "== c5q $args"
  timeout 600 python bench.py --config c5q --no-cpu --no-e2e --steps 3 $args 2>>gpurun_out/b.err | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print(j['ms_per_step'], j['roofline']['frac'], j['stages_ms'], j['tc'], j['uncertified_per_step'], j.get('tc_stats'))"
done
timeout 900 python bench.py --steps 3 --no-cpu --no-e2e --tc-stats 1 2>> gpurun_out/b.err | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print(j['ms_per_step'], j['roofline']['frac'], j['stages_ms'], j['tc'], j['uncertified_per_step'], j.get('tc_stats'))"
timeout 900 python bench.py --steps 3 --cpu-seconds 5 > gpurun_out/b_c5.json 2>> gpurun_out/b.err; tail -3 gpurun_out/b.err; cat gpurun_out/b_c5.json
