mkdir -p gpurun_out
for args in "--tc-mode 2" "--tc-mode 2 --tc-parts 2" "--tc-mode 2 --tc-parts 4" "--tc-mode 1 --tc-parts 4" "--tc-parts 4"; do
  echo "== c5q $args"
  timeout 600 python bench.py --config c5q --no-cpu --no-e2e --steps 3 $args 2>>gpurun_out/b.err | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print(j['ms_per_step'], j['roofline']['frac'], j['stages_ms']['ms_stage1'], j['uncertified_per_step'])"
done
tail -3 gpurun_out/b.err
