# Isolation modes of knn_tc_kernel (instrumented build): 2 = TMA + MMA only, 1 = + TMEM reads and min trees (no
# selection), 0 = full kernel.  c5q = one 256-query block per SM (37 888 x 1 M x 64), c5 = the full step.
for cfg in c5q c5; do for md in 2 1 0; do
timeout 600 python bench.py --config $cfg --no-cpu --no-e2e --steps 3 --tc-mode $md 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('$cfg mode $md (instrumented build): stage1 %.3f ms  frac of sustained peak %.3f  sm_mhz %s %s' % (j['stages_ms']['ms_stage1'], j['roofline']['frac'], j['clocks']['sm_mhz'], j['clocks']['reasons']))"
done; done
for cfg in c5q c5; do
timeout 600 python bench.py --config $cfg --no-cpu --no-e2e --steps 3 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    j=json.loads(l); print('$cfg production build: stage1 %.3f ms  frac of sustained peak %.3f  sm_mhz %s %s' % (j['stages_ms']['ms_stage1'], j['roofline']['frac'], j['clocks']['sm_mhz'], j['clocks']['reasons']))"
done
