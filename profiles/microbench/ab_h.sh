#!/bin/bash
# ab_h.sh <so> "<halves list>" "<flags list>" "<modes list>"
so=$1
cp build_ab/$so.so util_b200/csrc/libb200approx.so
for hv in $2; do for fl in $3; do for md in $4; do
  timeout 120 python bench.py --config c5q --steps 5 --warmup 3 --no-cpu --no-e2e --tc-halves $hv --tc-flags $fl --tc-mode $md 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        j = json.loads(l); print('$so halves $hv flags $fl mode $md', 'stage1 %.2f' % j['stages_ms']['ms_stage1'], 'frac %.3f' % j['roofline']['frac'], 'uncert', j['uncertified_per_step'], j['tc'])
"
done; done; done
