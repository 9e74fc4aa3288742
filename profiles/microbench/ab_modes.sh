#!/bin/bash
# ab_modes.sh <so> "<flags list>" "<modes list>": c5q stage-1 time per (tc-flags, tc-mode)
so=$1
cp build_ab/$so.so util_b200/csrc/libb200approx.so
for fl in $2; do for md in $3; do
  timeout 120 python bench.py --config c5q --steps 5 --warmup 3 --no-cpu --no-e2e --tc-flags $fl --tc-mode $md 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        j = json.loads(l); print('$so flags $fl mode $md', 'stage1 %.2f' % j['stages_ms']['ms_stage1'], 'frac %.3f' % j['roofline']['frac'], 'uncert', j['uncertified_per_step'])
"
done; done
