#!/bin/bash
# A/B of library builds on the same box: ab.sh <config> <so> [<so> ...]  (each .so from build_ab/)
cfg=$1; shift
for rep in 1 2; do
for so in "$@"; do
  cp build_ab/$so.so util_b200/csrc/libb200approx.so
  timeout 300 python bench.py --config $cfg --steps 5 --warmup 3 --no-cpu --no-e2e $AB_ARGS 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        j = json.loads(l); print('$so', '$cfg', 'stage1 %.2f' % j['stages_ms']['ms_stage1'], 'frac %.3f' % j['roofline']['frac'], 'uncert', j['uncertified_per_step'], 'exact %.2f' % j['stages_ms']['ms_exact'], 'value %.0f' % j['value'])
"
done
done
