#!/bin/bash
# ab_margin.sh <so> <config> "<margins>" [flags]
so=$1; cfg=$2
cp build_ab/$so.so util_b200/csrc/libb200approx.so
for mg in $3; do
  timeout 300 python bench.py --config $cfg --steps 3 --warmup 3 --no-cpu --no-e2e --margin $mg --tc-flags ${4:-0} 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        j = json.loads(l); print('$so $cfg margin $mg', 'stage1 %.2f' % j['stages_ms']['ms_stage1'], 'rerank %.2f' % j['stages_ms']['ms_rerank'], 'exact %.2f' % j['stages_ms']['ms_exact'], 'uncert', j['uncertified_per_step'], 'keep', j['tc']['keep'], 'cap', j['tc']['cap'], 'value %.0f' % j['value'])
"
done
