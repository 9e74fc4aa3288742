#!/bin/bash
# ab_cfg_modes.sh <so> <config> "<flags>" "<modes>"
so=$1; cfg=$2
cp build_ab/$so.so util_b200/csrc/libb200approx.so
for fl in $3; do for md in $4; do
  timeout 300 python bench.py --config $cfg --steps 3 --warmup 3 --no-cpu --no-e2e --tc-flags $fl --tc-mode $md 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        j = json.loads(l); print('$so $cfg flags $fl mode $md', 'stage1 %.2f' % j['stages_ms']['ms_stage1'], 'frac %.3f' % j['roofline']['frac'], 'uncert', j['uncertified_per_step'], 'clk', j['clocks']['sm_mhz'], j['clocks']['reasons'])
"
done; done
