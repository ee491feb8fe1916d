#!/bin/bash
# Developer A/B of k_resize_tail (pyramid levels >= ORBX_TAIL_FROM as phases of one cluster launch; 0 = off, the default).
# usage: tools/tail_ab.sh C1 C4      prints the batch pyramid stage time and the single-frame p50
for cfg in "$@"; do
  for e in "ORBX_TAIL_FROM=0" "ORBX_TAIL_FROM=4" "ORBX_TAIL_FROM=4 ORBX_TAIL_BATCH=1 ORBX_TAIL_CLUSTER=1" "ORBX_TAIL_FROM=4 ORBX_TAIL_BATCH=1 ORBX_TAIL_CLUSTER=2" "ORBX_TAIL_FROM=4 ORBX_TAIL_BATCH=1 ORBX_TAIL_CLUSTER=8"; do
    env $e python bench.py --config $cfg --steps 5 --warmup 3 --no-cpu-baseline --no-extras --sustain-s 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
l=d['latency']; k=d['roofline']['kernels']
print('$cfg', '$e', round(d['value']), 'pyramid=%.3f' % k['pyramid']['ms_per_launch_batch'], 'p50_us=%.1f' % l['p50_us'])
"
  done
done
