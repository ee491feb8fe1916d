#!/bin/bash
# Developer A/B: quadtree CTA size (ORBX_OCT_THREADS=128|256|512, "auto" = the library's rule) per BASELINE config;
# prints the batch stage times and the single-frame p50.   usage: tools/oct_ab.sh C1 C4 C5
for cfg in "$@"; do
  for t in auto 128 256 512; do
    if [ $t = auto ]; then unset ORBX_OCT_THREADS; else export ORBX_OCT_THREADS=$t; fi
    python bench.py --config $cfg --steps 5 --warmup 3 --no-cpu-baseline --no-extras --sustain-s 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
k=d['roofline']['kernels']
print('$cfg', 'threads=$t', round(d['value']), ' '.join(f\"{n}={v['ms_per_launch_batch']:.3f}\" for n,v in k.items()), 'p50_us=%.1f' % d['latency']['p50_us'])
"
  done
done
