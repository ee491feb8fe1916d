#!/bin/bash
# Developer A/B harness: runs bench.py once per kernel-variant library in _variants/ (built with
# `make -C orb-slam3-mac_b200/csrc OUT=... OBJDIR=... EXTRA=-D...`) and prints the stage times.
shopt -s nullglob
for lib in "" _variants/*.so; do
  if [ -n "$lib" ]; then export ORBX_B200_LIB=$PWD/$lib; else unset ORBX_B200_LIB; fi
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
k=d['roofline']['kernels']
print('${lib:-default}', round(d['value']), ' '.join(f\"{n}={v['ms_per_launch_batch']:.3f}\" for n,v in k.items()))
"
done
