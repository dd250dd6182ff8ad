#!/bin/bash
# A/B of several builds of the library on the same box: alternate short bench.py runs (SCG_LIB selects the .so).
# Usage (under gpurun): bash tools/ab_bench.sh <rounds> <lib.so>...
R=$1; shift
for r in $(seq 1 $R); do
  for L in "$@"; do
    SCG_LIB=$PWD/$L python bench.py --steps 100 --warmup 5 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$L', '%.4e' % d['value'], '%.4f ms' % d['ms_per_step'], 'frac %.3f' % d['roofline']['frac'])"
  done
done
