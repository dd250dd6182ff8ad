#!/bin/bash
# A/B of library builds on one box: bash tools/ab_quick.sh <rounds> <lib.so|default>...   (headline only, 100 timed launches each)
R=$1; shift
for r in $(seq 1 $R); do
  for L in "$@"; do
    if [ "$L" = default ]; then unset SCG_LIB; else export SCG_LIB=$PWD/$L; fi
    python bench.py --steps 100 --warmup 5 --no-cpu --configs '' --sustain-seconds 0 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$L', '%.4e' % d['value'], '%.4f ms' % d['ms_per_step'], 'frac %.3f' % d['roofline']['frac'])"
  done
done
