#!/bin/bash
# One full ncu capture of the rollout kernel of a config: bash tools/prof_config.sh <tag> <config> [k] [kernel regex]
TAG=$1; CFG=$2; K=${3:-64}; KRN=${4:-k_rollout}
OUT=gpurun_out; mkdir -p $OUT
CMD="python tools/profile_config.py --config $CFG --k $K --launches 3"
timeout 300 $CMD > $OUT/plain_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRN -s 1 -c 1 -o $OUT/prof_$TAG -f $CMD > $OUT/ncu_$TAG.log 2>&1
echo "ncu rc=$?"; tail -2 $OUT/plain_$TAG.log
