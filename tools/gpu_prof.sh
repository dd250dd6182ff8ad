#!/bin/bash
# quick sweep + one full ncu capture of the rollout kernel.  Usage: bash tools/gpu_prof.sh <tag> [profile_rollout args]
TAG=$1; shift
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python tools/quick_bench.py > $OUT/quick_$TAG.log 2>&1; echo "quick rc=$?"; cat $OUT/quick_$TAG.log
PROF="python tools/profile_rollout.py --k 100 --launches 3 $@"
timeout 300 $PROF > $OUT/plain_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_rollout -s 1 -c 1 -o $OUT/prof_$TAG -f $PROF > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; tail -3 $OUT/ncu_full_$TAG.log
