#!/bin/bash
# Round-end evidence: bash tools/capture_all.sh <tag> bench|A|B   (three gpurun calls: a call brings back at most 64 MiB)
#   bench: both bench arms, the ncu launch list of a short bench run, one full ncu capture of the headline launch;
#   A / B: one full ncu capture per other config (each after the same command ran clean).
TAG=$1; PART=${2:-bench}; OUT=gpurun_out; mkdir -p $OUT
if [ $PART = A ]; then for C in 3 4 5 5g; do bash tools/prof_config.sh ${TAG}_$C $C 64; done; exit 0; fi
if [ $PART = B ]; then for C in 2l pid step; do bash tools/prof_config.sh ${TAG}_$C $C 64; done; exit 0; fi
python bench.py --impl reference > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "ref arm rc=$?"
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu --sustain-seconds 0.2 --config-scale 0.25"
timeout 600 $SHORT > $OUT/${TAG}_short.json 2> $OUT/${TAG}_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches.csv $SHORT > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
HEAD="python tools/profile_rollout.py --k 1000 --launches 3"
timeout 300 $HEAD > $OUT/plain_${TAG}_2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_rollout -s 1 -c 1 -o $OUT/prof_${TAG}_2 -f $HEAD > $OUT/ncu_${TAG}_2.log 2>&1
echo "headline ncu rc=$?"
