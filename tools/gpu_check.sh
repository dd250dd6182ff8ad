#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench (both arms), ncu launch list + one full capture.
# Usage (from the repo root, under gpurun): bash tools/gpu_check.sh [tag] [launches|full]   (one ncu pass per box call)
TAG=${1:-r1}
NCU=${2:-launches}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > $OUT/clocks_$TAG.csv &
SMI=$!
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke_$TAG.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "bench ref rc=$?"
timeout 600 python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"
kill $SMI
PROF="python tools/profile_rollout.py --attitude fast_intrinsic --k 1000 --launches 3 --steps-api 20"
[ "$NCU" = launches ] && timeout 300 $PROF > $OUT/plain_$TAG.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/launches_$TAG.csv $PROF > $OUT/ncu_launches_$TAG.log 2>&1
echo "ncu launches rc=$?"
[ "$NCU" = full ] && timeout 300 $PROF > $OUT/plain2_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_rollout -s 1 -c 1 -o $OUT/prof_$TAG -f $PROF > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
tail -3 $OUT/pytest_gpu_$TAG.log; tail -2 $OUT/smoke_$TAG.log; cat $OUT/bench_ref_$TAG.json $OUT/bench_$TAG.json
