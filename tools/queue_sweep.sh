#!/bin/bash
# Sweep of the rollout ticket-queue chunk length on the bench workload (0 = static grid).  Usage: bash tools/queue_sweep.sh [chunks...]
for c in ${@:-0 32 64 96 128 250}; do
  if [ "$c" = 0 ]; then export SCG_QUEUE=0; else export SCG_QUEUE=-1 SCG_QUEUE_CHUNK=$c; fi
  python bench.py --steps 50 --warmup 3 2>/dev/null > /tmp/q.json
  python - "$c" <<'PY'
import json, sys
d = json.load(open('/tmp/q.json'))
print("chunk", sys.argv[1], "value %.4e" % d["value"], "ms %.4f" % d["ms_per_step"], "frac %.4f" % d["roofline"]["frac"], "e2e %.3e" % d["e2e"]["value"],
      "episodes", d["episode_stats"]["episodes"], "return", d["episode_stats"]["return"], flush=True)
PY
done
