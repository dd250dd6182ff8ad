#!/bin/bash
# Per-chunk cost of the ticket queue in isolation: 75,776 envs = 1,184 env blocks = the resident slots, so a queued launch
# never waits for a predecessor and differs from the static grid only by the range prologue / epilogue / hand-over.
for n in 75776 65536; do for q in 0 1; do for c in 32 64 128; do
  [ $q = 0 ] && [ $c != 64 ] && continue
  SCG_QUEUE=$q SCG_QUEUE_CHUNK=$c python bench.py --envs $n --steps 30 --warmup 3 2>/dev/null > /tmp/q.json
  python - "$n" "$q" "$c" <<'PY'
import json, sys
d = json.load(open('/tmp/q.json'))
print("envs", sys.argv[1], "queue", sys.argv[2], "chunk", sys.argv[3], "value %.4e" % d["value"], "ms %.4f" % d["ms_per_step"], "frac %.4f" % d["roofline"]["frac"], flush=True)
PY
done; done; done
