"""Quick device-timed numbers of the bench configs (no CPU baselines, short): python tools/bench_quick.py [--configs 3,4,5,2h,2l,2c]"""
import json
import subprocess
import sys
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
cfgs = sys.argv[sys.argv.index("--configs") + 1] if "--configs" in sys.argv else "3,4,5,2h,2l,2c"
out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "50", "--no-cpu", "--sustain-seconds", "0.3", "--configs", cfgs],
                     capture_output=True, text=True)
try:
    d = json.loads(out.stdout)
except Exception:
    print(out.stdout[-2000:], out.stderr[-3000:]); sys.exit(1)
print("headline %.4e (frac %.3f)  e2e %.3e  closed-loop e2e %.3e" % (d["value"], d["roofline"]["frac"], d["e2e"]["value"], d["e2e"]["closed_loop"]["value"]))
for c in d["configs"]:
    print("  %-14s %.4e  %8.2f ms  %s frac %.3f %s" % (c["id"], c["value"], c["ms_per_launch"], c["roofline"]["bound"], c["roofline"]["frac"], c.get("log_store", "")))
