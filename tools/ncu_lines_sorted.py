"""Compact, source-ordered view of tools/ncu_lines.py: python tools/ncu_lines_sorted.py <rep> <object.o> <warp_steps> [min_per_step=2.5]"""
import re
import subprocess
import sys
import os
rep, obj, ws = sys.argv[1], sys.argv[2], sys.argv[3]
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 2.5
out = subprocess.run([sys.executable, os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_lines.py"), rep, obj, "100000", ws],
                     capture_output=True, text=True).stdout.splitlines()
print("\n".join(out[:2]))
rows = []
for l in out[2:]:
    m = re.match(r"\s*(\S+)\s*:\s*(\d+)\s+exec\s+([\d.]+)%\s+([\d.]+)/step\s+samples\s+([\d.]+)%\s+threads\s+([\d.]+)", l)
    if m:
        rows.append((m.group(1), int(m.group(2)), float(m.group(4)), float(m.group(6)), float(m.group(5))))
rows.sort()
print(" | ".join(f"{f.replace('scg_', '').replace('.cuh', '')}:{l} {v:.1f} t{t:.0f}" for f, l, v, t, _ in rows if v >= thr))
print("listed %.1f of %.1f per warp-step" % (sum(r[2] for r in rows if r[2] >= thr), sum(r[2] for r in rows)))
