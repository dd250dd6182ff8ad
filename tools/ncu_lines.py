"""Where the executed instructions of a kernel go, by CUDA source line: joins the SASS page of an ncu report with the line table
of the object file it was built from (nvdisasm -g; build with -lineinfo).  No GPU needed.
    python tools/ncu_lines.py <rep> <object.o> [top=40] [warp_steps]
warp_steps = warps x control steps of the captured launch: adds instructions per warp-step."""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

rep, obj = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
wsteps = float(sys.argv[4]) if len(sys.argv) > 4 else 0.0
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = next(i for i, x in enumerate(rows) if x and x[0] == "Address")
h = rows[hdr]
data = [x for x in rows[hdr + 1:] if len(x) == len(h)]
ia, isrc, iex, isamp, ith = (h.index(k) for k in ("Address", "Source", "Instructions Executed", "# Samples", "Avg. Threads Executed"))
ops = [x[isrc].split()[0] if not x[isrc].startswith("@") else x[isrc].split()[1] for x in data]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
funcs, cur, line = {}, None, ("?", 0)
for l in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m:
        cur = m.group(1); funcs[cur] = []; line = ("?", 0); continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        line = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m and cur:
        txt = m.group(2).strip()
        op = txt.split()[1] if txt.startswith("@") else txt.split()[0]
        funcs[cur].append((op, line))
best = None
for name, ins in funcs.items():
    if len(ins) == len(ops) and all(a == b[0] for a, b in zip(ops[:200], ins[:200])):
        best = name
if best is None:
    cand = sorted(funcs, key=lambda n: abs(len(funcs[n]) - len(ops)))[:3]
    sys.exit(f"no function of {obj} matches the report's {len(ops)} instructions; closest: {[(c[:60], len(funcs[c])) for c in cand]}")
per, samp, thr = defaultdict(int), defaultdict(int), defaultdict(float)
tot = 0
for x, (_, ln) in zip(data, funcs[best]):
    ex = int(x[iex]); tot += ex
    per[ln] += ex; samp[ln] += int(x[isamp] or 0); thr[ln] += float(x[ith] or 0) * ex
ts = sum(samp.values()) or 1
print(f"{best[:110]}\n{len(ops)} SASS instructions, {tot} executed warp-instructions" + (f" = {tot / wsteps:.1f} per warp-step" if wsteps else ""))
for ln, ex in sorted(per.items(), key=lambda kv: -kv[1])[:top]:
    extra = f"  {ex / wsteps:7.1f}/step" if wsteps else ""
    print(f"  {ln[0]:18s}:{ln[1]:5d}  exec {100 * ex / tot:5.1f}%{extra}  samples {100 * samp[ln] / ts:5.1f}%  threads {thr[ln] / max(ex, 1):4.1f}")
