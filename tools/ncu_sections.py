"""Executed warp-instructions per loop of a kernel, from the source page of an ncu report (read here, no GPU).
Loops are found from backward branches; counts are exclusive of nested loops and normalised per warp-step.
    python tools/ncu_sections.py gpurun_out/prof_X.ncu-rep --steps 1000 [--warps 2048] [--dump 0x1b10:0x4b00]"""
import argparse
import csv
import io
import re
import subprocess

ap = argparse.ArgumentParser()
ap.add_argument("rep")
ap.add_argument("--steps", type=int, required=True, help="control steps per launch (K)")
ap.add_argument("--warps", type=int, default=2048)
ap.add_argument("--dump", default="", help="lo:hi offsets to list instruction by instruction")
ap.add_argument("--min-exec", type=float, default=0.0, help="with --dump: only instructions executed at least this often per warp-step")
a = ap.parse_args()
src = subprocess.run(["ncu", "-i", a.rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
start = next(i for i, x in enumerate(rows) if x and x[0] == "Address")
hd = rows[start]
data = [x for x in rows[start + 1:] if len(x) == len(hd)]
ia, isrc, isamp, iex, ith = (hd.index(k) for k in ("Address", "Source", "# Samples", "Instructions Executed", "Avg. Threads Executed"))
base = int(data[0][ia], 16)
ins = [(int(x[ia], 16) - base, x[isrc].strip(), int(x[iex]), int(x[isamp]), float(x[ith])) for x in data]
norm = float(a.steps * a.warps)
loops = []
for off, text, ex, _, _ in ins:
    m = re.search(r"\bBRA(?:\.\w+)*\s+(?:\S+,\s*)?0x([0-9a-f]+)", text)
    if m:
        tgt = int(m.group(1), 16) - base
        if 0 <= tgt <= off:
            loops.append((tgt, off))
loops = sorted(set(loops), key=lambda l: (l[1] - l[0]))
print(f"total executed per warp-step: {sum(i[2] for i in ins) / norm:.1f}   ({len(ins)} static instructions)")
claimed = [False] * len(ins)
offs = [i[0] for i in ins]
for lo, hi in loops:
    tot = excl = stat = 0
    for j, (off, text, ex, _, _) in enumerate(ins):
        if lo <= off <= hi:
            tot += ex
            if not claimed[j]:
                excl += ex
                stat += 1
                claimed[j] = True
    if tot / norm >= 1.0:
        print(f"  loop {lo:#7x}..{hi:#7x}: {tot / norm:8.1f} per warp-step inclusive, {excl / norm:8.1f} exclusive ({stat} static)")
if a.dump:
    lo, hi = (int(v, 16) for v in a.dump.split(":"))
    for off, text, ex, samp, th in ins:
        if lo <= off <= hi and ex / norm >= a.min_exec:
            print(f"+{off:#06x} {ex / norm:6.2f} {th:5.1f} s{samp:5d} {text[:90]}")
