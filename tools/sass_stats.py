"""SASS size / opcode histogram of the kernels in libscg_b200.so (no GPU needed).
    python tools/sass_stats.py [regex]         # list functions with instruction counts
    python tools/sass_stats.py -f <regex>      # opcode histogram + loop ranges of matching functions"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "safe_control_gym_b200", "libscg_b200.so")


def functions():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    name, body, res = None, [], {}
    for line in out.splitlines():
        m = re.match(r"\s+Function : (\S+)", line)
        if m:
            if name:
                res[name] = body
            name, body = m.group(1), []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", line)
        if m and name:
            body.append((int(m.group(1), 16), m.group(2).strip()))
    if name:
        res[name] = body
    return res


def demangle(n):
    return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip().split("(scg::Hot")[0]


if __name__ == "__main__":
    fs = functions()
    if len(sys.argv) > 2 and sys.argv[1] == "-f":
        pat = re.compile(sys.argv[2])
        for n, body in fs.items():
            d = demangle(n)
            if not pat.search(d):
                continue
            print(f"== {d}: {len(body)} instructions ({len(body) * 16 / 1024:.1f} KB)")
            ops = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0] for _, t in body)
            print("   " + ", ".join(f"{k} {v}" for k, v in ops.most_common(40)))
            # backward branches = loops
            for addr, t in body:
                m = re.search(r"BRA\S*\s+.*?(0x[0-9a-f]+)", t)
                if m and int(m.group(1), 16) < addr:
                    tgt = int(m.group(1), 16)
                    print(f"   loop {tgt:#x}..{addr:#x}: {(addr - tgt) // 16 + 1} instructions")
    else:
        pat = re.compile(sys.argv[1]) if len(sys.argv) > 1 else None
        rows = sorted(((len(b), demangle(n)) for n, b in fs.items()))
        for c, d in rows:
            if pat is None or pat.search(d):
                print(f"{c:7d}  {d}")
