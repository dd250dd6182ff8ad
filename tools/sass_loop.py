"""Print the SASS of one kernel of libscg_b200.so between two addresses (no GPU needed).
    python tools/sass_loop.py '<demangled-name regex>' 0x1090 0x15c0"""
import re, subprocess, sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sass_stats
pat = re.compile(sys.argv[1]); lo = int(sys.argv[2], 16); hi = int(sys.argv[3], 16)
for n, body in sass_stats.functions().items():
    if pat.search(sass_stats.demangle(n)):
        for a, t in body:
            if lo <= a <= hi:
                print(f"{a:6x} {t}")
        break
