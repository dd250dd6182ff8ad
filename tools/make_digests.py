"""profiles/<tag>_ncu_digests.md from the full ncu captures of tools/capture_all.sh (read here, no GPU):
    python tools/make_digests.py r2z
Per config: kernel, duration, registers, instructions per warp-step, issue / pipe utilisation, threads per instruction,
DRAM bytes, stall mix, and the source lines that cost the most instructions (joined with the line table of the build)."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
# config -> (what, object file, warp-steps of the captured launch)
CFG = [("2", "headline: Quadrotor 3D level0, 65,536 envs x 1,000 steps (bench launch)", "scg_k_3d_f.o", 2048 * 1000),
       ("2l", "level2-like randomisation, 65,536 envs x 64 steps", "scg_k_3d_f.o", 2048 * 64),
       ("pid", "closed loop (on-device PID), 65,536 envs x 64 steps", "scg_k_3d_f.o", 2048 * 64),
       ("step", "host-driven scg_step, 65,536 envs x 1 step", "scg_k_3d_f.o", 2048),
       ("3", "BASELINE config 3: cart-pole, 1,048,576 envs x 64 steps", "scg_k_cp_f.o", 32768 * 64),
       ("4", "BASELINE config 4: Quadrotor 2D, 262,144 envs x 64 steps", "scg_k_2d_f.o", 8192 * 64),
       ("5", "BASELINE config 5, gates off: Quadrotor 3D logged, 1,048,576 envs x 64 steps", "scg_k_3d_f_tlog.o", 32768 * 64),
       ("5g", "BASELINE config 5, gates on: Quadrotor 3D logged + gate geometry, 1,048,576 envs x 64 steps", "scg_k_3d_f_tlog.o", 32768 * 64)]
out = ["# ncu digests, captures `%s` (tools/capture_all.sh; `ncu --set full --clock-control none --import-source on`, one launch each)" % tag, "",
       "Instruction counts are warp-level (`smsp__inst_executed.sum`) divided by the warp-steps of the captured launch.", ""]
for cid, what, obj, ws in CFG:
    rep = os.path.join(ROOT, "gpurun_out", "prof_%s_%s.ncu-rep" % (tag, cid))
    if not os.path.exists(rep):
        continue
    summ = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), rep], capture_output=True, text=True).stdout.splitlines()
    keep = [l for l in summ if l.startswith("kernel:") or l.startswith("  ") and not l.lstrip().startswith("+")]
    inst = [float(l.split()[-2]) for l in summ if "smsp__inst_executed.sum" in l]
    lines = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), rep, os.path.join(ROOT, "safe_control_gym_b200", "build", obj),
                            "100000", str(ws)], capture_output=True, text=True).stdout.splitlines()
    rows = []
    for l in lines:
        m = re.match(r"\s*(\S+)\s*:\s*(\d+)\s+exec\s+([\d.]+)%\s+([\d.]+)/step\s+samples\s+([\d.]+)%\s+threads\s+([\d.]+)", l)
        if m:
            rows.append((float(m.group(4)), m.group(1), int(m.group(2)), float(m.group(5)), float(m.group(6))))
    rows.sort(reverse=True)
    out += ["## %s -- %s" % (cid, what), "", "```"] + keep
    if inst:
        out.append("  instructions per warp-step                                                  %.1f" % (inst[0] / ws))
    out += ["```", "", "Most expensive source lines (instructions per warp-step, share of stall samples, active threads):", "",
            "| file:line | instr / warp-step | samples | threads |", "|---|---|---|---|"]
    out += ["| %s:%d | %.1f | %.1f %% | %.0f |" % (f, ln, v, smp, t) for v, f, ln, smp, t in rows[:18]]
    out.append("")
open(os.path.join(ROOT, "profiles", "%s_ncu_digests.md" % tag), "w").write("\n".join(out))
print("\n".join(out[:60]))
