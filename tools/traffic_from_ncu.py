"""profiles/traffic.json from a full ncu capture of the bench-shaped rollout launch (read here, no GPU):
    python tools/traffic_from_ncu.py gpurun_out/prof_X.ncu-rep"""
import csv
import io
import json
import os
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
h, units, row = r[0], r[1], r[2]


def val(name):
    v, u = float(row[h.index(name)]), units[h.index(name)]
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)


out = {"source": os.path.basename(rep), "kernel": row[h.index("Kernel Name")],
       "k_env_rollout_dram_bytes_per_launch": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
       "dram_read_bytes": val("dram__bytes_read.sum"), "dram_write_bytes": val("dram__bytes_write.sum"),
       "duration_under_ncu": row[h.index("gpu__time_duration.sum")] + " " + units[h.index("gpu__time_duration.sum")]}
for key, name in (("k_env_rollout_fma_pipe_frac", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                  ("k_env_rollout_issue_active_frac", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                  ("k_env_rollout_alu_pipe_frac", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                  ("k_env_rollout_xu_pipe_frac", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active")):
    if name in h:
        out[key] = float(row[h.index(name)]) / 100.0
if "smsp__thread_inst_executed_per_inst_executed.ratio" in h:
    out["k_env_rollout_threads_per_instruction"] = float(row[h.index("smsp__thread_inst_executed_per_inst_executed.ratio")])
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
json.dump(out, open(os.path.join(root, "profiles", "traffic.json"), "w"), indent=1)
print(out)
