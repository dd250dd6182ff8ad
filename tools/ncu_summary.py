"""Summarise an ncu report (read here, no GPU): key raw metrics, stall reasons, and executed-instruction /
sample shares per 128-instruction bucket of the kernel.   python tools/ncu_summary.py gpurun_out/prof_X.ncu-rep [launch_idx]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
h, row = r[0], r[2 + idx]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.max", "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "sm__sass_thread_inst_executed_op_ffma_pred_on.sum", "sm__sass_thread_inst_executed_op_fadd_pred_on.sum",
        "sm__sass_thread_inst_executed_op_fmul_pred_on.sum", "smsp__sass_thread_inst_executed_op_fp32_pred_on.sum"]
print("kernel:", row[h.index("Kernel Name")][:100])
for w in want:
    if w in h:
        print(f"  {w:75s} {row[h.index(w)]} {r[1][h.index(w)]}")
st = []
for i, name in enumerate(h):
    if name.startswith("smsp__pcsamp_warps_issue_stalled_") and not name.endswith("_not_issued"):
        try:
            st.append((float(row[i]), name.replace("smsp__pcsamp_warps_issue_stalled_", "")))
        except ValueError:
            pass
tot = sum(v for v, _ in st) or 1
print("  stalls: " + ", ".join(f"{n} {100 * v / tot:.1f}%" for v, n in sorted(st, reverse=True)[:9]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
starts = [i for i, x in enumerate(rows) if x and x[0] == "Address"]
hd = rows[starts[idx]]
end = starts[idx + 1] - 1 if idx + 1 < len(starts) else len(rows)
data = [x for x in rows[starts[idx] + 1:end] if len(x) == len(hd)]
ia, isrc, isamp, iex, ith = (hd.index(k) for k in ("Address", "Source", "# Samples", "Instructions Executed", "Avg. Threads Executed"))
base = int(data[0][ia], 16)
tex = sum(int(x[iex]) for x in data) or 1
ts = sum(int(x[isamp]) for x in data) or 1
print(f"  source page: {len(data)} instructions, {tex} executed, {ts} samples")
B = 128
for b in range(0, len(data), B):
    ch = data[b:b + B]
    ex = sum(int(x[iex]) for x in ch)
    s = sum(int(x[isamp]) for x in ch)
    th = sum(float(x[ith]) * int(x[iex]) for x in ch) / max(ex, 1)
    if ex > tex * 0.005 or s > ts * 0.005:
        print(f"   +{int(ch[0][ia], 16) - base:#7x}: exec {100 * ex / tex:5.1f}%  samples {100 * s / ts:5.1f}%  threads {th:5.1f}  {ch[0][isrc].strip()[:50]}")
