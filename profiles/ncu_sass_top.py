"""Top stall sites of one kernel from `ncu -i X.ncu-rep --page source --csv --print-source sass`, plus headline metrics
from `--page raw --csv`.  usage: python profiles/ncu_sass_top.py X.ncu-rep [top_n] [kernel-name regex]"""
import csv, subprocess, sys, io

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
pick = ["-k", "regex:" + sys.argv[3], "-c", "1"] if len(sys.argv) > 3 else []
raw = subprocess.run(["ncu", "-i", rep] + pick + ["--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, v = rows[0], rows[-1]
d = dict(zip(h, v))
print(d.get("Kernel Name", "")[:100])
for k in h:
    if "issue_stalled" in k and k.endswith("per_issue_active.ratio"):
        try:
            if float(d[k].replace(",", "")) > 0.3:
                print(" ", k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), d[k])
        except ValueError:
            pass
for k in ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
          "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
          "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
          "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active"]:
    if k in d:
        print(" ", k, d[k])
src = subprocess.run(["ncu", "-i", rep] + pick + ["--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h, body = rows[1], rows[2:]
col = {n: h.index(n) for n in ["Source", "# Samples", "Instructions Executed", "stall_short_sb", "stall_barrier", "stall_mio", "stall_long_sb",
                               "stall_wait", "stall_math", "L1 Wavefronts Shared", "L1 Wavefronts Shared Excessive"]}
tot = sum(int(r[col["# Samples"]]) for r in body)
print("total samples", tot, "instructions", len(body))
agg = {}
for r in body:
    for n in ["stall_short_sb", "stall_barrier", "stall_mio", "stall_long_sb", "stall_wait", "stall_math"]:
        agg[n] = agg.get(n, 0) + int(r[col[n]])
print(" ", {k: round(v / tot, 3) for k, v in agg.items()})
top = sorted(range(len(body)), key=lambda i: -int(body[i][col["# Samples"]]))[:top_n]
for i in sorted(top):
    r = body[i]
    print(f"{i:5d} {r[col['Source']].strip()[:56]:56s} {r[col['# Samples']]:>6s} sh {r[col['stall_short_sb']]:>5s} bar {r[col['stall_barrier']]:>5s} "
          f"mio {r[col['stall_mio']]:>5s} long {r[col['stall_long_sb']]:>5s} wf {r[col['L1 Wavefronts Shared']]:>8s} exc {r[col['L1 Wavefronts Shared Excessive']]:>8s}")
