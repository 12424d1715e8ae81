"""Copies one run of profiles/experiments/r02_evidence.sh (gpurun_out/ev/) into the tracked files of this directory:
bench lines, the launch list, the ncu --set full summaries and traffic.json (DRAM bytes per launch / per C4 step).
    python profiles/refresh_from_evidence.py [gpurun_out/ev]"""
import collections
import csv
import json
import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ev = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(HERE), "gpurun_out", "ev")
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3,
         "usecond": 1e-3, "msecond": 1, "nsecond": 1e-6, "second": 1e3}

shutil.copy(os.path.join(ev, "bench_n1.json"), os.path.join(HERE, "r02_bench_n1.json"))
shutil.copy(os.path.join(ev, "bench_reference.json"), os.path.join(HERE, "r02_bench_reference.json"))
shutil.copy(os.path.join(ev, "launches_default_bench.csv"), os.path.join(HERE, "r02_launches_default_bench.csv"))
for w in ("c2", "c3", "c4", "c5"):
    out = subprocess.run([sys.executable, os.path.join(HERE, "ncu_summary.py"), os.path.join(ev, "full_%s.ncu-rep" % w)],
                         capture_output=True, text=True, check=True).stdout
    open(os.path.join(HERE, "r02_%s_ncu_full.txt" % w), "w").write(out)


def grab(path):
    txt = open(path).read()
    total = 0.0
    for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        m = re.search(name + r" ([\d.]+) (\w+)", txt)
        total += float(m.group(1)) * SCALE[m.group(2)]
    return int(total)


rows = list(csv.reader(open(os.path.join(HERE, "r02_launches_default_bench.csv"))))
start = [k for k, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[start]
col = {n: h.index(n) for n in ("ID", "Kernel Name", "Metric Name", "Metric Unit", "Metric Value")}
launches = collections.OrderedDict()
for r in rows[start + 1:]:
    if len(r) <= col["Metric Value"]:
        continue
    d = launches.setdefault(int(r[col["ID"]]), {"name": r[col["Kernel Name"]]})
    d[r[col["Metric Name"]]] = float(r[col["Metric Value"]].replace(",", "")) * SCALE[r[col["Metric Unit"]]]
seq = list(launches.values())
# a C4 step at 5e8 rows = [fill, insert (> 1 ms) ...] x 2 ranges + the merge; steps start at every other big insert
big = [k for k, d in enumerate(seq) if "agg_insert" in d["name"] and d["gpu__time_duration.sum"] > 1.0]
assert len(big) >= 4 and len(big) % 2 == 0, big
step_len = big[2] - big[0]
first = big[-2] - 1
assert "fill_u64" in seq[first]["name"], seq[first]["name"]
step = seq[first:first + step_len]
c4_bytes = int(sum(d["dram__bytes_read.sum"] + d["dram__bytes_write.sum"] for d in step))
c4_ms = sum(d["gpu__time_duration.sum"] for d in step)

path = os.path.join(HERE, "traffic.json")
t = json.load(open(path))
for w in ("c2", "c3", "c5"):
    t[w]["bytes"] = grab(os.path.join(HERE, "r02_%s_ncu_full.txt" % w))
t["c4"]["bytes"] = c4_bytes
json.dump(t, open(path, "w"), indent=1)
print("c4 step: %d launches, %.2f ms under ncu, %.2f GB of DRAM traffic" % (len(step), c4_ms, c4_bytes / 1e9))
print({w: t[w]["bytes"] for w in t if w != "_comment"})
