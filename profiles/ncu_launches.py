#!/usr/bin/env python
"""Summarises an `ncu --csv` launch list (gpu__time_duration.sum [+ dram bytes]) per launch: name, ms, GB read / written.
usage: ncu_launches.py file.csv [last_n]"""
import collections
import csv
import sys


def load(path):
    rows = list(csv.reader(open(path)))
    hdr, data = None, []
    for r in rows:
        if r and r[0] == "ID":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            data.append(dict(zip(hdr, r)))
    by = collections.OrderedDict()
    for d in data:
        e = by.setdefault(d["ID"], {"name": d["Kernel Name"]})
        v = float(d["Metric Value"].replace(",", ""))
        u = d["Metric Unit"]
        if d["Metric Name"] == "gpu__time_duration.sum":
            e["ms"] = v / {"ns": 1e6, "us": 1e3, "ms": 1.0, "s": 1e-3}.get(u, 1e6)
        else:
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
            e["rd" if "read" in d["Metric Name"] else "wr"] = v * scale / 1e9
    return list(by.values())


if __name__ == "__main__":
    ls = load(sys.argv[1])
    if len(sys.argv) > 2:
        ls = ls[-int(sys.argv[2]):]
    tot = 0.0
    for e in ls:
        tot += e.get("ms", 0)
        print(f"{e['name'][:70]:70s} {e.get('ms', 0):9.3f} ms  rd {e.get('rd', 0):8.3f} GB  wr {e.get('wr', 0):8.3f} GB")
    print(f"{'total':70s} {tot:9.3f} ms")
