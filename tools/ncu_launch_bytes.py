#!/usr/bin/env python
"""Per kernel: device time and DRAM bytes per build from an ncu CSV with gpu__time_duration.sum and
dram__bytes_{read,write}.sum (tools/build_bench.py runs 2 builds: initial + 1 rebuild).  Usage: ncu_launch_bytes.py csv [builds]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
builds = float(sys.argv[2]) if len(sys.argv) > 2 else 2.0
h = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
H = rows[h]
ki, mi, vi, ui = H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[h + 1:]:
    name = r[ki].split("(")[0].replace("void ", "")
    if "vkrt" not in name:
        continue
    name = name.split("::")[-1]
    v = float(r[vi].replace(",", ""))
    a = agg.setdefault(name, {"n": 0, "t": 0.0, "b": 0.0})
    if r[mi] == "gpu__time_duration.sum":
        a["n"] += 1
        a["t"] += v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
    else:
        a["b"] += v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[r[ui]]
tot = sum(a["t"] for a in agg.values())
print(f"# {tot / builds:.1f} us of kernels per build (ncu: serialised, cold caches)")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["t"]):
    print(f"{a['t'] / builds:9.1f} us/build {a['n']:4d} launches  dram {a['b'] / builds / 1e9:6.2f} GB/build  -> {a['b'] / 1e3 / max(a['t'], 1e-9):7.0f} GB/s  {k}")
