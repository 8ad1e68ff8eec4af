#!/usr/bin/env python
"""Turns an ncu CSV of the trace kernels of ONE bench.py frame into profiles/rNN_traffic.json, the file bench.py's
roofline block reads (`traffic`, `dram_frac`, `issue_frac`).  The JSON names the git commit and a hash of the library
sources it was captured on; bench.py prints `traffic: null` when the hash of the sources it runs differs.

Capture (one GPU, after the same command has exited 0 without ncu):
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum \
      --clock-control none -k regex:'k_trace_(any|closest)' --launch-skip 4 --launch-count 4 --csv --log-file gpurun_out/traffic.csv \
      python bench.py --workload config3 --steps 1 --warmup 3 --no-cpu-baseline
(the first 4 matching launches are the counting pass; the next 4 are one warm-up frame: closest d0, any d0, closest d1, any d1)
Usage: ncu_traffic.py traffic.csv config3 profiles/r02_traffic.json
"""
import csv
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def csrc_sha():
    d = os.path.join(ROOT, "vkrt_b200", "csrc")
    h = hashlib.sha256()
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh", ".hpp")) or name == "Makefile":
            h.update(name.encode())
            h.update(open(os.path.join(d, name), "rb").read())
    return h.hexdigest()[:16]


def main():
    path, workload, out = sys.argv[1], sys.argv[2], sys.argv[3]
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 5]
    h = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    H = rows[h]
    ii, ki, mi, vi, ui = H.index("ID"), H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value"), H.index("Metric Unit")
    launches = {}
    for r in rows[h + 1:]:
        e = launches.setdefault(int(r[ii]), {"kernel": r[ki].split("(")[0].replace("void ", "").split("::")[-1]})
        v = float(r[vi].replace(",", ""))
        u = r[ui]
        if r[mi] == "gpu__time_duration.sum":
            e["gpu_time_ms"] = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
        elif r[mi].startswith("dram__bytes"):
            e[r[mi].split(".")[0].replace("dram__bytes_", "dram_") + "_bytes"] = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        elif r[mi] == "smsp__inst_executed.sum":
            e["warp_inst"] = v
        elif r[mi] == "smsp__thread_inst_executed.sum":
            e["thread_inst"] = v
    order = [launches[k] for k in sorted(launches)]
    names = []
    seen = {"any": 0, "closest": 0}
    for e in order:
        kind = "any" if "k_trace_any" in e["kernel"] else "closest"
        names.append(f"{kind}_d{seen[kind]}")
        seen[kind] += 1
    try:
        commit = subprocess.check_output(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], text=True).strip()
    except Exception:
        commit = os.environ.get("VKRT_COMMIT", "unknown")
    doc = json.load(open(out)) if os.path.exists(out) else {}
    doc["_comment"] = ("per launch of the trace kernels of one frame: ncu gpu__time_duration, dram__bytes_read+write, smsp__inst_executed "
                       "(warp instructions), smsp__thread_inst_executed; --clock-control none; times are cold-cache and serialised, "
                       "bench.py divides the byte and instruction counts by its own live launch times")
    doc[workload] = {"commit": commit, "csrc_sha": csrc_sha(), "launches": dict(zip(names, order))}
    json.dump(doc, open(out, "w"), indent=1)
    print(json.dumps(doc[workload], indent=1))


if __name__ == "__main__":
    main()
