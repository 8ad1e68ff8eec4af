#!/usr/bin/env python
"""Per source line: warp instructions executed and stall samples from `ncu -i X.ncu-rep --page source --csv`,
for one kernel (substring match on the function name).  Usage: ncu_source_lines.py rep kernel_substr [top]"""
import csv
import subprocess
import sys

rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source=cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
agg = {}
cur_file, cur_fn, hdr = None, None, None
use = False
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif r[0] == "Function Name":
        cur_fn = r[1]
        use = pat in cur_fn
    elif r[0] == "Line No":
        hdr = r
    elif use and hdr and r[0] not in ("", "Line No"):
        try:
            ie = float(r[hdr.index("Instructions Executed")])
            te = float(r[hdr.index("Thread Instructions Executed")])
            ss = float(r[hdr.index("# Samples")])
        except ValueError:
            continue
        k = (cur_fn[:60], cur_file, int(r[0]), r[1].strip()[:110])
        a = agg.setdefault(k, [0, 0, 0])
        a[0] += ie
        a[1] += te
        a[2] += ss
fns = sorted({k[0] for k in agg})
for fn in fns:
    items = [(k, v) for k, v in agg.items() if k[0] == fn]
    tot = sum(v[0] for _, v in items)
    tots = sum(v[2] for _, v in items)
    print(f"## {fn}: {tot:.3e} warp instructions, {tots:.0f} samples")
    for k, v in sorted(items, key=lambda kv: -kv[1][0])[:top]:
        print(f"{100*v[0]/tot:5.1f}% inst {100*v[2]/max(tots,1):5.1f}% smp  thr/inst {v[1]/max(v[0],1):5.1f}  {k[1]}:{k[2]}  {k[3]}")
