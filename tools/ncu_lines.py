#!/usr/bin/env python
"""Per-source-line totals from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`:
warp instructions executed and stall samples per (file, line), largest first.
usage: ncu_lines.py source.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = ""
hdr = None
tot = {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        i_inst = hdr.index("Instructions Executed")
        i_thr = hdr.index("Thread Instructions Executed")
        i_samp = hdr.index("# Samples")
        continue
    if hdr is None or not r[0].strip().isdigit():
        continue
    try:
        inst, thr, samp = int(r[i_inst]), int(r[i_thr]), int(r[i_samp])
    except ValueError:
        continue
    k = (cur_file, int(r[0]), r[1].strip()[:110])
    a = tot.setdefault(k, [0, 0, 0])
    a[0] += inst; a[1] += thr; a[2] += samp
all_inst = sum(v[0] for v in tot.values()) or 1
all_samp = sum(v[2] for v in tot.values()) or 1
print(f"total warp instructions {all_inst}, samples {all_samp}")
print(f"{'inst%':>6} {'samp%':>6} {'thr/inst':>8}  file:line  source")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100 * v[0] / all_inst:6.2f} {100 * v[2] / all_samp:6.2f} {v[1] / max(v[0], 1):8.1f}  {k[0]}:{k[1]}  {k[2]}")
