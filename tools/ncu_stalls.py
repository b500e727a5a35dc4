#!/usr/bin/env python
"""Stall-sample totals per source line from `ncu --page source --csv --print-source cuda,sass`.
usage: ncu_stalls.py source.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
hdr = None; cur = ''; tot = {}
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur = r[1].split('/')[-1]; continue
    if r[0] == 'Line No':
        hdr = r; continue
    if hdr is None or not r[0].strip().isdigit() or len(r) != len(hdr):
        continue
    names = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    k = (cur, int(r[0]), r[1].strip()[:70])
    d = tot.setdefault(k, {})
    try:
        d['samp'] = d.get('samp', 0) + int(r[hdr.index('# Samples')])
    except ValueError:
        continue
    for n in names:
        try:
            d[n] = d.get(n, 0) + int(r[hdr.index(n)])
        except ValueError:
            pass
S = sum(d.get('samp', 0) for d in tot.values()) or 1
agg = {}
for d in tot.values():
    for n, v in d.items():
        if n != 'samp':
            agg[n] = agg.get(n, 0) + v
print(S, [(n, round(100 * v / S, 1)) for n, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]])
for k, d in sorted(tot.items(), key=lambda kv: -kv[1].get('samp', 0))[:top]:
    t3 = sorted(((v, n) for n, v in d.items() if n != 'samp'), reverse=True)[:2]
    print(f"{100 * d['samp'] / S:6.2f}  {k[0]}:{k[1]}  {k[2]}  {t3}")
