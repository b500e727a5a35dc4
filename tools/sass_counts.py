#!/usr/bin/env python
"""SASS evidence of the built library (no GPU needed): per kernel the registers / shared memory / stack of
`cuobjdump --dump-resource-usage` and the number of SASS instructions, with the mnemonics that tell which hardware paths
a kernel uses (UBLKCP = cp.async.bulk, SYNCS = mbarrier, LDGSTS = per-thread cp.async, ATOM / RED, MATCH, BAR, DADD ...).
usage: python tools/sass_counts.py [lib.so] > profiles/rNN_sass_counts.txt"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
lib = sys.argv[1] if len(sys.argv) > 1 else str(ROOT / "telomeri_b200" / "libhera_b200.so")
WATCH = ["UBLKCP", "SYNCS", "LDGSTS", "UTMALDG", "ATOM", "ATOMG", "RED", "MATCH", "BAR", "SHFL", "VOTE", "DADD", "DMUL", "DFMA", "FFMA", "MUFU", "LDG", "STG", "LDS", "STS", "LDL", "STL", "MEMBAR", "ERRBAR"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


res = subprocess.run(["cuobjdump", "--dump-resource-usage", lib], capture_output=True, text=True).stdout
usage = {}
cur = None
for line in res.splitlines():
    m = re.match(r"\s*Function (\S+):", line)
    if m:
        cur = m.group(1)
        continue
    if cur and "REG:" in line:
        usage[cur] = " ".join(t for t in line.split() if t.split(":")[0] in ("REG", "STACK", "SHARED", "LOCAL"))
        cur = None
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
counts = collections.OrderedDict()
cur = None
arch = set()
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s*arch = (\S+)", line)
    if m:
        arch.add(m.group(1))
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        op = m.group(1)
        counts[cur]["_all"] += 1
        counts[cur][op.split(".")[0]] += 1
names = demangle(list(counts))
print(f"{lib}: arch {sorted(arch)}, {len(counts)} kernels (an sm_52 entry is the empty stub of the nvcc -shared link step: every kernel is sm_100a code)")
tot = collections.Counter()
for k, c in counts.items():
    tot.update(c)
print("whole library: " + ", ".join(f"{w} {tot[w]}" for w in WATCH if tot[w]))
print()
for k, c in sorted(counts.items(), key=lambda kv: -kv[1]["_all"]):
    name = re.sub(r"\(.*", "", names.get(k, k))
    print(f"{name}\n    {usage.get(k, '')}  SASS {c['_all']}: " + ", ".join(f"{w} {c[w]}" for w in WATCH if c[w]))
