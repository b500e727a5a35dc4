#!/usr/bin/env python
"""Times paf_parse_kernel on the rr text of the bench dataset for every library under telomeri_b200/build/variants
(tools/build_variant.sh: other tile sizes / threads / CTAs per SM) and the regular build."""
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
SNIPPET = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r)
import bench
from telomeri_b200 import capi
ds, _ = bench.make_dataset(bench.HEADLINE)
t = torch.from_numpy(np.fromfile(ds / "rr_overlaps.paf", dtype=np.uint8)).to("cuda:0")
torch.cuda.synchronize()
h = capi.Hera(0)
out = []
for _ in range(4):
    h.graph_reset()
    r = h.load_device_view_ptr(t.data_ptr(), t.numel(), False)
    out.append(round(r.parse_ms, 2))
print(out, r.records, r.kept)
h.close()
''' % str(ROOT)

if __name__ == "__main__":
    import bench
    bench.make_dataset(bench.HEADLINE)
    libs = [None] + sorted((ROOT / "telomeri_b200" / "build" / "variants").glob("parse_*.so"))
    for lib in libs:
        env = dict(os.environ)
        if lib:
            env["HERA_B200_LIB"] = str(lib)
        p = subprocess.run([sys.executable, "-c", SNIPPET], env=env, capture_output=True, text=True)
        print(lib.name if lib else "regular build (320 threads, 30720-byte tiles, 3 CTAs/SM)", (p.stdout.strip().splitlines() or [p.stderr[-300:]])[-1], flush=True)
