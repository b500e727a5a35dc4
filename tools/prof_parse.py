#!/usr/bin/env python
"""Small driver for ncu: the graph-build stages on the bench dataset, a few repetitions (no walks)."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
import torch

import bench
from telomeri_b200 import capi

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
ds, meta = bench.make_dataset(bench.WORKLOAD["seed"])
dev = torch.device("cuda", 0)
rr = torch.from_numpy(np.fromfile(ds / "rr_overlaps.paf", dtype=np.uint8)).to(dev)
cr = torch.from_numpy(np.fromfile(ds / "cr_overlaps.paf", dtype=np.uint8)).to(dev)
torch.cuda.synchronize()
h = capi.Hera(0)
h.set_option("flush_l2", 1)
for _ in range(reps):
    h.graph_reset()
    r1 = h.load_device_view_ptr(rr.data_ptr(), rr.numel(), False)
    r2 = h.load_device_view_ptr(cr.data_ptr(), cr.numel(), True)
    info = h.finalize(["es", "os"])
    print(f"parse {r1.parse_ms:.3f}+{r2.parse_ms:.3f} ms intern {info.intern_ms:.3f} csr {info.csr_ms:.3f} rank {info.rank_ms:.3f}")
h.close()
