#!/usr/bin/env python
"""Graph-build probe on the bench dataset: HERA_DEBUG_TIMING=1 python tools/prof_graph.py prints the device time of the
phases of hera_graph_finalize (stderr) and the stage times of the loads."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch

import bench
from telomeri_b200 import capi

dev = torch.device("cuda", 0)
ds, meta = bench.make_dataset(bench.HEADLINE)
rr = bench.Slice(ds / "rr_overlaps.paf", 0, 1, dev, capi, torch)
cr = bench.Slice(ds / "cr_overlaps.paf", 0, 1, dev, capi, torch)
h = capi.Hera(0)
for ctas in ([int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [2]):
  h.set_option("verify_ctas_per_sm", ctas)
  print("verify CTAs per SM:", ctas)
  for rep in range(3):
    h.graph_reset()
    h.timer_start()
    r1 = h.load_device_view_ptr(rr.dev.data_ptr(), rr.avail, False)
    r2 = h.load_device_view_ptr(cr.dev.data_ptr(), cr.avail, True)
    info = h.finalize(bench.METRICS)
    t = h.timer_stop()
    print(f"rep {rep}: graph {t:.2f} ms, parse {r1.parse_ms:.2f} + {r2.parse_ms:.2f} ms, intern {info.intern_ms:.2f}, csr {info.csr_ms:.2f}, ranks {info.rank_ms:.2f}")
h.close()
