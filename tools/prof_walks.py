#!/usr/bin/env python
"""Walk-stage probe on the bench dataset (configs[3] unless HERA_BENCH_WORKLOAD says otherwise): every walk call alone and
all four together, under a few settings of the walk options.  python tools/prof_walks.py"""
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
h.load_device_view_ptr(rr.dev.data_ptr(), rr.avail, False)
h.load_device_view_ptr(cr.dev.data_ptr(), cr.avail, True)
info = h.finalize(bench.METRICS)
print(f"nodes {info.nodes} anchors {info.anchors} slots {info.slots}")
pp = capi.default_path_params(rb_att=bench.RB_ATT)
calls = [("mc", "es"), ("mc", "os"), ("det", "es"), ("det", "os")]
for warps in (32, 16):
    h.set_option("walk_warps_per_sm", warps)
    out = []
    for kind, m in calls:
        h.paths_reset()
        s = h.monte_carlo(m, pp, bench.WALK_SEED) if kind == "mc" else h.deterministic(m)
        out.append(f"{kind}_{m} {s.ms:.1f} ms ({s.walks} walks, {s.steps / 1e6:.0f}M entries)")
    print(f"alone, {warps} warps/SM: " + " | ".join(out))
for variant, warps in ((0, 32), (0, 40)):
    h.set_option("walk_variant", variant)
    h.set_option("walk_warps_per_sm", warps)
    h.set_option("walk_hints_reset", 1)
    for rep in range(3):
        h.paths_reset()
        h.timer_start()
        st = h.build_paths(calls, pp, bench.WALK_SEED)
        t = h.timer_stop()
        print(f"together, variant {variant}, {warps} warps/SM, rep {rep}: stage {t:.1f} ms, kernels {[round(s.ms, 1) for s in st]}")
h.close()
