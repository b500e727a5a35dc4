#!/usr/bin/env python
"""Host-side timeline of one hot-path step on the bench dataset (perf_counter around every C-ABI call)."""
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
import torch

import bench
from telomeri_b200 import capi

ds, meta = bench.make_dataset(bench.WORKLOAD["seed"])
dev = torch.device("cuda", 0)
rr = torch.from_numpy(np.fromfile(ds / "rr_overlaps.paf", dtype=np.uint8)).to(dev)
cr = torch.from_numpy(np.fromfile(ds / "cr_overlaps.paf", dtype=np.uint8)).to(dev)
torch.cuda.synchronize()
h = capi.Hera(0)
pp = capi.default_path_params(rb_att=bench.WORKLOAD["rb_att"])
batched = "--separate" not in sys.argv
for a in sys.argv[1:]:
    if a.startswith("--variant="):
        h.set_option("walk_variant", int(a.split("=")[1]))
    if a.startswith("--warps="):
        h.set_option("walk_warps_per_sm", int(a.split("=")[1]))
for rep in range(5):
    t = [time.perf_counter()]
    def lap(name):
        t.append(time.perf_counter())
        return f"{name} {1e3 * (t[-1] - t[-2]):.2f}"
    out = []
    h.graph_reset(); h.paths_reset(); out.append(lap("reset"))
    h.timer_start()
    h.load_device_view_ptr(rr.data_ptr(), rr.numel(), False); out.append(lap("load_rr"))
    h.load_device_view_ptr(cr.data_ptr(), cr.numel(), True); out.append(lap("load_cr"))
    info = h.finalize(["es", "os"]); out.append(lap("finalize"))
    g_ms = h.timer_stop()
    if batched:
        h.build_paths([("mc", "es"), ("mc", "os"), ("det", "es"), ("det", "os")], pp, 1); out.append(lap("build"))
    else:
        for m in ("es", "os"):
            h.monte_carlo(m, pp, 1)
        for m in ("es", "os"):
            h.deterministic(m)
        out.append(lap("walk calls"))
    h.unique(); out.append(lap("unique"))
    h.groups(pp); out.append(lap("groups"))
    print(f"rep {rep}: graph(ev) {g_ms:.2f} | " + " | ".join(out))
h.close()
