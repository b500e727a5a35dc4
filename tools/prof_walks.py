#!/usr/bin/env python
"""Walk-stage probe on the bench dataset: per-call kernel times for a few settings of resident warps."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
import torch

import bench
from telomeri_b200 import capi

settings = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [32]
ds, meta = bench.make_dataset(bench.WORKLOAD["seed"])
dev = torch.device("cuda", 0)
rr = torch.from_numpy(np.fromfile(ds / "rr_overlaps.paf", dtype=np.uint8)).to(dev)
cr = torch.from_numpy(np.fromfile(ds / "cr_overlaps.paf", dtype=np.uint8)).to(dev)
torch.cuda.synchronize()
h = capi.Hera(0)
h.load_device_view_ptr(rr.data_ptr(), rr.numel(), False)
h.load_device_view_ptr(cr.data_ptr(), cr.numel(), True)
h.finalize(["es", "os"])
pp = capi.default_path_params(rb_att=bench.WORKLOAD["rb_att"])
for w in settings:
    h.set_option("walk_warps_per_sm", w)
    for rep in range(2):
        h.paths_reset()
        out = []
        for m in ("es", "os"):
            s = h.monte_carlo(m, pp, 20261104)
            out.append(f"mc_{m} {s.ms:.2f} ms walks {s.walks} found {s.found} entries {s.steps / 1e6:.0f}M")
        for m in ("es", "os"):
            s = h.deterministic(m)
            out.append(f"det_{m} {s.ms:.2f} ms walks {s.walks} found {s.found}")
        h.timer_start()
        n = h.unique()
        t = h.timer_stop()
    d = h.export_paths()["mc_draws_per_walk"].astype(np.int64)
    half = d.size // 2
    for name, x in (("es", d[:half]), ("os", d[half:])):
        per_anchor = x.reshape(-1, bench.WORKLOAD["rb_att"])
        print(f"  draws per walk {name}: mean {x.mean():.1f} max {x.max()} p99 {np.percentile(x, 99):.0f}; per-anchor mean {per_anchor.mean(axis=1).round(0).tolist()}")
    print(f"warps/SM {w}: " + " | ".join(out) + f" | unique {n} {t:.2f} ms")
h.close()
