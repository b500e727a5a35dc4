#!/usr/bin/env python
"""One step of the hot path on the bench dataset (configs[3] unless HERA_BENCH_WORKLOAD says otherwise), nothing else:
the command tools/profile_round.sh runs under ncu (launch list, --set full captures)."""
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
pp = capi.default_path_params(rb_att=bench.RB_ATT)
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1
for _ in range(steps):
    s = bench.hot_path_step(h, pp, rr, cr, True, False)
print({k: s[k] for k in ("graph_ms", "parse_ms", "intern_ms", "csr_ms", "rank_ms", "paths_ms", "groups_ms", "records", "kept", "walks", "launches")})
h.close()
