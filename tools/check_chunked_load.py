#!/usr/bin/env python
"""Quick check on a GPU box (no torch import): hera_graph_load_bytes with forced small chunks gives the tables of
hera_graph_load_file on the `small` dataset, twice in a row on one context.  python tools/check_chunked_load.py"""
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
t0 = time.perf_counter()
import common  # noqa: E402
from telomeri_b200 import capi, pipeline  # noqa: E402

ds = common.make_dataset("small", Path("/tmp/hera_chk/small"))
print(f"dataset ready +{time.perf_counter() - t0:.1f}s", flush=True)
with capi.Hera(0) as h:
    ref = pipeline.run_hot_path(ds, [], stages="graph", hera=h, loader="file")
    for chunk in (8 * 30720, 0, 8 * 30720):
        h.graph_reset()
        h.set_option("load_chunk_bytes", chunk)
        got = pipeline.run_hot_path(ds, [], stages="graph", hera=h, loader="bytes")
        for k, v in ref.items():
            if isinstance(v, np.ndarray):
                assert np.array_equal(v, got[k]), (chunk, k)
        print(f"chunk {chunk}: {sum(isinstance(v, np.ndarray) for v in ref.values())} tables equal, h2d {got['timings_ms']['h2d']} ms, parse {got['timings_ms']['parse']} ms +{time.perf_counter() - t0:.1f}s", flush=True)
print("OK")
