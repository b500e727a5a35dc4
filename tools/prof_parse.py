#!/usr/bin/env python
"""Parse-kernel probe on the bench dataset: is the time of paf_parse_kernel a property of the kernel or of what ran before it?"""
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


def loads(h, tag, n=3, flush=0):
    h.set_option("flush_l2", flush)
    out = []
    for _ in range(n):
        h.graph_reset()
        r1 = h.load_device_view_ptr(rr.dev.data_ptr(), rr.avail, False)
        out.append(round(r1.parse_ms, 2))
    print(tag, out, flush=True)


h = capi.Hera(0)
loads(h, "fresh context, loads only:")
loads(h, "same, L2 flushed before the kernel:", flush=1)
h.graph_reset()
h.load_device_view_ptr(rr.dev.data_ptr(), rr.avail, False)
h.load_device_view_ptr(cr.dev.data_ptr(), cr.avail, True)
h.finalize(bench.METRICS)
loads(h, "after one finalize:")
pp = capi.default_path_params(rb_att=bench.RB_ATT)
h.graph_reset()
h.load_device_view_ptr(rr.dev.data_ptr(), rr.avail, False)
h.load_device_view_ptr(cr.dev.data_ptr(), cr.avail, True)
h.finalize(bench.METRICS)
h.build_paths([("mc", "es"), ("mc", "os"), ("det", "es"), ("det", "os")], pp, 1)
loads(h, "after finalize + walks:")
torch.cuda.empty_cache()
loads(h, "after torch.cuda.empty_cache():")
h.close()
h = capi.Hera(0)
loads(h, "second fresh context:")
free, total = torch.cuda.mem_get_info()
print(f"device memory free {free >> 20} MiB of {total >> 20} MiB")
h.close()
