"""Drives the C ABI in the call order of the reference's main() (src/Main.cpp:211-342) and returns
every intermediate table under the section names of the heradump format (oracle/hera_dump.hpp),
so tests can compare the CUDA path with the oracle array by array.  Orchestration only."""
from pathlib import Path

import numpy as np

from . import capi


def parse_options(opts):
    """['--mode','min','--rb-att','300', ...] (the reference's flags in the harness spelling) -> dict."""
    o = {"mode": "avg", "min_oll": 0, "min_olp": 0.0, "max_ohl": 2 ** 64 - 1, "max_ohp": 1.0, "rb_att": 2000, "bt_att": 60,
         "len_thr": 4700000, "w_size": 1000, "r_thr": 0.9, "test_load_num": 0, "metrics": ["es", "os"]}
    it = iter(opts)
    for k in it:
        v = next(it)
        key = k.lstrip("-").replace("-", "_")
        if key == "metrics":
            o[key] = v.split(",")
        elif key == "mode":
            o[key] = v
        elif key in ("min_olp", "max_ohp", "r_thr"):
            o[key] = float(v)
        else:
            o[key] = int(v)
    return o


def run_hot_path(ds_dir, opts=(), seed=0, device=0, stages="all", hera=None, batched=False, exports=True, loader="file"):
    """exports=False: make every (collective) call but copy no table back -- the other ranks of a gang at large sizes.
    loader="device_view": the PAF files are put into device tensors of exactly their size and parsed in place
    (hera_graph_load_device_view) instead of going through hera_graph_load_file; loader="bytes": host arrays through
    hera_graph_load_bytes."""
    o = parse_options(list(opts))
    ds_dir = Path(ds_dir)
    own = hera is None
    h = hera or capi.Hera(device)
    out = {}
    try:
        h.set_filter(capi.default_filter(o["mode"], o["min_oll"], o["min_olp"], o["max_ohl"], o["max_ohp"]))
        if loader == "device_view":
            import torch
            held = []   # the library reads node names from these buffers until the graph is finalized / exported

            def load(name, anchors):
                data = np.fromfile(ds_dir / name, dtype=np.uint8)
                t = torch.from_numpy(data).to(torch.device("cuda", device)) if data.size else torch.empty(0, dtype=torch.uint8, device=torch.device("cuda", device))
                held.append(t)
                return h.load_device_view_ptr(t.data_ptr() if data.size else 0, int(data.size), anchors, o["test_load_num"])
            out["_held"] = held
        elif loader == "bytes":   # host arrays through hera_graph_load_bytes (copied and parsed chunk by chunk)
            def load(name, anchors):
                data = np.fromfile(ds_dir / name, dtype=np.uint8)
                return h.load_host_ptr(data.ctypes.data if data.size else 0, int(data.size), anchors, o["test_load_num"])
        else:
            def load(name, anchors):
                return h.load_file(ds_dir / name, anchors, o["test_load_num"])
        r1 = load("rr_overlaps.paf", False)
        r2 = load("cr_overlaps.paf", True)
        out["load_ctr"] = np.array([r1.records, r2.records], dtype=np.uint64)
        out["load_cum_edges"] = np.array([r1.kept_total, r2.kept_total], dtype=np.uint64)
        info = h.finalize(o["metrics"])
        # device times of the stages (CUDA events inside the library), for the full-size runs of tools/run_config.py
        tm = out["timings_ms"] = {"parse": [r1.parse_ms, r2.parse_ms], "h2d": [r1.h2d_ms, r2.h2d_ms], "paf_bytes": [int(r1.bytes), int(r2.bytes)],
                                  "intern": info.intern_ms, "csr": info.csr_ms, "ranks": info.rank_ms, "slots": int(info.slots)}
        if exports:
            out.update(h.export_graph())
            out.update(slot_sections(out))
        elif h.size > 1:
            h.gather_rows()   # the other ranks' export_graph is collective
        if stages == "graph":
            return out
        pp = capi.default_path_params(o["rb_att"], o["bt_att"], o["len_thr"], o["w_size"], o["r_thr"])
        h.timer_start()
        if batched:   # one call, all walk kernels in flight together (hera_paths_build)
            stats = h.build_paths([("mc", m) for m in o["metrics"]] + [("det", m) for m in o["metrics"]], pp, seed)
        else:         # the reference's call sequence (src/Main.cpp:237-243)
            stats = []
            for m in o["metrics"]:
                stats.append(h.monte_carlo(m, pp, seed))
            for m in o["metrics"]:
                stats.append(h.deterministic(m))
        tm["walk_stage"] = h.timer_stop()
        out["walk_stats"] = stats
        tm["walks"] = [round(st.ms, 3) for st in stats]
        tm["walk_entries_examined"] = int(sum(st.steps for st in stats))
        if exports:
            out.update(h.export_paths("p_"))
        h.unique()
        if exports:
            u = h.export_paths("u_")
            out.update({k: v for k, v in u.items() if k.startswith("u_")})
        if stages == "paths":
            return out
        out.update(h.groups(pp))
        tm["groups"] = out["ms"]
        return out
    finally:
        if own:
            h.close()


def slot_sections(g):
    """a_* sections (each adjacency slot as the owner's list stores it) derived from the export:
    pure re-indexing of exported arrays for comparison with the oracle dump."""
    rec, side = g["adj_rec"], g["adj_side"].astype(bool)
    pick = lambda a, b: np.where(side, g[b][rec], g[a][rec])
    return {
        "a_q": pick("e_q", "e_t"), "a_t": pick("e_t", "e_q"), "a_qs": pick("e_qs", "e_ts"), "a_qe": pick("e_qe", "e_te"),
        "a_ts": pick("e_ts", "e_qs"), "a_te": pick("e_te", "e_qe"), "a_os": g["e_os"][rec], "a_si": g["e_si"][rec],
        "a_es": g["e_es"][rec], "a_strand": g["e_strand"][rec],
    }
