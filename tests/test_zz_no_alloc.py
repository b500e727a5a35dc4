"""hera_paf_parse_noalloc (the allocation-free variant of the tokenizer / filter / score kernel, scratch in a file-scope
device array) against the existing entry point hera_graph_load_device_view on the same text, through the C ABI, on several
calls in a row (the calls share the scratch array: each must clear what the previous one left).  Bit-exact -- the
tolerance of every parity test here.  Collected last on purpose."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu

U32 = ("qs", "qe", "ts", "te", "qlen", "tlen", "rec")
F32 = ("os", "si", "es")
U64 = ("qhash", "thash", "qname", "tname")


def existing_entry_point(h, text_dev, n, anchors, mode):
    """Rows of ONE text through hera_graph_load_device_view + finalize + export (a fresh graph per text)."""
    from telomeri_b200 import capi
    h.graph_reset()
    h.set_filter(capi.default_filter(mode))
    r = h.load_device_view_ptr(text_dev.data_ptr(), n, anchors)
    h.finalize(["es"])
    g = h.export_graph()
    return r, g


def no_alloc(torch, text_dev, n, anchors, mode, cols, result):
    from telomeri_b200 import capi
    capacity = n // 24 + 2
    capi.paf_parse_noalloc(0, text_dev.data_ptr(), n, anchors, capi.default_filter(mode), {k: v.data_ptr() for k, v in cols.items()}, capacity,
                           result.data_ptr(), stream=0)
    torch.cuda.synchronize()
    res = result.cpu().numpy().astype(np.uint64)
    kept = int(res[1])
    return res, {k: v[:kept].cpu().numpy() for k, v in cols.items()}


def test_no_alloc_parse_matches_existing_entry_point(datasets):
    import torch
    from telomeri_b200 import capi
    dev = torch.device("cuda", 0)
    ds = datasets["small"]
    texts = []
    for name, anchors in (("rr_overlaps.paf", False), ("cr_overlaps.paf", True)):
        data = np.fromfile(ds / name, dtype=np.uint8)
        texts.append((name, anchors, torch.from_numpy(data).to(dev), int(data.size), data))
    biggest = max(t[3] for t in texts)
    assert biggest <= capi.NOALLOC_MAX_TEXT_BYTES
    cap = biggest // 24 + 2
    # outputs are the caller's: allocated ONCE here, reused by every call (the entry point itself allocates nothing)
    cols = {k: torch.empty(cap, dtype=torch.int32, device=dev) for k in U32}
    cols.update({k: torch.empty(cap, dtype=torch.float32, device=dev) for k in F32})
    cols["flags"] = torch.empty(cap, dtype=torch.uint8, device=dev)
    cols.update({k: torch.empty(cap, dtype=torch.int64, device=dev) for k in U64})
    result = torch.zeros(4, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()

    with capi.Hera(0) as h:
        # rr (long text, many tiles), cr (short text: the chains of the previous call must not leak in), rr again, another filter
        for name, anchors, text_dev, n, data, mode in [t + ("avg",) for t in (texts[0], texts[1], texts[0])] + [texts[0] + ("min",)]:
            r, g = existing_entry_point(h, text_dev, n, anchors, mode)
            for k in cols:
                cols[k].fill_(-1 if k not in F32 and k != "flags" else 7)   # stale contents must not survive
            res, got = no_alloc(torch, text_dev, n, anchors, mode, cols, result)
            what = f"{name} ({mode})"
            assert int(res[0]) == r.records and int(res[1]) == r.kept, f"{what}: records / kept {res[:2]} vs {r.records} / {r.kept}"
            assert int(res[2]) == 2 ** 64 - 1, f"{what}: unexpected irregular record {res[2]}"
            assert r.kept > 100
            for k in ("qs", "qe", "ts", "te"):
                assert np.array_equal(got[k].view(np.uint32), g["e_" + k]), f"{what}: column {k}"
            for k in F32:
                assert np.array_equal(got[k].view(np.uint32), g["e_" + k].view(np.uint32)), f"{what}: column {k} (bits)"
            assert np.array_equal(got["flags"] & 1, g["e_strand"]), f"{what}: strand flag"
            assert np.array_equal((got["flags"] >> 1) & 1, np.full(r.kept, int(anchors), np.uint8)), f"{what}: anchors flag"
            # record indices ascend, lengths are the creating records' lengths of the nodes
            assert np.all(np.diff(got["rec"].view(np.uint32).astype(np.int64)) > 0) and int(got["rec"].view(np.uint32)[-1]) < r.records
            # name references point at the names of the nodes the existing path assigned
            names = g["node_names"].tobytes()
            off = g["node_name_off"]
            for col, ids in (("qname", g["e_q"]), ("tname", g["e_t"])):
                ref = got[col].view(np.uint64)
                for i in (0, 1, r.kept // 2, r.kept - 1):
                    o, ln = int(ref[i] >> np.uint64(16)) & ((1 << 44) - 1), int(ref[i] & np.uint64(0xffff))
                    assert data[o:o + ln].tobytes() == names[int(off[ids[i]]):int(off[ids[i] + 1])], f"{what}: {col}[{i}]"
            # equal names <=> equal hashes on this data set (node ids from the existing path)
            qh = got["qhash"].view(np.uint64)
            th = got["thash"].view(np.uint64)
            both_h = np.concatenate([qh, th])
            both_id = np.concatenate([g["e_q"], g["e_t"]])
            assert np.unique(both_h).size == np.unique(both_id).size, f"{what}: name hashes"


def test_no_alloc_rejects_what_it_cannot_hold():
    import torch
    from telomeri_b200 import capi
    dev = torch.device("cuda", 0)
    small = torch.zeros(4096, dtype=torch.uint8, device=dev)
    cols = {k: small.data_ptr() for k in U32 + F32 + ("flags",) + U64}
    result = torch.zeros(4, dtype=torch.int64, device=dev)
    with pytest.raises(capi.HeraError) as e:
        capi.paf_parse_noalloc(0, small.data_ptr(), capi.NOALLOC_MAX_TEXT_BYTES + 1, False, capi.default_filter(), cols, 1 << 40, result.data_ptr())
    assert e.value.code == 60 and "HERA_NOALLOC_MAX_TEXT_BYTES" in str(e.value)
    with pytest.raises(capi.HeraError) as e:
        capi.paf_parse_noalloc(0, small.data_ptr(), 4096, False, capi.default_filter(), cols, 10, result.data_ptr())
    assert e.value.code == 61
    with pytest.raises(capi.HeraError) as e:
        capi.paf_parse_noalloc(0, small.data_ptr() + 1, 100, False, capi.default_filter(), cols, 1000, result.data_ptr())
    assert e.value.code == 101
