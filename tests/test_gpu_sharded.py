"""The sharded hot path: PAF parsed by byte range, kept overlaps all-gathered, walks
dealt out by walk id, paths all-gathered -- every rank must end up with exactly the tables one GPU
produces, i.e. the oracle's.  Runs the ranks as a local gang (one process, one thread per rank) so it
also works on a single-GPU box; with >= 2 GPUs the ranks sit on different devices and the copies are
peer copies.  The NCCL transport is covered by test_gpu_nccl.py (needs >= 2 GPUs)."""
import numpy as np
import pytest

import common
from test_gpu_parity import GRAPH_SECTIONS, GROUP_SECTIONS, METRIC_ID, PATH_SECTIONS

pytestmark = pytest.mark.gpu


def gang_devices(n):
    import torch
    k = max(1, torch.cuda.device_count())
    return [r % k for r in range(n)]


def run_gang(ds, opts, seed, nranks, stages="all", batched=False):
    from telomeri_b200 import capi, pipeline
    with capi.LocalGang(nranks, gang_devices(nranks)) as gang:
        return gang.each(lambda r, h: pipeline.run_hot_path(ds, opts, seed=seed, stages=stages, hera=h, batched=batched))


@pytest.mark.parametrize("nranks", [2, 3])
@pytest.mark.parametrize("case", ["small_default", "mixed_sqrt", "bundled_min", "small_windows"] + common.WINDOW_CASES + common.WRAP_CASES)
def test_sharded_matches_oracle(case, nranks, oracle_bin, datasets, tmp_path):
    from telomeri_b200 import pipeline
    golden = common.load_golden(case)
    ds = datasets[golden["dataset"]]
    ora = common.run_dump(oracle_bin, ds, tmp_path / "o.bin", golden["options"], extra=["--rng", "philox", "--seed", "99"])
    outs = run_gang(ds, golden["options"], 99, nranks, batched=(nranks == 3))
    opts = pipeline.parse_options(golden["options"])
    ranks = [f"rank_{METRIC_ID[m]}" for m in opts["metrics"]]
    for r, gpu in enumerate(outs):
        common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS + ranks + PATH_SECTIONS + GROUP_SECTIONS, f"{case} rank {r}/{nranks}")
        if case in common.WINDOW_CASES:
            common.assert_enters_window_code(gpu, golden["options"], f"{case} rank {r}/{nranks}")


def test_more_ranks_than_work(oracle_bin, tmp_path):
    """Slices that own no token at all, records cut by every slice boundary."""
    rec = "r1 1000 10 990 + ctg1 5000 0 980 970 980 0 tp:A:S cm:i:1 s1:i:1 dv:f:0.1\n"
    rec2 = "r2 1000 10 990 - ctg1 5000 100 1080 970 980 0 tp:A:S cm:i:1 s1:i:1 dv:f:0.1\n"
    texts = {"two": rec + rec2, "empty": "", "blank": " \n\t ", "many": (rec + rec2) * 500, "twelve": rec + " ".join(rec2.split()[:12]),
             "eleven": (rec + rec2) * 7 + " ".join(rec2.split()[:11])}
    for name, text in texts.items():
        d = tmp_path / name
        d.mkdir()
        (d / "rr_overlaps.paf").write_text("")
        (d / "cr_overlaps.paf").write_text(text)
        ora = common.run_dump(oracle_bin, d, d / "o.bin", ["--mode", "max"], extra=["--stage", "graph"])
        for nranks in (2, 5):
            for r, gpu in enumerate(run_gang(d, ["--mode", "max"], 0, nranks, stages="graph")):
                common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS, f"{name} rank {r}/{nranks}")


def test_irregular_file_falls_back_to_the_whole_file_reader(oracle_bin, tmp_path):
    """A record formatted extraction fails on: hera_graph_load_shard refuses (code 20), hera_graph_load_file
    reloads the file unsharded on every rank; the result is still the reference's."""
    from telomeri_b200 import capi
    rec = "r1 1000 10 990 + ctg1 5000 0 980 970 980 0 tp:A:S cm:i:1 s1:i:1 dv:f:0.1\n"
    rec2 = "r2 1000 10 990 - ctg1 5000 100 1080 970 980 0 tp:A:S cm:i:1 s1:i:1 dv:f:0.1\n"
    text = (rec + rec2) * 40 + rec.replace("\n", " s2:i:5\n") + (rec2 + rec) * 40     # a 17-column line in the middle
    d = tmp_path / "irregular"
    d.mkdir()
    (d / "rr_overlaps.paf").write_text("")
    (d / "cr_overlaps.paf").write_text(text)
    ora = common.run_dump(oracle_bin, d, d / "o.bin", ["--mode", "max"], extra=["--stage", "graph"])
    for r, gpu in enumerate(run_gang(d, ["--mode", "max"], 0, 2, stages="graph")):
        common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS, f"irregular rank {r}")
    # the slice-level call reports it instead
    data = np.frombuffer(text.encode(), dtype=np.uint8)
    with capi.LocalGang(2, gang_devices(2)) as gang:
        def load(r, h):
            h.set_filter(capi.default_filter("max"))
            b, e, a = capi.shard_range(data.size, r, 2)
            sl = np.ascontiguousarray(data[b:a])
            try:
                h.load_shard_ptr(sl.ctypes.data, False, e - b, a - b, int(data[b - 1]) if b else -1, a == data.size, True)
            except capi.HeraError as err:
                return err.code
            return 0
        assert gang.each(load) == [20, 20]


@pytest.mark.parametrize("on_device", [1, 2], ids=["copied", "in_place"])
def test_shard_from_device_memory(on_device, oracle_bin, datasets, tmp_path):
    """hera_graph_load_shard with the slice already resident in HBM: copied into the library's buffer (1) or parsed in
    place (2, what bench.py times -- a slice ends in the middle of the file, the bytes behind it are not readable)."""
    import torch
    from telomeri_b200 import capi
    ds = datasets["small"]
    ora = common.run_dump(oracle_bin, ds, tmp_path / "o.bin", [], extra=["--stage", "graph"])
    files = [np.fromfile(ds / "rr_overlaps.paf", dtype=np.uint8), np.fromfile(ds / "cr_overlaps.paf", dtype=np.uint8)]
    nranks = 3
    devs = gang_devices(nranks)
    with capi.LocalGang(nranks, devs) as gang:
        def run(r, h):
            h.set_filter(capi.default_filter())
            res = []
            keep = []
            for data, anchors in zip(files, (False, True)):
                b, e, a = capi.shard_range(data.size, r, nranks)
                t = torch.from_numpy(data[b:a].copy()).to(f"cuda:{devs[r]}") if a > b else torch.zeros(1, dtype=torch.uint8, device=f"cuda:{devs[r]}")
                torch.cuda.synchronize(devs[r])
                keep.append(t)
                res.append(h.load_shard_ptr(t.data_ptr(), on_device, e - b, a - b, int(data[b - 1]) if b else -1, a == data.size, anchors))
            h.finalize(["es", "os"])
            g = h.export_graph()
            g["load_ctr"] = np.array([x.records for x in res], dtype=np.uint64)
            g["load_cum_edges"] = np.array([x.kept_total for x in res], dtype=np.uint64)
            return g
        for r, gpu in enumerate(gang.each(run)):
            common.assert_sections_equal(ora, gpu, ["load_ctr", "load_cum_edges", "node_len", "node_anchor", "node_names", "e_q", "e_t", "e_os", "e_es",
                                                    "adj_off", "adj_rec", "adj_side", "canon", "rank_0", "rank_1"], f"device shard rank {r}")
