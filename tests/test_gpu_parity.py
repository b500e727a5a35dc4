"""Parity of the CUDA path (through the C ABI) with the oracle on the same inputs: bit-exact for
every table of the hot path.  Runs on the B200 box (`-m gpu`)."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu

GRAPH_SECTIONS = ["load_ctr", "load_cum_edges", "node_len", "node_anchor", "node_name_off", "node_names",
                  "e_q", "e_t", "e_qs", "e_qe", "e_ts", "e_te", "e_os", "e_si", "e_es", "e_strand",
                  "adj_off", "adj_rec", "adj_side", "a_q", "a_t", "a_qs", "a_qe", "a_ts", "a_te", "a_os", "a_si", "a_es", "a_strand", "canon"]
PATH_SECTIONS = ["call_kind", "call_metric", "call_found", "mc_draws_per_walk", "p_off", "p_nodes", "p_slots", "p_len",
                 "u_off", "u_nodes", "u_slots", "u_len"]
GROUP_SECTIONS = ["pair_a1", "pair_a2", "pair_off", "pair_sorted", "grp_pair", "grp_size0", "grp_pig_off", "grp_pig",
                  "grp_consensus", "grp_valid", "pair_consensus"]
METRIC_ID = {"es": 0, "os": 1, "es_sqrt": 2, "os_sqrt": 3}

CASES = common.golden_cases()


def run_both(case, oracle_bin, datasets, tmp_path, seed):
    from telomeri_b200 import pipeline
    golden = common.load_golden(case)
    ds = datasets[golden["dataset"]]
    ora = common.run_dump(oracle_bin, ds, tmp_path / "o.bin", golden["options"], extra=["--rng", "philox", "--seed", str(seed)])
    gpu = pipeline.run_hot_path(ds, golden["options"], seed=seed)
    return golden, ora, gpu


@pytest.mark.parametrize("case", CASES)
def test_hot_path_matches_oracle(case, oracle_bin, datasets, tmp_path):
    golden, ora, gpu = run_both(case, oracle_bin, datasets, tmp_path, seed=1234)
    opts = __import__("telomeri_b200.pipeline", fromlist=["x"]).parse_options(golden["options"])
    ranks = [f"rank_{METRIC_ID[m]}" for m in opts["metrics"]]
    common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS + ranks, case + " graph")
    common.assert_sections_equal(ora, gpu, PATH_SECTIONS, case + " paths")
    common.assert_sections_equal(ora, gpu, GROUP_SECTIONS, case + " groups")
    if case in common.WINDOW_CASES:     # the CUDA result itself shows the window / border / multi-group code ran
        common.assert_enters_window_code(gpu, golden["options"], case + " (CUDA)")
        common.assert_enters_window_code(ora, golden["options"], case + " (oracle)")
    if case in common.WINDOW_CASES + common.WRAP_CASES and golden["known"]["call_found"][:len(opts["metrics"])] == [0] * len(opts["metrics"]):
        # no Monte-Carlo path survives len_threshold here, so nothing depends on the RNG: every section, groups and
        # consensus included, must equal the digests of the UNMODIFIED reference
        for k in PATH_SECTIONS + GROUP_SECTIONS:
            if k in golden["sections"] and k != "mc_draws_per_walk":
                got = {"dtype": str(gpu[k].dtype), "n": int(gpu[k].size), "sha": common.section_digest(gpu[k])}
                assert got == golden["sections"][k], f"{case}: {k} differs from the reference golden"
    # the graph part does not depend on the RNG: it must also equal the reference-generated golden digests
    for k in GRAPH_SECTIONS:
        if k in golden["sections"]:
            got = {"dtype": str(gpu[k].dtype), "n": int(gpu[k].size), "sha": common.section_digest(gpu[k])}
            assert got == golden["sections"][k], f"{case}: {k} differs from the reference golden"
    # deterministic walks do not depend on the RNG either
    assert gpu["call_found"].tolist()[len(opts["metrics"]):] == golden["known"]["call_found"][len(opts["metrics"]):]


@pytest.mark.parametrize("case", ["mixed_default", "bundled_min_win42"])
def test_device_digests_match_the_exported_tables(case, datasets):
    """hera_digest (what bench.py compares across ranks and with tests/golden/bench_digests.json) hashes exactly the
    tables the parity tests compare with the oracle: restated with numpy over the exported arrays."""
    from telomeri_b200 import capi, pipeline
    golden = common.load_golden(case)
    with capi.Hera(0) as h:
        out = pipeline.run_hot_path(datasets[golden["dataset"]], golden["options"], seed=3, hera=h, batched=True)
        dg = h.digest()
    nbr = out["a_q"] | (out["node_anchor"][out["a_q"]].astype(np.uint32) << np.uint32(31))
    delta = (out["a_qs"].astype(np.int64) - out["a_ts"].astype(np.int64)).astype(np.int32).view(np.uint32)
    walk_view = np.stack([nbr, out["a_es"].view(np.uint32), out["a_os"].view(np.uint32), delta], axis=1).reshape(-1)
    expect = {"node_len": out["node_len"], "node_anchor": out["node_anchor"], "e_q": out["e_q"], "e_t": out["e_t"],
              "adj_off": out["adj_off"].astype(np.uint32), "adj_pay": out["adj_rec"] * np.uint32(2) + out["adj_side"], "walk_view": walk_view,
              "canon": out["canon"], "rank_0": out["rank_0"], "rank_1": out["rank_1"],
              "u_off": out["u_off"], "u_nodes": out["u_nodes"], "u_slots": out["u_slots"], "u_len": out["u_len"]}
    for k in ("pair_a1", "pair_a2", "pair_sorted", "grp_size0", "grp_pig", "grp_consensus", "grp_valid", "pair_consensus"):
        expect[k] = out[k]
    for k, arr in expect.items():
        assert dg[k] == capi.digest_words(arr), k
    assert dg["rank_2"] == 0 and dg["rank_3"] == 0 and dg["u_len"] != 0 and dg["grp_pig"] != 0


def test_seed_changes_monte_carlo_only(oracle_bin, datasets, tmp_path):
    from telomeri_b200 import pipeline
    golden = common.load_golden("mixed_default")
    ds = datasets["mixed"]
    a = pipeline.run_hot_path(ds, golden["options"], seed=1, stages="paths")
    b = pipeline.run_hot_path(ds, golden["options"], seed=2, stages="paths")
    a2 = pipeline.run_hot_path(ds, golden["options"], seed=1, stages="paths")
    common.assert_sections_equal(a, a2, PATH_SECTIONS, "same seed, two runs")
    assert not np.array_equal(a["p_len"], b["p_len"])
    nmc = int(a["call_found"][:2].sum()), int(b["call_found"][:2].sum())
    assert a["call_found"].tolist()[2:] == b["call_found"].tolist()[2:]
    assert np.array_equal(a["p_len"][nmc[0]:], b["p_len"][nmc[1]:])


def test_reader_edge_cases_gpu(oracle_bin, tmp_path):
    """Token-stream semantics through the CUDA tokenizer, against the oracle."""
    from telomeri_b200 import pipeline
    rec = "r1 1000 10 990 + ctg1 5000 0 980 970 980 0 tp:A:S cm:i:1 s1:i:1 dv:f:0.1\n"
    rec2 = "r2 1000 10 990 - ctg1 5000 100 1080 970 980 0 tp:A:S cm:i:1 s1:i:1 dv:f:0.1\n"
    long_name = "r" * 700
    texts = {
        "plain": rec + rec2,
        "crlf_tabs": rec.replace(" ", "\r\n") + rec2.replace(" ", "\t\t"),
        "seventeen": rec.replace("\n", " s2:i:5\n") + rec2 + rec,
        "twelve": rec + " ".join(rec2.split()[:12]),
        "eleven": rec + " ".join(rec2.split()[:11]),
        "nonnumeric": rec + rec2.replace("1080", "x80") + rec,
        "overflow": rec + rec2.replace("5000", "99999999999") + rec,
        "trailing_garbage": rec + rec2.replace(" 0 tp", " 0x tp") + rec,   # mapq "0x": record taken, stream desynchronises after it
        "strand_two_chars": rec + rec2.replace(" - ", " -+ ") + rec,
        "signed": rec.replace(" 10 ", " +10 ") + rec2,
        "long_names": rec.replace("r1", long_name) + rec2 + rec.replace("ctg1", "ctg" + long_name),
        "empty": "",
        "only_space": " \n\t ",
        "many": (rec + rec2) * 3000,
    }
    for name, text in texts.items():
        d = tmp_path / name
        d.mkdir()
        (d / "rr_overlaps.paf").write_text("")
        (d / "cr_overlaps.paf").write_text(text)
        ora = common.run_dump(oracle_bin, d, d / "o.bin", ["--mode", "max"], extra=["--stage", "graph"])
        gpu = pipeline.run_hot_path(d, ["--mode", "max"], stages="graph")
        common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS, name)
    # same file as rr then cr: nodes created by the rr load stay reads
    d = tmp_path / "both"
    d.mkdir()
    (d / "rr_overlaps.paf").write_text(rec + rec2)
    (d / "cr_overlaps.paf").write_text(rec + rec2)
    ora = common.run_dump(oracle_bin, d, d / "o.bin", ["--mode", "max"], extra=["--stage", "graph"])
    gpu = pipeline.run_hot_path(d, ["--mode", "max"], stages="graph")
    common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS, "both")
    assert gpu["node_anchor"].sum() == 0


@pytest.mark.parametrize("digit", [9, 10, 11])
def test_radix_sort_digit_widths(digit, oracle_bin, datasets, tmp_path):
    """radix_sort_pairs picks 8- to 11-bit digits from the key width (two 10-bit passes for configs[3]'s 20-bit keys); small
    inputs only ever meet 8 bits, so the other widths are forced here: CSR order, pair order and groups must not change."""
    from telomeri_b200 import capi, pipeline
    golden = common.load_golden("mixed_default")
    ds = datasets[golden["dataset"]]
    ora = common.run_dump(oracle_bin, ds, tmp_path / "o.bin", golden["options"], extra=["--rng", "philox", "--seed", "3"])
    with capi.Hera(0) as h:
        h.set_option("radix_digit_bits", digit)
        gpu = pipeline.run_hot_path(ds, golden["options"], seed=3, hera=h, batched=True)
    common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS + ["rank_0", "rank_1"] + PATH_SECTIONS + GROUP_SECTIONS, f"digit {digit}")


def test_optimistic_row_reservation_and_its_overflow(oracle_bin, datasets, tmp_path):
    """Large loads do not reserve bytes / 24 rows: the first one counts its tokens, later ones on the same context try with the
    rows the context already holds and are repeated (with the count) when the text keeps more rows than that.  Forced here at
    small size with the "parse_count_rows" option: a context sized by 58 % of a file then loads the whole file."""
    from telomeri_b200 import capi
    ds = datasets["small"]
    full = (ds / "rr_overlaps.paf").read_bytes()
    cut = full.rfind(b"\n", 0, int(len(full) * 0.58)) + 1
    d = tmp_path / "prefix"
    d.mkdir()
    (d / "rr_overlaps.paf").write_bytes(full[:cut])
    (d / "cr_overlaps.paf").write_bytes(b"")
    ora_full = common.run_dump(oracle_bin, ds, tmp_path / "f.bin", [], extra=["--stage", "graph"])
    ora_pre = common.run_dump(oracle_bin, d, tmp_path / "p.bin", [], extra=["--stage", "graph"])
    cols = ["node_len", "e_q", "e_t", "e_qs", "e_te", "e_os", "e_es", "adj_off", "adj_rec", "canon", "rank_0"]
    with capi.Hera(0) as h:
        h.set_option("parse_count_rows", 1000)
        h.set_filter(capi.default_filter())
        for rep, (path, ora) in enumerate([(d, ora_pre), (ds, ora_full), (ds, ora_full), (d, ora_pre)]):
            # 1: counted; 2: optimistic, overflows (more rows than the prefix left room for), repeated; 3: optimistic, fits; 4: optimistic, fits
            h.graph_reset()
            r1 = h.load_file(path / "rr_overlaps.paf", False)
            r2 = h.load_file(path / "cr_overlaps.paf", True)
            assert [r1.records, r2.records] == ora["load_ctr"].tolist() and [r1.kept_total, r2.kept_total] == ora["load_cum_edges"].tolist(), rep
            h.finalize(["es"])
            common.assert_sections_equal(ora, h.export_graph(), cols, f"load {rep}")


def test_host_text_is_parsed_chunk_by_chunk_as_it_lands(oracle_bin, datasets, tmp_path):
    """hera_graph_load_bytes copies a host text in chunks on a copy stream and parses every chunk with its own launch as soon
    as it has landed (tokens and kept rows of the earlier chunks are carried on the device).  Forced here at small size with
    "load_chunk_bytes" = 8 tiles: dozens of chunks per file, records across every chunk border, texts that end on / next to
    a border, an irregular record in a late chunk (exact reader from there), --test-load-num, and the row reservations of
    large loads (counted = the text is awaited whole; optimistic; optimistic with overflow and repeat)."""
    from telomeri_b200 import capi, pipeline
    chunk = 8 * 30720
    # whole datasets, walks and groups included
    for name in ("small", "mixed"):
        ds = datasets[name]
        assert (ds / "rr_overlaps.paf").stat().st_size > 3 * chunk
        ora = common.run_dump(oracle_bin, ds, tmp_path / f"{name}.bin", [], extra=["--rng", "philox", "--seed", "5"])
        with capi.Hera(0) as h:
            h.set_option("load_chunk_bytes", chunk)
            gpu = pipeline.run_hot_path(ds, [], seed=5, hera=h, loader="bytes")
        common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS + PATH_SECTIONS + GROUP_SECTIONS, f"{name} in chunks")
    # ends of text around the chunk borders, an irregular record, a record limit
    rec = "read%05d 9000 10 8990 + ctg%d 50000 %d %d 8000 8980 0 tp:A:S cm:i:1 s1:i:1 dv:f:0.1\n"
    body = "".join(rec % (i, i % 3, i % 7, 8980 + i % 7) for i in range(11000))
    assert len(body) > 3 * chunk + 4096
    texts = {f"cut{k}{d:+d}": body[:k * chunk + d] for k in (1, 2, 3) for d in (-1, 0, 1, 16, 1500)}
    late = body.find("\n", 2 * chunk + 9000) + 1
    texts["irregular"] = body[:late] + "x" + body[late:]
    texts["no_newline"] = body.rstrip("\n")
    with capi.Hera(0) as h:
        h.set_option("load_chunk_bytes", chunk)
        for name, text in texts.items():
            for opts in ([], ["--test-load-num", "7000"]) if name in ("cut3+0", "irregular") else ([],):
                d = tmp_path / (name + "_" + str(len(opts)))
                d.mkdir()
                (d / "rr_overlaps.paf").write_text(text, newline="")
                (d / "cr_overlaps.paf").write_text(body[:5000], newline="")
                ora = common.run_dump(oracle_bin, d, d / "o.bin", ["--mode", "max"] + opts, extra=["--stage", "graph"])
                h.graph_reset()
                gpu = pipeline.run_hot_path(d, ["--mode", "max"] + opts, stages="graph", hera=h, loader="bytes")
                common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS, f"{name} {opts}")
    # the reservations of large loads
    ds = datasets["small"]
    full = (ds / "rr_overlaps.paf").read_bytes()
    pre = full[:full.rfind(b"\n", 0, int(len(full) * 0.58)) + 1]
    d = tmp_path / "prefix"
    d.mkdir()
    (d / "rr_overlaps.paf").write_bytes(pre)
    (d / "cr_overlaps.paf").write_bytes(b"")
    ora_full = common.run_dump(oracle_bin, ds, tmp_path / "f.bin", [], extra=["--stage", "graph"])
    ora_pre = common.run_dump(oracle_bin, d, tmp_path / "p.bin", [], extra=["--stage", "graph"])
    with capi.Hera(0) as h:
        h.set_option("load_chunk_bytes", chunk)
        h.set_option("parse_count_rows", 1000)
        for rep, (path, ora) in enumerate([(d, ora_pre), (ds, ora_full), (ds, ora_full), (d, ora_pre)]):   # counted; overflow + repeat; fits; fits
            h.graph_reset()
            gpu = pipeline.run_hot_path(path, [], stages="graph", hera=h, loader="bytes")
            common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS, f"reservation {rep}")


def test_names_of_65536_bytes_are_refused(tmp_path):
    """Name references hold 16 bits of length: a kept record with a longer name fails the load (code 22) instead of
    producing a truncated node name; the reference has no such limit (documented in INTEGRATION.md)."""
    from telomeri_b200 import capi
    rec = "r1 1000 10 990 + ctg1 5000 0 980 970 980 0 tp:A:S cm:i:1 s1:i:1 dv:f:0.1\n"
    text = rec + rec.replace("r1", "r" * 70000) + rec
    f = tmp_path / "long.paf"
    f.write_text(text)
    with capi.Hera(0) as h:
        h.set_filter(capi.default_filter("max"))
        with pytest.raises(capi.HeraError) as e:
            h.load_file(f, True)
        assert e.value.code == 22
        h.graph_reset()
        (tmp_path / "ok.paf").write_text(rec)
        assert h.load_file(tmp_path / "ok.paf", True).kept == 1   # the context stays usable


def test_device_view_loads_match_oracle(oracle_bin, tmp_path):
    """hera_graph_load_device_view parses the caller's device buffer in place: nothing behind byte n is readable, so
    texts whose end falls on / next to a tile, halo or 16-byte chunk boundary (the tokenizer stages 30 KiB tiles + 1 KiB
    halo in 16-byte chunks) must give the oracle's tables, as must the stream edge cases."""
    from telomeri_b200 import pipeline
    rec = "read%05d 9000 10 8990 + ctg1 50000 %d %d 8000 8980 0 tp:A:S cm:i:1 s1:i:1 dv:f:0.1\n"
    body = "".join(rec % (i, i % 7, 8980 + i % 7) for i in range(1200))   # ~ 3.1 tiles
    texts = {"whole": body, "no_newline": body.rstrip("\n"), "crlf": body[:5000].replace(" ", "\r\n"), "nonnumeric": body[:3000] + "x" + body[3000:],
             "one": rec % (1, 0, 8980), "space": " \n"}
    tile, halo = 30720, 1024
    for cut in (tile - 17, tile - 1, tile, tile + 1, tile + 15, tile + 16, tile + halo - 1, tile + halo, tile + halo + 1, 2 * tile, 2 * tile + 5, 3 * tile - 16, 3 * tile):
        texts[f"cut{cut}"] = body[:cut]
    assert len(body) > 3 * tile
    for name, text in texts.items():
        d = tmp_path / name
        d.mkdir()
        (d / "rr_overlaps.paf").write_text("")
        (d / "cr_overlaps.paf").write_text(text, newline="")
        ora = common.run_dump(oracle_bin, d, d / "o.bin", ["--mode", "max"], extra=["--stage", "graph"])
        gpu = pipeline.run_hot_path(d, ["--mode", "max"], stages="graph", loader="device_view")
        common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS, name)
    # a whole dataset through the views, walks and groups included
    d = common.make_dataset("small", tmp_path / "small")
    ora = common.run_dump(oracle_bin, d, d / "o.bin", [], extra=["--rng", "philox", "--seed", "5"])
    gpu = pipeline.run_hot_path(d, [], seed=5, loader="device_view")
    common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS + PATH_SECTIONS + GROUP_SECTIONS, "small through views")


def test_warp_parse_variant_matches_oracle(oracle_bin, tmp_path, monkeypatch):
    """paf_parse_warp_kernel (parse_variant = 1: one 3 KiB sub-tile per warp, no CTA barrier; opt-in, see DESIGN.md K1+K2)
    must give the same tables as the default kernel: stream edge cases, in-place loads cut at every boundary, a dataset."""
    monkeypatch.setenv("HERA_PARSE_VARIANT", "1")
    for name, fn in (("edge", test_reader_edge_cases_gpu), ("view", test_device_view_loads_match_oracle)):
        d = tmp_path / name
        d.mkdir()
        fn(oracle_bin, d)


def test_walk_bitmap_in_global_memory(oracle_bin, datasets, tmp_path):
    """Graphs above 65536 nodes keep the visited bitmaps of the Monte-Carlo walks in global memory; force that path."""
    from telomeri_b200 import capi, pipeline
    golden = common.load_golden("small_default")
    ds = datasets[golden["dataset"]]
    ora = common.run_dump(oracle_bin, ds, tmp_path / "o.bin", golden["options"], extra=["--rng", "philox", "--seed", "77"])
    with capi.Hera(0) as h:
        h.set_option("walk_smem_bitmap", 0)
        gpu = pipeline.run_hot_path(ds, golden["options"], seed=77, hera=h)
    common.assert_sections_equal(ora, gpu, PATH_SECTIONS + GROUP_SECTIONS, "global bitmap")


@pytest.mark.parametrize("batched", [False, True])
def test_walk_scratch_overflow_is_rerun_with_full_capacity(batched, oracle_bin, datasets, tmp_path):
    """The per-warp walk scratch holds 16384 nodes by default; a walk that visits more makes the call run again with room
    for every node.  Forced here with a capacity of 8 nodes: most walks overflow, the paths must still be the oracle's;
    the second run of the same context goes straight to the full capacity (and the adaptive warp shares are in use)."""
    from telomeri_b200 import capi, pipeline
    golden = common.load_golden("small_default")
    ds = datasets[golden["dataset"]]
    ora = common.run_dump(oracle_bin, ds, tmp_path / "o.bin", golden["options"], extra=["--rng", "philox", "--seed", "78"])
    with capi.Hera(0) as h:
        h.set_option("walk_cap_small", 8)
        for _ in range(2):
            h.graph_reset()
            h.paths_reset()
            gpu = pipeline.run_hot_path(ds, golden["options"], seed=78, hera=h, batched=batched)
            common.assert_sections_equal(ora, gpu, PATH_SECTIONS + GROUP_SECTIONS, "small walk scratch")


@pytest.mark.parametrize("case", ["small_default", "mixed_sqrt", "bundled_min"])
def test_batched_build_equals_separate_calls(case, oracle_bin, datasets, tmp_path):
    """hera_paths_build (all walk kernels in flight together) leaves paths_ exactly as the separate calls do."""
    from telomeri_b200 import pipeline
    golden = common.load_golden(case)
    ds = datasets[golden["dataset"]]
    ora = common.run_dump(oracle_bin, ds, tmp_path / "o.bin", golden["options"], extra=["--rng", "philox", "--seed", "31"])
    gpu = pipeline.run_hot_path(ds, golden["options"], seed=31, batched=True)
    common.assert_sections_equal(ora, gpu, PATH_SECTIONS + GROUP_SECTIONS, case + " batched")


def test_rank_and_canon_of_long_lists(oracle_bin, tmp_path):
    """K4b has code paths by list length (<= 64: all pairs, <= 512: one warp per node with sampled buckets, <= 4096: the
    same with the whole CTA, longer: counting from global memory); hubs of 40 / 300 / 700 / 2600 / 4500 reads with many
    equal scores, parallel edges (the same read twice) and all four metrics go through each of them."""
    import random
    from telomeri_b200 import pipeline
    rng = random.Random(5)
    lines = []
    for hub, n in (("ctgA", 40), ("ctgB", 700), ("ctgC", 2600), ("ctgD", 300), ("ctgE", 4500)):
        for k in range(n):
            name = f"{hub}r{rng.randrange(n * 3 // 4)}"          # repeats: parallel edges -> canon != slot
            qs = rng.choice((10, 20, 30))
            matches = rng.choice((900, 940, 960, 970))            # few distinct scores: long runs of ties
            ts = rng.randrange(0, 3) * 100
            lines.append(f"{name} 1000 {qs} 990 {rng.choice('+-')} {hub} 50000 {ts} {ts + 980} {matches} 980 0 tp:A:S cm:i:1 s1:i:1 dv:f:0.1\n")
    d = tmp_path / "hubs"
    d.mkdir()
    (d / "rr_overlaps.paf").write_text("")
    (d / "cr_overlaps.paf").write_text("".join(lines))
    for metrics in ("es,os", "es_sqrt,os_sqrt"):
        opts = ["--mode", "max", "--metrics", metrics]
        ora = common.run_dump(oracle_bin, d, d / "o.bin", opts, extra=["--stage", "graph"])
        gpu = pipeline.run_hot_path(d, opts, stages="graph")
        ranks = [f"rank_{METRIC_ID[m]}" for m in metrics.split(",")]
        deg = np.diff(ora["adj_off"].astype(np.int64))
        assert deg.max() > 4096 and ((deg > 512) & (deg <= 4096)).any() and ((deg > 64) & (deg <= 512)).any()
        assert (ora["canon"] != np.arange(ora["canon"].size)).any()
        common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS + ranks, f"hubs {metrics}")

