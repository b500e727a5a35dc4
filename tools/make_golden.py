#!/usr/bin/env python
"""Regenerates tests/golden/*.json from the UNMODIFIED reference (oracle/_ref/ref_dump, built from
/root/reference by oracle/Makefile).  Each fixture records, per dump section, dtype / length /
sha256 of the reference's bytes, plus a few human-readable known answers.
Run here (the reference tree is needed); the JSON files are committed and travel to the GPU box.
"""
import hashlib
import json
import subprocess
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))
import common  # noqa: E402
from heradump import text  # noqa: E402

CASES = [
    # (fixture name, dataset, ref_dump options, with_fasta)
    ("bundled_avg", "bundled", [], False),
    ("bundled_sum", "bundled", ["--mode", "sum"], False),
    ("bundled_max", "bundled", ["--mode", "max"], False),
    ("bundled_min", "bundled", ["--mode", "min"], False),
    ("bundled_none", "bundled", ["--mode", "max", "--min-oll", "2000", "--min-olp", "0.05", "--max-ohl", "1000"], False),
    ("bundled_min_cut", "bundled", ["--mode", "min", "--test-load-num", "9000", "--rb-att", "300"], False),
    ("small_default", "small", [], True),
    ("small_windows", "small", ["--len-thr", "200000", "--w-size", "500", "--rb-att", "400"], True),
    # cases that ENTER the window / border / multi-group code (reference src/PathManager.cpp:527-580, src/PathWindow.cpp,
    # src/Main.cpp:306-332); "known" records windowed_pairs and groups > pairs, and the tests assert both
    ("bundled_min_win42", "bundled", ["--mode", "min", "--len-thr", "500000", "--w-size", "10000", "--rb-att", "300"], False),
    ("bundled_min_win7", "bundled", ["--mode", "min", "--len-thr", "1000000", "--w-size", "50000", "--rb-att", "300", "--r-thr", "0.5"], False),
    ("mixed_win13", "mixed", ["--len-thr", "50000", "--w-size", "1000", "--rb-att", "300"], True),
    ("small_win", "small", ["--len-thr", "500", "--w-size", "100", "--rb-att", "50"], True),
    ("star_wrap_neg", "star_wrap_neg", ["--len-thr", "3000000000", "--rb-att", "1"], False),
    ("star_wrap_pos", "star_wrap_pos", ["--len-thr", "3000000000", "--rb-att", "1"], False),
    ("star_border", "star_border", ["--len-thr", "150", "--w-size", "100", "--rb-att", "5"], False),
    ("star_tournament", "star_tournament", ["--len-thr", "150", "--w-size", "100", "--rb-att", "5"], False),
    ("mixed_default", "mixed", ["--rb-att", "500"], True),
    ("mixed_sqrt", "mixed", ["--rb-att", "300", "--metrics", "es_sqrt,os_sqrt", "--mode", "max", "--max-ohp", "0.9"], True),
]


def len_thr(opts):
    return int(opts[opts.index("--len-thr") + 1]) if "--len-thr" in opts else 4700000


def main():
    common.build_oracle()
    ref = common.REF_DIR / "ref_dump"
    with tempfile.TemporaryDirectory() as tmp:
        tmp = Path(tmp)
        dirs = {"bundled": common.bundled_dataset(tmp / "bundled")}
        for name in common.DATASETS:
            dirs[name] = common.make_dataset(name, tmp / name)
        for name in common.STAR_DATASETS:
            dirs[name] = common.make_star_dataset(name, tmp / name)
        for name, ds, opts, with_fasta in CASES:
            dump = common.run_dump(ref, dirs[ds], tmp / f"{name}.bin", opts, with_fasta)
            fixture = {
                "generated_by": "tools/make_golden.py from oracle/_ref/ref_dump (unmodified reference classes, std::mt19937 Monte Carlo)",
                "dataset": ds, "options": opts,
                "known": {
                    "load_ctr": dump["load_ctr"].tolist(), "load_cum_edges": dump["load_cum_edges"].tolist(),
                    "nodes": int(dump["node_len"].size), "anchors": int(dump["node_anchor"].sum()),
                    "call_found": dump["call_found"].tolist(), "unique_paths": int(dump["u_len"].size),
                    "pairs": int(dump["pair_a1"].size), "groups": int(dump["grp_pair"].size),
                    "windowed_pairs": common.windowed_pairs(dump, len_thr(opts)),
                    "max_group_sum": max([int(dump["u_len"][dump["grp_pig"][int(a):int(b)]].sum())
                                          for a, b in zip(dump["grp_pig_off"][:-1], dump["grp_pig_off"][1:])] or [0]),
                    "have_scaffold": int(dump["have_scaffold"][0]),
                    "graph_stats": text(dump["graph_stats"]), "path_stats": text(dump["path_stats"]),
                },
                "sections": common.digest_dump(dump),
            }
            fa = Path(str(tmp / f"{name}.bin") + ".fasta")
            if with_fasta and fa.exists():
                fixture["fasta_md5"] = hashlib.md5(fa.read_bytes()).hexdigest()
            with open(common.GOLDEN / f"{name}.json", "w") as f:
                json.dump(fixture, f, indent=1)
            print(name, fixture["known"]["load_cum_edges"], fixture["known"]["call_found"], fixture["known"]["unique_paths"],
                  fixture["known"]["groups"], fixture["known"]["have_scaffold"], fixture.get("fasta_md5"))


if __name__ == "__main__":
    main()
