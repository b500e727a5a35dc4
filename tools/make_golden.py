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
    ("mixed_default", "mixed", ["--rb-att", "500"], True),
    ("mixed_sqrt", "mixed", ["--rb-att", "300", "--metrics", "es_sqrt,os_sqrt", "--mode", "max", "--max-ohp", "0.9"], True),
]


def main():
    common.build_oracle()
    ref = common.REF_DIR / "ref_dump"
    with tempfile.TemporaryDirectory() as tmp:
        tmp = Path(tmp)
        dirs = {"bundled": common.bundled_dataset(tmp / "bundled")}
        for name in common.DATASETS:
            dirs[name] = common.make_dataset(name, tmp / name)
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
