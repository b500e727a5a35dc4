"""Shared helpers for the test-suite: build/run the oracle binaries, datasets, dump comparison."""
import gzip
import hashlib
import json
import os
import subprocess
from pathlib import Path

import numpy as np

from heradump import read_dump

ROOT = Path(__file__).resolve().parent.parent
ORACLE_DIR = ROOT / "oracle"
ORACLE_BIN = ORACLE_DIR / "bin" / "hera_oracle"
REF_DIR = ORACLE_DIR / "_ref"
GOLDEN = ROOT / "tests" / "golden"
GEN_BIN = ROOT / "tools" / "bin" / "gen_dataset"
REFERENCE_TREE = Path("/root/reference")

# Datasets used across tests: name -> gen_dataset options (deterministic).
DATASETS = {
    # SURVEY.md Appendix A shape: 600 kb genome, 3 contigs, 1 800 reads of ~10 kb.
    "small": ["--genome", "600000", "--read-len", "10000", "--coverage", "30", "--contigs", "3", "--gap", "50000",
              "--families", "2", "--copies", "2", "--seed", "7"],
    # a little strand / jitter variety, more contigs
    "mixed": ["--genome", "400000", "--read-len", "8000", "--coverage", "20", "--contigs", "4", "--gap", "30000",
              "--families", "3", "--copies", "2", "--seed", "11", "--jitter", "40", "--minus-frac", "0.3"],
}


# Hand-made "star" datasets: start anchor ctgB -> read r_i -> end anchor ctgA, one read per path, so the path lengths
# len(ctgB) + qs_i (reference src/Path.cpp:6-16: the last edge is excluded) and their frequencies are chosen here.
# Each read gives exactly one greedy path per metric (reference src/PathManager.cpp:179-430), whatever the RNG.
STAR_DATASETS = {
    # --rb-att 1: 4 greedy + 2 Monte-Carlo paths of ~1.2e9 from ctgB to ctgA; std::accumulate(..., 0, ...) wraps int32
    # (reference src/PathGroup.cpp:66-68): 7.2e9 mod 2^32 ~ 2.905e9 >= 2^31 -> negative int -> avg ~ 2^64/6 -> consensus = last path
    "star_wrap_neg": {"lb": 1200000000, "qs": [4000, 4100]},
    # 6 greedy + 2 Monte-Carlo paths of ~1.1e9: 8.8e9 mod 2^32 ~ 2.1e8 > 0 -> avg far too small -> consensus = first path
    "star_wrap_pos": {"lb": 1100000000, "qs": [4000, 4100, 4200]},
    # windows of 100 from min_len: {4000 x2, 4050 x4} {4100 x8} {4200 x10}: valley = window 0, its lowest-frequency
    # length is min_len itself -> border <= first length -> the first group is {v[0]} although v[1] has the same length
    # (reference src/PathManager.cpp:568-579 starts the scan at begin + 1)
    "star_border": {"lb": 100000, "qs": [4000] + [4050] * 2 + [4100] * 4 + [4200] * 5},
    # three groups with a consensus between one anchor pair, so the inter-group loop "2 * longer <= shorter" (reference
    # src/Main.cpp:318-330) runs and takes both branches.  Negative qs = a forked read: its extension-score walk goes
    # r -> ctgA, its overlap-score walk r -> s -> ctgA (4000 longer), so those node sequences occur once (valid number 1)
    # while a plain read's two walks are equal (valid number 2).  Lengths: 4000 x10 | 4150 x2, 4300 x10 | 4450 x2, 4600 x5, 8600 x5
    "star_tournament": {"lb": 100000, "qs": [4000] * 5 + [4150] + [4300] * 5 + [4450] + [-4600] * 5},
}


def make_star_dataset(name, out_dir):
    spec = STAR_DATASETS[name]
    out_dir = Path(out_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    lb, la, rl = spec["lb"], 50000, 10000
    tail = "0\ttp:A:S\tcm:i:1\ts1:i:1\tdv:f:0.1000\n"
    lines = []
    forked = [i for i, qs in enumerate(spec["qs"]) if qs < 0]
    spec = dict(spec, qs=[abs(q) for q in spec["qs"]])
    for i, qs in enumerate(spec["qs"]):   # read suffix over the prefix of ctgB: ctgB becomes an anchor as the target of a cr record
        ol = rl - qs
        lines.append(f"r{i:03d}\t{rl}\t{qs}\t{rl}\t+\tctgB\t{lb}\t0\t{ol}\t{ol * 3 // 10}\t{ol}\t" + tail)
    for i, qs in enumerate(spec["qs"]):   # suffix of ctgA over the read prefix: ctgA is the query ("ctg...") -> anchor
        lines.append(f"ctgA\t{la}\t{la - 5000}\t{la}\t+\tr{i:03d}\t{rl}\t0\t5000\t5000\t5000\t" + tail)
    for i in forked:                      # s_i: best overlap score at r_i, and itself best connected to ctgA
        lines.append(f"s{i:03d}\t{rl}\t4000\t{rl}\t+\tr{i:03d}\t{rl}\t0\t6000\t5400\t6000\t" + tail)
        lines.append(f"ctgA\t{la}\t{la - 8000}\t{la}\t+\ts{i:03d}\t{rl}\t0\t8000\t8000\t8000\t" + tail)
    (out_dir / "rr_overlaps.paf").write_text("")
    (out_dir / "cr_overlaps.paf").write_text("".join(lines))
    if lb <= 1000000:   # sequences for the Scaffolder (two lines per record, reference src/Scaffolder.cpp:24-38)
        rng = np.random.default_rng(len(spec["qs"]))
        seq = lambda n: "".join(np.array(list("ACGT"))[rng.integers(0, 4, n)])
        (out_dir / "contigs.fasta").write_text(f">ctgA\n{seq(la)}\n>ctgB\n{seq(lb)}\n")
        names = [f"r{i:03d}" for i in range(len(spec["qs"]))] + [f"s{i:03d}" for i in forked]
        (out_dir / "reads.fasta").write_text("".join(f">{n}\n{seq(rl)}\n" for n in names))
    return out_dir


def windowed_pairs(dump, len_thr):
    """Number of anchor pairs whose paths enter the window branch of constructGroups (max - min >= len_threshold,
    reference src/PathManager.cpp:521-527)."""
    n = 0
    off, srt, ln = dump["pair_off"], dump["pair_sorted"], dump["u_len"]
    for p in range(len(dump["pair_a1"])):
        ids = srt[int(off[p]):int(off[p + 1])]
        n += int(int(ln[ids[-1]]) - int(ln[ids[0]]) >= len_thr)
    return n


def build_oracle():
    subprocess.run(["make", "-s", "-C", str(ORACLE_DIR), "bin/hera_oracle"], check=True)
    if REFERENCE_TREE.exists():
        subprocess.run(["make", "-s", "-C", str(ORACLE_DIR), "ref"], check=True)


def build_generator():
    src = ROOT / "tools" / "gen_dataset.cpp"
    if not GEN_BIN.exists() or GEN_BIN.stat().st_mtime < src.stat().st_mtime:
        GEN_BIN.parent.mkdir(parents=True, exist_ok=True)
        subprocess.run(["g++", "-O2", "-std=c++17", "-o", str(GEN_BIN), str(src)], check=True)


def make_dataset(name, out_dir):
    out_dir = Path(out_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    build_generator()
    subprocess.run([str(GEN_BIN), "--out", str(out_dir)] + DATASETS[name], check=True, stdout=subprocess.DEVNULL)
    return out_dir


def bundled_dataset(out_dir):
    """Config 1: the reference's bundled contig-read PAF plus an empty read-read PAF."""
    out_dir = Path(out_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    with gzip.open(GOLDEN / "ecoli_cr_overlaps.paf.gz", "rb") as f:
        (out_dir / "cr_overlaps.paf").write_bytes(f.read())
    (out_dir / "rr_overlaps.paf").write_bytes(b"")
    return out_dir


def run_dump(binary, ds_dir, out_path, options=(), with_fasta=False, extra=()):
    ds_dir = Path(ds_dir)
    cmd = [str(binary)] + list(options) + list(extra) + [str(ds_dir / "rr_overlaps.paf"), str(ds_dir / "cr_overlaps.paf"), str(out_path)]
    if with_fasta:
        cmd += [str(ds_dir / "reads.fasta"), str(ds_dir / "contigs.fasta"), str(out_path) + ".fasta"]
    subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL)
    return read_dump(out_path)


def section_digest(arr):
    return hashlib.sha256(np.ascontiguousarray(arr).tobytes()).hexdigest()[:32]


def digest_dump(dump, skip=("replay_",)):
    return {k: {"dtype": str(v.dtype), "n": int(v.size), "sha": section_digest(v)} for k, v in dump.items()
            if not k.startswith(tuple(skip))}


def assert_sections_equal(a, b, names=None, what=""):
    names = names if names is not None else [k for k in a if k in b]
    assert names, "nothing to compare"
    for k in names:
        assert k in a and k in b, f"{what}: section {k} missing"
        x, y = a[k], b[k]
        assert x.dtype == y.dtype and x.shape == y.shape, f"{what}: {k} dtype/shape {x.dtype}{x.shape} vs {y.dtype}{y.shape}"
        if not np.array_equal(x.view(np.uint8), y.view(np.uint8)):
            bad = np.nonzero(x != y)[0][:5]
            raise AssertionError(f"{what}: section {k} differs at {bad}: {x[bad]} vs {y[bad]}")


def golden_cases():
    """Fixture names under tests/golden (bench_digests.json is bench.py's digest golden, not a fixture)."""
    return sorted(p.stem for p in GOLDEN.glob("*.json") if not p.stem.startswith(("bench_", "config_")))


def load_golden(name):
    with open(GOLDEN / f"{name}.json") as f:
        return json.load(f)


def assert_matches_golden(dump, golden, what=""):
    for k, meta in golden["sections"].items():
        assert k in dump, f"{what}: section {k} missing"
        got = {"dtype": str(dump[k].dtype), "n": int(dump[k].size), "sha": section_digest(dump[k])}
        assert got == meta, f"{what}: section {k} differs from golden: {got} vs {meta}"


# Fixtures that must enter the window / border / multi-group code; the tests assert it so that a change of generator or
# options cannot silently turn them back into one-group cases.
WINDOW_CASES = ["bundled_min_win42", "bundled_min_win7", "mixed_win13", "small_win", "star_border", "star_tournament"]
WRAP_CASES = ["star_wrap_neg", "star_wrap_pos"]


def option_value(opts, key, default):
    return type(default)(opts[opts.index(key) + 1]) if key in opts else default


def groups_with_consensus_per_pair(dump):
    n = np.zeros(len(dump["pair_a1"]), dtype=np.int64)
    for g, p in enumerate(dump["grp_pair"]):
        n[int(p)] += int(dump["grp_consensus"][g] >= 0)
    return n


def assert_enters_window_code(dump, options, what=""):
    """The dump (oracle, reference or CUDA path) went through PathWindow / getBorderPathLengths / the group split."""
    wp = windowed_pairs(dump, option_value(options, "--len-thr", 4700000))
    assert wp > 0, f"{what}: no anchor pair reaches the window branch"
    assert dump["grp_pair"].size > dump["pair_a1"].size, f"{what}: no pair was split into several groups"
