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


def load_golden(name):
    with open(GOLDEN / f"{name}.json") as f:
        return json.load(f)


def assert_matches_golden(dump, golden, what=""):
    for k, meta in golden["sections"].items():
        assert k in dump, f"{what}: section {k} missing"
        got = {"dtype": str(dump[k].dtype), "n": int(dump[k].size), "sha": section_digest(dump[k])}
        assert got == meta, f"{what}: section {k} differs from golden: {got} vs {meta}"
