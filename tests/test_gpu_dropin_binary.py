"""The drop-in command-line binary (reference class surface + CLI over the C ABI) against the
reference binary and the oracle: stdout narrative and scaffold FASTA.  Runs on the B200 box."""
import hashlib
import re
import subprocess
from pathlib import Path

import pytest

import common

pytestmark = pytest.mark.gpu

BIN = common.ROOT / "telomeri_b200" / "bin"
REF = common.REF_DIR / "Telomeri"


def narrative(text):
    """stdout without wall-clock figures."""
    text = re.sub(r"Done \([0-9.e+-]+s\)", "Done (T)", text)
    return re.sub(r"Total time: [0-9.e+-]+s", "Total time: T", text)


def run(exe, ds, out, opts=(), env=None):
    cmd = [str(exe)] + list(opts) + [str(ds / "rr_overlaps.paf"), str(ds / "cr_overlaps.paf"), str(ds / "reads.fasta"), str(ds / "contigs.fasta"), str(out)]
    return subprocess.run(cmd, capture_output=True, text=True, env=env)


@pytest.fixture(scope="module")
def binaries():
    from telomeri_b200 import build
    build.build_library()
    build.build_host()
    return [b for b in (BIN / "Telomeri", BIN / "Telomeri_refmain") if b.exists()]


@pytest.mark.skipif(not REF.exists(), reason="oracle/_ref/Telomeri not built")
@pytest.mark.parametrize("ds_name", ["small", "mixed"])
def test_deterministic_run_equals_reference_binary(ds_name, binaries, datasets, tmp_path):
    """--rb-att 0 switches Monte Carlo off, so the UNMODIFIED reference binary must agree on everything:
    every stdout line (counts, stats, consensus lengths, scaffold node names) and the FASTA bytes."""
    ds = datasets[ds_name]
    ref = run(REF, ds, tmp_path / "out.fa", ["--rb-att", "0"])
    want = (tmp_path / "out.fa").read_bytes() if ref.returncode == 0 else None
    for exe in binaries:
        out = tmp_path / "out.fa"   # same name: it is echoed on stdout
        out.unlink(missing_ok=True)
        got = run(exe, ds, out, ["--rb-att", "0"])
        assert got.returncode == ref.returncode, got.stderr
        assert narrative(got.stdout) == narrative(ref.stdout)
        if want is not None:
            assert out.read_bytes() == want


@pytest.mark.parametrize("case", ["small_default", "mixed_default"])
def test_full_run_equals_oracle_fasta(case, binaries, datasets, oracle_bin, tmp_path):
    """With Monte Carlo on (Philox, seed from HERA_SEED) the scaffold FASTA must equal the oracle's, which in
    turn equals the reference replaying the same draws (tests/test_oracle_vs_reference.py)."""
    import os
    golden = common.load_golden(case)
    ds = datasets[golden["dataset"]]
    common.run_dump(oracle_bin, ds, tmp_path / "o.bin", golden["options"], with_fasta=True, extra=["--rng", "philox", "--seed", "42"])
    want = Path(str(tmp_path / "o.bin") + ".fasta")
    cli = [o.replace("_", "-") for o in golden["options"]]
    for exe in binaries:
        out = tmp_path / f"{exe.name}.fa"
        got = run(exe, ds, out, cli, env=dict(os.environ, HERA_SEED="42"))
        assert got.returncode == 0, got.stderr
        assert hashlib.md5(out.read_bytes()).hexdigest() == hashlib.md5(want.read_bytes()).hexdigest()


def test_cli_errors_match_reference_contract(binaries, datasets, tmp_path):
    ds = datasets["small"]
    exe = binaries[0]
    assert subprocess.run([str(exe)], capture_output=True).returncode == 1                  # too few arguments
    bad = subprocess.run([str(exe), "nope.paf", "nope2.paf", "a", "b", "c"], capture_output=True, text=True)
    assert bad.returncode == 1 and "Cannot open file: nope.paf" in bad.stderr
    none = run(exe, ds, tmp_path / "x.fa", ["--filter-max", "--min-oll", "2000000"])
    assert none.returncode == 1 and "No consensus found!" in none.stderr
