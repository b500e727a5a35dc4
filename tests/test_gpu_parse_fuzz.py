"""Randomised PAF texts through the CUDA tokenizer against the oracle (the reference reads a
whitespace token stream, 16 tokens per record).  The texts are built to hit what the tiled, pipelined parse
kernel has to get right: records and names that straddle tile and halo boundaries, tiles with more records than
threads, runs of white space longer than a mask word, files around multiples of the tile size, records that make
formatted extraction fail in the middle of a file, and the same texts cut into byte ranges for 2 and 3 ranks."""
import random

import numpy as np
import pytest

import common
from test_gpu_parity import GRAPH_SECTIONS

pytestmark = pytest.mark.gpu

TILE = 30720


def random_text(rng, target_bytes, flavour):
    out = []
    size = 0
    k = 0
    while size < target_bytes:
        k += 1
        qname = "r%d" % rng.randrange(400) if flavour != "long_names" else "read_" + "x" * rng.randrange(1, 900) + str(rng.randrange(50))
        tname = ("ctg%d" % rng.randrange(4)) if rng.random() < 0.3 else "r%d" % rng.randrange(400)
        qlen = rng.randrange(900, 20000)
        tlen = rng.randrange(900, 20000)
        qs = rng.randrange(0, qlen // 3)
        qe = rng.randrange(qlen // 2, qlen)
        ts = rng.randrange(0, tlen // 3)
        te = rng.randrange(tlen // 2, tlen)
        fields = [qname, str(qlen), str(qs), str(qe), rng.choice("+-"), tname, str(tlen), str(ts), str(te),
                  str(rng.randrange(1, min(qe - qs, te - ts) + 1)), str(max(qe - qs, te - ts)), str(rng.randrange(0, 61)),
                  "tp:A:S", "cm:i:%d" % rng.randrange(100), "s1:i:%d" % rng.randrange(1000), "dv:f:0.%d" % rng.randrange(100)]
        if flavour == "tiny":       # more records per tile than threads
            fields = ["a%d" % rng.randrange(30), "9", "1", "8", "+", "b%d" % rng.randrange(30), "9", "1", "8", "7", "7", "0", "x", "y", "z", "w"]
        if flavour == "signed" and rng.random() < 0.2:
            fields[2] = "+" + fields[2]
            fields[7] = "000000000" + fields[7]
        if flavour == "irregular" and k == 150:
            fields[3] = fields[3] + "abc"   # int token with trailing garbage: the stream leaves the 16-token grid here
        seps = []
        for _ in range(15):
            r = rng.random()
            if flavour == "wide_space" and r < 0.05:
                seps.append(" " * rng.randrange(33, 200))
            elif r < 0.1:
                seps.append("\t")
            elif r < 0.15:
                seps.append("  \t ")
            else:
                seps.append(" ")
        line = "".join(f + s for f, s in zip(fields, seps + [""])) + rng.choice(["\n", "\n", "\r\n", "\n\n", " \n"])
        out.append(line)
        size += len(line)
    return "".join(out)


CASES = [("plain", 2 * TILE + 17), ("plain", 3 * TILE), ("plain", TILE - 1), ("tiny", 2 * TILE + 5), ("long_names", 4 * TILE),
         ("wide_space", 3 * TILE + 999), ("signed", 2 * TILE), ("irregular", 3 * TILE + 1)]


@pytest.mark.parametrize("idx", range(len(CASES)))
def test_random_texts_match_oracle(idx, oracle_bin, tmp_path):
    from telomeri_b200 import capi, pipeline
    from test_gpu_sharded import gang_devices
    flavour, size = CASES[idx]
    rng = random.Random(1000 + idx)
    text = random_text(rng, size, flavour)
    if flavour == "plain":       # cut to the exact byte size: the last record is incomplete more often than not
        text = text[:size]
    d = tmp_path / f"{flavour}_{size}"
    d.mkdir()
    (d / "rr_overlaps.paf").write_text(text[: len(text) // 2], newline="")
    (d / "cr_overlaps.paf").write_text(text, newline="")
    opts = ["--mode", "max", "--max-ohp", "50.0"]
    ora = common.run_dump(oracle_bin, d, d / "o.bin", opts, extra=["--stage", "graph"])
    gpu = pipeline.run_hot_path(d, opts, stages="graph")
    common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS, f"{flavour} {size}")
    assert int(ora["load_ctr"][1]) > 0
    for nranks in (2, 3):
        with capi.LocalGang(nranks, gang_devices(nranks)) as gang:
            outs = gang.each(lambda r, h: pipeline.run_hot_path(d, opts, stages="graph", hera=h))
        for r, out in enumerate(outs):
            common.assert_sections_equal(ora, out, GRAPH_SECTIONS, f"{flavour} {size} rank {r}/{nranks}")
