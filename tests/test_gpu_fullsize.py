"""BASELINE.json's larger configurations through the CUDA path, bit-exact against the oracle on the same
generated input: configs[1] (E. coli scale, the bench workload, ~1.6 M PAF records) on one GPU and as a
gang of two ranks, configs[2] (yeast scale, 120 contigs, ~8 M records) as a gang of two (BASELINE: 1/2 GPU).
Monte-Carlo attempts are reduced (--rb-att) so that the CPU oracle finishes in seconds; everything else
is the reference default.  Size-independent properties are checked on top (CSR symmetry, rank
permutations, unique idempotence)."""
import subprocess

import numpy as np
import pytest

import common
from test_gpu_parity import GRAPH_SECTIONS, GROUP_SECTIONS, PATH_SECTIONS
from test_gpu_sharded import gang_devices

pytestmark = pytest.mark.gpu

CONFIGS = {
    "ecoli": (["--genome", "4600000", "--read-len", "15000", "--coverage", "50", "--contigs", "10", "--gap", "40000", "--families", "10", "--copies", "3",
               "--min-ovl", "2000", "--no-fasta", "--seed", "2"], ["--rb-att", "100"]),
    "yeast": (["--genome", "12000000", "--read-len", "20000", "--coverage", "60", "--contigs", "120", "--gap", "30000", "--families", "56", "--copies", "4",
               "--min-ovl", "1000", "--no-fasta", "--seed", "3"], ["--rb-att", "20"]),
}


@pytest.fixture(scope="module")
def big(tmp_path_factory, oracle_bin):
    cache = {}

    def get(name):
        if name not in cache:
            d = tmp_path_factory.mktemp(name)
            gen, opts = CONFIGS[name]
            common.build_generator()
            subprocess.run([str(common.GEN_BIN), "--out", str(d)] + gen, check=True, stdout=subprocess.DEVNULL)
            ora = common.run_dump(oracle_bin, d, d / "o.bin", opts, extra=["--rng", "philox", "--seed", "17"])
            (d / "o.bin").unlink()
            cache[name] = (d, opts, ora)
        return cache[name]
    return get


def check_properties(g):
    S, E, N = g["adj_rec"].size, g["e_q"].size, g["node_len"].size
    assert S == 2 * E and g["adj_off"][-1] == S and np.all(np.diff(g["adj_off"].astype(np.int64)) >= 0)
    # every kept overlap sits once in its target's list (side 0) and once in its query's list (side 1), in record order
    owner = np.repeat(np.arange(N, dtype=np.uint32), np.diff(g["adj_off"].astype(np.int64)))
    side = g["adj_side"].astype(bool)
    assert np.array_equal(owner[~side], g["e_t"][g["adj_rec"][~side]]) and np.array_equal(owner[side], g["e_q"][g["adj_rec"][side]])
    assert np.array_equal(np.sort(g["adj_rec"][~side]), np.arange(E, dtype=np.uint32))
    same_owner = owner[1:] == owner[:-1]
    assert np.all(g["adj_rec"][1:][same_owner] >= g["adj_rec"][:-1][same_owner])
    for m, key in ((0, "a_es"), (1, "a_os")):
        r = g[f"rank_{m}"]
        assert np.array_equal(np.sort(r), np.arange(S, dtype=np.uint32))           # a permutation of the slots ...
        assert np.array_equal(owner[r], owner)                                      # ... inside every node's segment ...
        v = g[key][r]
        assert np.all((v[1:] <= v[:-1]) | ~same_owner)                              # ... in descending metric order


def test_ecoli_scale_single_gpu(big):
    from telomeri_b200 import pipeline
    ds, opts, ora = big("ecoli")
    gpu = pipeline.run_hot_path(ds, opts, seed=17, batched=True)
    common.assert_sections_equal(ora, gpu, GRAPH_SECTIONS + ["rank_0", "rank_1"] + PATH_SECTIONS + GROUP_SECTIONS, "configs[1]")
    check_properties(gpu)
    assert gpu["load_ctr"].sum() > 1_500_000


def test_ecoli_scale_two_ranks(big):
    from telomeri_b200 import capi, pipeline
    ds, opts, ora = big("ecoli")
    with capi.LocalGang(2, gang_devices(2)) as gang:
        outs = gang.each(lambda r, h: pipeline.run_hot_path(ds, opts, seed=17, hera=h, batched=True) if r == 0 else
                         {k: v for k, v in pipeline.run_hot_path(ds, opts, seed=17, hera=h, batched=True).items() if k in ("u_len", "pair_consensus", "e_os")})
    common.assert_sections_equal(ora, outs[0], GRAPH_SECTIONS + ["rank_0", "rank_1"] + PATH_SECTIONS + GROUP_SECTIONS, "configs[1] rank 0/2")
    common.assert_sections_equal(ora, outs[1], ["u_len", "pair_consensus", "e_os"], "configs[1] rank 1/2")


def test_yeast_scale_two_ranks(big):
    from telomeri_b200 import capi, pipeline
    ds, opts, ora = big("yeast")
    with capi.LocalGang(2, gang_devices(2)) as gang:
        outs = gang.each(lambda r, h: pipeline.run_hot_path(ds, opts, seed=17, hera=h, batched=True) if r == 0 else
                         {k: v for k, v in pipeline.run_hot_path(ds, opts, seed=17, hera=h, batched=True).items() if k in ("u_len", "pair_consensus", "e_os")})
    common.assert_sections_equal(ora, outs[0], GRAPH_SECTIONS + ["rank_0", "rank_1"] + PATH_SECTIONS + GROUP_SECTIONS, "configs[2] rank 0/2")
    common.assert_sections_equal(ora, outs[1], ["u_len", "pair_consensus", "e_os"], "configs[2] rank 1/2")
    check_properties(outs[0])
    assert outs[0]["load_ctr"].sum() > 6_000_000 and outs[0]["node_anchor"].sum() == 120
