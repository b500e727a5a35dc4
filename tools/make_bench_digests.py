#!/usr/bin/env python
"""Writes tests/golden/bench_digests.json: for every bench workload (configs[1], configs[3]) the hera_digest words of a
single-GPU run of the CUDA path that was compared, table by table and bit for bit, with the CPU oracle on the same
dataset, options (--rb-att 2000, metrics ES + OS) and Philox seed as bench.py uses.  bench.py (any number of GPUs)
then proves parity of ITS run by comparing digests -- the oracle needs minutes at this size, the digest microseconds.
Run on the GPU box (the oracle walks run on HERA_ORACLE_THREADS host threads):
    python tools/make_bench_digests.py [ecoli] [dmel]"""
import json
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import bench  # noqa: E402
import common  # noqa: E402
from telomeri_b200 import capi, pipeline  # noqa: E402
from test_gpu_parity import GRAPH_SECTIONS, GROUP_SECTIONS, PATH_SECTIONS  # noqa: E402


def main():
    keys = [k for k in sys.argv[1:] if k in bench.WORKLOADS] or ["ecoli", "dmel"]
    common.build_oracle()
    os.environ.setdefault("HERA_ORACLE_THREADS", str(os.cpu_count() or 1))
    out_path = bench.DIGEST_GOLDEN
    golden = json.loads(out_path.read_text()) if out_path.exists() else {}
    opts = ["--rb-att", str(bench.RB_ATT)]
    names = GRAPH_SECTIONS + ["rank_0", "rank_1"] + PATH_SECTIONS + GROUP_SECTIONS
    for key in keys:
        ds, meta = bench.make_dataset(key)
        t0 = time.perf_counter()
        dump = ds / "oracle.bin"
        ora = common.run_dump(common.ORACLE_BIN, ds, dump, opts, extra=["--rng", "philox", "--seed", str(bench.WALK_SEED)])
        dump.unlink()
        t_ora = time.perf_counter() - t0
        t0 = time.perf_counter()
        with capi.Hera(0) as h:
            gpu = pipeline.run_hot_path(ds, opts, seed=bench.WALK_SEED, hera=h, batched=True)
            digest = h.digest()
        t_gpu = time.perf_counter() - t0
        common.assert_sections_equal(ora, gpu, names, key)
        golden[key] = {
            "digests": {k: str(v) for k, v in digest.items()},
            "verified": f"tools/make_bench_digests.py: {len(names)} tables of a 1-GPU run bit-exact against oracle/hera_oracle "
                        f"(--rng philox --seed {bench.WALK_SEED} --rb-att {bench.RB_ATT}) on {bench.WORKLOADS[key]['name']}",
            "records": int(gpu["load_ctr"].sum()), "kept_overlaps": int(gpu["e_q"].size), "nodes": int(gpu["node_len"].size),
            "paths": int(gpu["p_len"].size), "unique_paths": int(gpu["u_len"].size), "anchor_pairs": int(gpu["pair_a1"].size),
            "groups": int(gpu["grp_pair"].size), "seconds": {"oracle": round(t_ora, 1), "cuda_path_incl_file_io_and_export": round(t_gpu, 1)},
        }
        print(key, {k: golden[key][k] for k in ("records", "kept_overlaps", "nodes", "paths", "unique_paths", "anchor_pairs", "groups", "seconds")}, flush=True)
        out_path.write_text(json.dumps(golden, indent=1) + "\n")
        del ora, gpu


if __name__ == "__main__":
    main()
