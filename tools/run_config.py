#!/usr/bin/env python
"""One BASELINE.json configuration end to end on the GPU box: generate the dataset, run the CPU oracle and the
CUDA path (optionally as a local gang of N ranks) on it, compare every table bit for bit, print timings.
  python tools/run_config.py dmel [--ranks N] [--rb-att K]
  python tools/run_config.py chr1 --rb-att 1000
Used for the configurations that are too large for the test-suite (configs[3], configs[4])."""
import argparse
import json
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

CONFIGS = {
    "yeast": ["--genome", "12000000", "--read-len", "20000", "--coverage", "60", "--contigs", "120", "--gap", "30000", "--families", "56", "--copies", "4", "--min-ovl", "1000", "--seed", "3"],
    # configs[4]: human-chr1 scale, high repeat (272 families x 30 copies of 5-20 kb: ~40 % of the records are
    # repeat-induced), 30x 25 kb reads -> 402 M read-read records (42 GB of PAF); run with --rb-att 1000
    "chr1": ["--genome", "248000000", "--read-len", "25000", "--coverage", "30", "--contigs", "120", "--gap", "30000", "--families", "272", "--copies", "30", "--min-ovl", "1000", "--seed", "5"],
    "dmel": ["--genome", "140000000", "--read-len", "20000", "--coverage", "60", "--contigs", "400", "--gap", "30000", "--families", "650", "--copies", "4", "--min-ovl", "1000", "--seed", "4"],
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config")
    ap.add_argument("--ranks", type=int, default=1)
    ap.add_argument("--rb-att", type=int, default=10)
    ap.add_argument("--out", default="/tmp/hera_cfg")
    ap.add_argument("--no-oracle", action="store_true", help="time the CUDA path only (stage times), no comparison")
    ap.add_argument("--no-exports", action="store_true", help="with --no-oracle: do not copy the tables back to the host")
    args = ap.parse_args()
    import common
    from telomeri_b200 import capi, pipeline
    from test_gpu_parity import GRAPH_SECTIONS, GROUP_SECTIONS, PATH_SECTIONS
    d = Path(args.out) / args.config
    d.mkdir(parents=True, exist_ok=True)

    def note(msg):
        print(f"[run_config +{time.perf_counter() - t_start:.0f}s] {msg}", file=sys.stderr, flush=True)
    t_start = time.perf_counter()
    common.build_generator()
    common.build_oracle()
    t0 = time.perf_counter()
    meta = subprocess.run([str(common.GEN_BIN), "--out", str(d), "--no-fasta"] + CONFIGS[args.config], check=True, capture_output=True, text=True).stdout
    t_gen = time.perf_counter() - t0
    note(f"generated {meta.strip()}")
    opts = ["--rb-att", str(args.rb_att)]
    ora, t_ora = None, 0.0
    if not args.no_oracle:
        t0 = time.perf_counter()
        ora = common.run_dump(common.ORACLE_BIN, d, d / "o.bin", opts, extra=["--rng", "philox", "--seed", "17"])
        t_ora = time.perf_counter() - t0
        (d / "o.bin").unlink()
        note(f"oracle done in {t_ora:.0f} s")
    t0 = time.perf_counter()
    if args.ranks == 1:
        gpu = pipeline.run_hot_path(d, opts, seed=17, batched=True, exports=not (args.no_oracle and args.no_exports))
    else:
        import torch
        k = max(1, torch.cuda.device_count())
        with capi.LocalGang(args.ranks, [r % k for r in range(args.ranks)]) as gang:
            outs = gang.each(lambda r, h: pipeline.run_hot_path(d, opts, seed=17, hera=h, batched=True) if r == 0 else
                             pipeline.run_hot_path(d, opts, seed=17, hera=h, batched=True, exports=False) and None)
        gpu = outs[0]
    t_gpu = time.perf_counter() - t0
    note(f"CUDA path done in {t_gpu:.0f} s; comparing")
    names = GRAPH_SECTIONS + ["rank_0", "rank_1"] + PATH_SECTIONS + GROUP_SECTIONS
    st = gpu["walk_stats"]
    if args.no_oracle:
        tm = gpu["timings_ms"]
        rec = int(gpu["load_ctr"].sum())
        graph_ms = sum(tm["parse"]) + tm["intern"] + tm["csr"] + tm["ranks"]
        print(json.dumps({"config": args.config, "ranks": args.ranks, "generator": json.loads(meta), "records": rec, "kept_overlaps": int(gpu["load_cum_edges"][-1]),
                          "walks": int(sum(s.walks for s in st)), "accepted_paths": int(sum(s.found for s in st)), "stage_ms": tm,
                          "graph_build_ms": round(graph_ms, 2), "overlaps_per_s_graph_build": round(rec / graph_ms * 1e3),
                          "walks_per_s": round(sum(s.walks for s in st) / tm["walk_stage"] * 1e3),
                          "parse_GBps": [round(b / ms / 1e6, 1) if ms else None for b, ms in zip(tm["paf_bytes"], tm["parse"])],
                          "h2d_GBps_incl_file_read": [round(b / ms / 1e6, 1) if ms else None for b, ms in zip(tm["paf_bytes"], tm["h2d"])],
                          "seconds": {"generate": round(t_gen, 1), "cuda_path_incl_file_io": round(t_gpu, 1)}}))
        return
    common.assert_sections_equal(ora, gpu, names, args.config)
    print(json.dumps({"config": args.config, "ranks": args.ranks, "generator": json.loads(meta), "records": int(gpu["load_ctr"].sum()),
                      "kept_overlaps": int(gpu["e_q"].size), "nodes": int(gpu["node_len"].size), "anchors": int(gpu["node_anchor"].sum()),
                      "walks": int(sum(s.walks for s in st)), "paths": int(gpu["p_len"].size), "unique_paths": int(gpu["u_len"].size),
                      "anchor_pairs": int(gpu["pair_a1"].size), "sections_compared": len(names), "bit_exact": True,
                      "seconds": {"generate": round(t_gen, 1), "oracle_cpu": round(t_ora, 1), "cuda_path_incl_file_io_and_export": round(t_gpu, 1)}}))


if __name__ == "__main__":
    main()
