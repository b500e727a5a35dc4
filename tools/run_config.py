#!/usr/bin/env python
"""One BASELINE.json configuration end to end on the GPU box.  Used for the configurations that are too large for the
test-suite (configs[3], configs[4]).

  single GPU, against the oracle (every table bit for bit), recording the digests of the verified run:
      python tools/run_config.py chr1 --rb-att 1000 --write-digests
  several ranks of one process on the available GPUs (local gang), against the oracle:
      python tools/run_config.py dmel --ranks 4
  one process per GPU over NCCL (the north_star target run), checked against the recorded digests:
      python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \\
          tools/run_config.py chr1 --rb-att 1000 --nccl
The digests (hera_digest, include/hera_b200.h) live in tests/golden/config_digests.json."""
import argparse
import json
import os
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
DIGESTS = ROOT / "tests" / "golden" / "config_digests.json"
SEED = 17


def generate(common, d, config):
    d.mkdir(parents=True, exist_ok=True)
    if (d / "done").exists():
        return json.loads((d / "done").read_text()), 0.0
    common.build_generator()
    t0 = time.perf_counter()
    meta = subprocess.run([str(common.GEN_BIN), "--out", str(d), "--no-fasta"] + CONFIGS[config], check=True, capture_output=True, text=True).stdout
    (d / "done").write_text(meta)
    return json.loads(meta), time.perf_counter() - t0


def stage_times(gpu):
    tm = gpu["timings_ms"]
    st = gpu["walk_stats"]
    rec = int(gpu["load_ctr"].sum())
    graph_ms = sum(tm["parse"]) + tm["intern"] + tm["csr"] + tm["ranks"]
    return {"stage_ms": {k: v for k, v in tm.items()}, "graph_build_ms_sum_of_stages": round(graph_ms, 2),
            "overlaps_per_s_graph_build": round(rec / graph_ms * 1e3), "walks": int(sum(s.walks for s in st)),
            "accepted_paths": int(sum(s.found for s in st)), "walks_per_s": round(sum(s.walks for s in st) / tm["walk_stage"] * 1e3),
            "parse_GBps": [round(b / ms / 1e6, 1) if ms else None for b, ms in zip(tm["paf_bytes"], tm["parse"])]}


def run_nccl(args, common, capi, pipeline):
    import torch
    import torch.distributed as dist
    dist.init_process_group("gloo")   # hands out the NCCL id and the verdicts; the data path is the library's own communicator
    rank, world = dist.get_rank(), dist.get_world_size()
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    d = Path(args.out) / args.config
    gen_s = 0.0
    if rank == 0:
        meta, gen_s = generate(common, d, args.config)
    dist.barrier()
    meta = json.loads((d / "done").read_text())
    ids = [capi.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, 0)
    opts = ["--rb-att", str(args.rb_att)]
    h = capi.Hera(local)
    h.comm_init_nccl(rank, world, ids[0])
    t0 = time.perf_counter()
    gpu = pipeline.run_hot_path(d, opts, seed=SEED, hera=h, batched=True, exports=False)
    t_gpu = time.perf_counter() - t0
    digest = {k: str(v) for k, v in h.digest().items()}
    h.close()
    want = json.loads(DIGESTS.read_text()).get(f"{args.config}_rb{args.rb_att}", {}) if DIGESTS.exists() else {}
    ok = bool(want) and want["digests"] == digest
    verdicts = [None] * world
    dist.all_gather_object(verdicts, {"rank": rank, "ok": ok, "digest": digest, "seconds": round(t_gpu, 1), "times": stage_times(gpu)})
    if rank == 0:
        same = all(v["digest"] == verdicts[0]["digest"] for v in verdicts)
        print(json.dumps({"config": args.config, "transport": "nccl", "ranks": world, "generator": meta, "records": int(gpu["load_ctr"].sum()),
                          "kept_overlaps": int(gpu["load_cum_edges"][-1]), "rb_att": args.rb_att,
                          "ranks_agree": same, "bit_exact_vs_oracle_verified_digests": all(v["ok"] for v in verdicts), "verified_by": want.get("verified"),
                          "rank0": verdicts[0]["times"], "stage_ms_max_over_ranks": {k: max(v["times"]["stage_ms"][k] for v in verdicts) for k in ("intern", "csr", "ranks", "walk_stage", "groups")},
                          "parse_ms_max_over_ranks": [max(v["times"]["stage_ms"]["parse"][i] for v in verdicts) for i in range(2)],
                          "seconds": {"generate": round(gen_s, 1), "cuda_path_incl_file_io_max_over_ranks": max(v["seconds"] for v in verdicts)}}))
        if not (same and all(v["ok"] for v in verdicts)):
            bad = {k: (verdicts[0]["digest"][k], want.get("digests", {}).get(k)) for k in verdicts[0]["digest"] if verdicts[0]["digest"][k] != want.get("digests", {}).get(k)}
            print("MISMATCH", json.dumps(bad), file=sys.stderr)
    dist.barrier()
    dist.destroy_process_group()
    if not ok:
        sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config")
    ap.add_argument("--ranks", type=int, default=1)
    ap.add_argument("--rb-att", type=int, default=10)
    ap.add_argument("--out", default="/tmp/hera_cfg")
    ap.add_argument("--nccl", action="store_true", help="one process per GPU (launch with torch.distributed.run); checks the recorded digests")
    ap.add_argument("--write-digests", action="store_true", help="after the comparison with the oracle, record the digests of this run")
    ap.add_argument("--no-oracle", action="store_true", help="time the CUDA path only (stage times), no comparison")
    args = ap.parse_args()
    import common
    from telomeri_b200 import capi, pipeline
    from test_gpu_parity import GRAPH_SECTIONS, GROUP_SECTIONS, PATH_SECTIONS
    if args.nccl:
        return run_nccl(args, common, capi, pipeline)
    d = Path(args.out) / args.config

    def note(msg):
        print(f"[run_config +{time.perf_counter() - t_start:.0f}s] {msg}", file=sys.stderr, flush=True)
    t_start = time.perf_counter()
    common.build_oracle()
    os.environ.setdefault("HERA_ORACLE_THREADS", str(os.cpu_count() or 1))
    meta, t_gen = generate(common, d, args.config)
    note(f"generated {meta}")
    opts = ["--rb-att", str(args.rb_att)]
    ora, t_ora = None, 0.0
    if not args.no_oracle:
        t0 = time.perf_counter()
        ora = common.run_dump(common.ORACLE_BIN, d, d / "o.bin", opts, extra=["--rng", "philox", "--seed", str(SEED)])
        t_ora = time.perf_counter() - t0
        (d / "o.bin").unlink()
        note(f"oracle done in {t_ora:.0f} s")
    t0 = time.perf_counter()
    digest = None
    if args.ranks == 1:
        with capi.Hera(0) as h:
            gpu = pipeline.run_hot_path(d, opts, seed=SEED, hera=h, batched=True, exports=not args.no_oracle)
            digest = {k: str(v) for k, v in h.digest().items()}
    else:
        import torch
        k = max(1, torch.cuda.device_count())
        with capi.LocalGang(args.ranks, [r % k for r in range(args.ranks)]) as gang:
            outs = gang.each(lambda r, h: pipeline.run_hot_path(d, opts, seed=SEED, hera=h, batched=True, exports=(r == 0 and not args.no_oracle)))
        gpu = outs[0]
    t_gpu = time.perf_counter() - t0
    note(f"CUDA path done in {t_gpu:.0f} s")
    out = {"config": args.config, "ranks": args.ranks, "generator": meta, "records": int(gpu["load_ctr"].sum()), "kept_overlaps": int(gpu["load_cum_edges"][-1]),
           "rb_att": args.rb_att}
    out.update(stage_times(gpu))
    names = GRAPH_SECTIONS + ["rank_0", "rank_1"] + PATH_SECTIONS + GROUP_SECTIONS
    if ora is not None:
        common.assert_sections_equal(ora, gpu, names, args.config)
        out.update({"nodes": int(gpu["node_len"].size), "anchors": int(gpu["node_anchor"].sum()), "paths": int(gpu["p_len"].size),
                    "unique_paths": int(gpu["u_len"].size), "anchor_pairs": int(gpu["pair_a1"].size), "sections_compared": len(names), "bit_exact": True})
        if args.write_digests and digest:
            golden = json.loads(DIGESTS.read_text()) if DIGESTS.exists() else {}
            golden[f"{args.config}_rb{args.rb_att}"] = {
                "digests": digest, "generator": "tools/gen_dataset --no-fasta " + " ".join(CONFIGS[args.config]), "options": opts, "seed": SEED,
                "verified": f"tools/run_config.py: {len(names)} tables of a 1-GPU run bit-exact against oracle/hera_oracle (--rng philox --seed {SEED})",
                "records": out["records"], "kept_overlaps": out["kept_overlaps"], "nodes": out["nodes"], "paths": out["paths"], "anchor_pairs": out["anchor_pairs"]}
            DIGESTS.write_text(json.dumps(golden, indent=1) + "\n")
    out["seconds"] = {"generate": round(t_gen, 1), "oracle_cpu": round(t_ora, 1), "oracle_threads": int(os.environ["HERA_ORACLE_THREADS"]),
                      "cuda_path_incl_file_io_and_export": round(t_gpu, 1)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
