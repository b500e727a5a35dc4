#!/usr/bin/env python
"""Benchmark of the HERA repeat-resolution hot path (BASELINE.json metric: overlaps/s graph build +
paths/s extension, % HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one pass of the hot path over one synthetic dataset of BASELINE.json configs[1]
(E. coli scale): PAF text -> filtered overlaps -> node ids -> CSR + rank permutations ->
Monte-Carlo and greedy walks (2 metrics each) -> unique -> anchor-pair groups and consensus.
`value` is graph-build overlaps/s with the PAF text already resident in HBM; `paths_per_s` the walk
rate; `e2e` the same step through the C ABI from pinned host buffers with the result read back.
N > 1 (torchrun, one rank per GPU): the SHARDED path on ONE dataset N times the size (genome, contigs
and repeat families scaled by N -> per-GPU work fixed, weak scaling): every rank tokenizes its byte
range of the PAF files, kept overlaps are all-gathered over NCCL, every rank holds the full CSR, walks
are dealt out by walk id and the accepted paths all-gathered; time = max over ranks.
"""
import argparse
import json
import os
import re
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOAD = {
    "name": "configs[1]: synthetic E. coli-scale 4.6 Mbp, 10 repeat families, 50x 15 kb reads",
    "gen": ["--genome", "4600000", "--read-len", "15000", "--coverage", "50", "--contigs", "10", "--gap", "40000",
            "--families", "10", "--copies", "3", "--min-ovl", "2000", "--no-fasta"],
    "seed": 2,
    "rb_att": 2000,          # reference default (src/Main.cpp:64)
    "metrics": ["es", "os"],  # the two calls main() enables (src/Main.cpp:237-243)
}
REF_SAMPLE = {"rr_records": 400000, "rb_att": 100}   # bounded sample timed on the host cores
DATA_ROOT = Path(os.environ.get("HERA_BENCH_DATA", "/tmp/hera_bench"))
QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


def ensure_tools():
    from telomeri_b200 import build as b
    b.build_tools()
    return ROOT / "tools" / "bin" / "gen_dataset"


def gen_args(scale):
    """configs[1] generator options with genome, contigs and repeat families multiplied by `scale`."""
    a = list(WORKLOAD["gen"])
    for key, base in (("--genome", 4600000), ("--contigs", 10), ("--families", 10)):
        a[a.index(key) + 1] = str(base * scale)
    return a


def make_dataset(seed, scale=1):
    d = DATA_ROOT / (f"ecoli_seed{seed}" if scale == 1 else f"ecoli_x{scale}_seed{seed}")
    if not (d / "done").exists():
        d.mkdir(parents=True, exist_ok=True)
        gen = ensure_tools()
        out = subprocess.run([str(gen), "--out", str(d), "--seed", str(seed)] + gen_args(scale), check=True, capture_output=True, text=True).stdout
        (d / "done").write_text(out)
    return d, json.loads((d / "done").read_text())


def make_ref_sample(ds):
    """First REF_SAMPLE['rr_records'] lines of rr_overlaps.paf + the whole cr file."""
    s = ds / "ref_sample"
    if not (s / "done").exists():
        s.mkdir(exist_ok=True)
        n = 0
        with open(ds / "rr_overlaps.paf", "rb") as fi, open(s / "rr_overlaps.paf", "wb") as fo:
            for line in fi:
                fo.write(line)
                n += 1
                if n >= REF_SAMPLE["rr_records"]:
                    break
        (s / "cr_overlaps.paf").write_bytes((ds / "cr_overlaps.paf").read_bytes())
        for f in ("reads.fasta", "contigs.fasta"):
            (s / f).write_text(">none\nA\n")
        cr = sum(1 for _ in open(s / "cr_overlaps.paf", "rb"))
        (s / "done").write_text(json.dumps({"rr_records": n, "cr_records": cr}))
    return s, json.loads((s / "done").read_text())


def bind_to_gpu_cpus(index):
    """Best effort: run this process (and first-touch its pinned buffers) on the CPUs NVML reports as local to the GPU,
    so that the e2e leg's host<->device copies do not cross a socket.  No-op where NVML or the affinity call is missing."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


class ClockSampler:
    """nvidia-smi clocks line of the profiling recipe, sampled while the timed region runs."""

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={QUERY}", "--format=csv,noheader,nounits", "-lms", "200", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 9 for i in range(4) if r[5 + i].lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def run_reference_sample(ds, steps, warmup):
    """Times the reference's own CPU implementation (oracle/_ref/Telomeri when it was built from
    /root/reference, else the oracle port) on the bounded sample.  Returns per-step dicts."""
    sample, meta = make_ref_sample(ds)
    records = meta["rr_records"] + meta["cr_records"]
    ref_bin = ROOT / "oracle" / "_ref" / "Telomeri"
    kind = "reference" if ref_bin.exists() else "port"
    out = []
    for i in range(warmup + steps):
        if kind == "reference":
            cmd = [str(ref_bin), "--rb-att", str(REF_SAMPLE["rb_att"]), str(sample / "rr_overlaps.paf"), str(sample / "cr_overlaps.paf"),
                   str(sample / "reads.fasta"), str(sample / "contigs.fasta"), str(sample / "out.fasta")]
            t0 = time.perf_counter()
            p = subprocess.run(cmd, capture_output=True, text=True)
            wall = time.perf_counter() - t0
            laps = [float(x) for x in re.findall(r"^Done \(([0-9.e+-]+)s\)", p.stdout, flags=re.M)]
            found = [int(x) for x in re.findall(r"^Found (\d+) paths\.", p.stdout, flags=re.M)]
            load_s, paths_s = (laps + [wall, wall])[:2]
        else:
            ensure_tools()
            cmd = [str(ROOT / "oracle" / "bin" / "hera_oracle"), "--rng", "mt19937", "--rb-att", str(REF_SAMPLE["rb_att"]), "--stage", "paths", "--timing",
                   str(sample / "rr_overlaps.paf"), str(sample / "cr_overlaps.paf"), str(sample / "o.bin")]
            p = subprocess.run(cmd, capture_output=True, text=True, check=True)
            t = json.loads(p.stdout.strip().splitlines()[-1])
            load_s, paths_s, found = t["graph_s"], t["paths_s"], []
        if i >= warmup:
            out.append({"load_s": load_s, "paths_s": paths_s, "found": sum(found)})
    load = statistics.mean(s["load_s"] for s in out)
    paths = statistics.mean(s["paths_s"] for s in out)
    return {"kind": kind, "records": records, "load_s": load, "paths_s": paths, "overlaps_per_s": records / load,
            "sample": f"first {meta['rr_records']} rr records + all {meta['cr_records']} cr records of the workload, --rb-att {REF_SAMPLE['rb_att']}, "
                      f"reference laps: load {load:.2f}s, walks {paths:.2f}s"}


def hot_path_step(h, capi, pp, rr, cr, device_resident, read_back):
    """One step through the C ABI.  rr / cr: (pointer, nbytes) of the whole file, or -- sharded run -- a Slice of it.
    Returns stage times (ms, CUDA events on the library's stream), counts and the kernel-launch delta."""
    launches0 = h.launches
    h.graph_reset()
    h.paths_reset()

    def load(src, anchors):
        if isinstance(src, Slice):
            return h.load_shard_ptr(src.ptr(device_resident), 2 if device_resident else 0, src.owned, src.avail, src.prev, src.at_eof, anchors)
        # resident text is parsed in place (hera_graph_load_device_view): the buffers live for the whole run
        return (h.load_device_view_ptr if device_resident else h.load_host_ptr)(src[0], src[1], anchors)
    h.comm_barrier()   # sharded runs: a late host thread of another rank is not part of this rank's stage times
    h.timer_start()
    r1 = load(rr, False)
    r2 = load(cr, True)
    info = h.finalize(WORKLOAD["metrics"])
    t_graph = h.timer_stop()
    h.comm_barrier()
    h.timer_start()
    # the four build calls of main() as one hera_paths_build call: their kernels share the GPU
    stats = h.build_paths([("mc", m) for m in WORKLOAD["metrics"]] + [("det", m) for m in WORKLOAD["metrics"]], pp, 20261104)
    walks = sum(s.walks for s in stats)
    steps = sum(s.steps for s in stats)
    found = sum(s.found for s in stats)
    n_unique = h.unique()
    t_paths = h.timer_stop()
    h.timer_start()
    g = h.groups(pp)
    t_groups = h.timer_stop()
    d2h = 0
    if read_back:  # the result a caller consumes: groups/consensus tables + the unique paths
        paths = h.export_paths("u_", copy=False)   # views into the pinned mirrors the library copied the result to
        d2h = sum(v.nbytes for v in g.values() if hasattr(v, "nbytes")) + sum(v.nbytes for v in paths.values())
    return {"graph_ms": t_graph, "paths_ms": t_paths, "groups_ms": t_groups, "parse_ms": r1.parse_ms + r2.parse_ms,
            "exchange_ms": r1.exchange_ms + r2.exchange_ms, "parse_rr_ms": r1.parse_ms, "intern_ms": info.intern_ms, "csr_ms": info.csr_ms, "rank_ms": info.rank_ms,
            "records": r1.records + r2.records, "kept": r2.kept_total, "nodes": info.nodes, "walks": walks, "walk_entries": steps, "found": found,
            "unique_paths": n_unique, "pairs": len(g["pair_a1"]), "launches": h.launches - launches0, "d2h": d2h,
            "rr_bytes": r1.bytes, "rr_kept": r1.kept}


def parse_traffic():
    """DRAM bytes of one paf_parse_kernel launch on the rr text from the committed `ncu --set full` capture (per launch, like `achieved`)."""
    try:
        t = json.loads((ROOT / "profiles" / "parse_kernel_traffic.json").read_text())
        return int(t["dram_bytes_read"]) + int(t["dram_bytes_write"])
    except (OSError, KeyError, ValueError):
        return None


class Slice:
    """This rank's byte range of a PAF file (hera_shard_range): pinned host copy + device copy."""

    def __init__(self, path, rank, world, dev, capi, torch, np):
        n = path.stat().st_size
        b, e, a = capi.shard_range(n, rank, world)
        self.owned, self.avail, self.at_eof, self.file_bytes = e - b, a - b, a == n, n
        self.prev = -1
        if b:
            with open(path, "rb") as f:
                f.seek(b - 1)
                self.prev = f.read(1)[0]
        self.host = torch.empty(max(self.avail, 1), dtype=torch.uint8, pin_memory=True)
        if self.avail:
            self.host.numpy()[:self.avail] = np.fromfile(path, dtype=np.uint8, count=self.avail, offset=b)
        self.dev = self.host.to(dev)

    def ptr(self, device_resident):
        return self.dev.data_ptr() if device_resident else self.host.data_ptr()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    config = {"workload": WORKLOAD["name"], "generator": "tools/gen_dataset " + " ".join(WORKLOAD["gen"]) + f" --seed {WORKLOAD['seed']}",
              "filter": "reference defaults (AVG)", "rb_att": WORKLOAD["rb_att"], "metrics": WORKLOAD["metrics"],
              "parallelism": (f"{world} ranks, one dataset {world}x the size: PAF sharded by byte range, kept overlaps pulled from peer memory over NVLink "
                              f"(NCCL rendezvous), CSR replicated, walks dealt out by walk id, paths all-gathered (NCCL); ranks are aligned by a barrier before the "
                              f"graph stage and before the path stage of every step (stage times exclude the wait for a late host thread; e2e wall time does not)")
                             if world > 1 else "single GPU",
              "l2": "PAF text (173 MB per GPU) is larger than L2 (126 MB); the roofline leg additionally writes 512 MB before the parse kernel"}
    if world > 1:
        config["generator"] = "tools/gen_dataset " + " ".join(gen_args(world)) + f" --seed {WORKLOAD['seed']}"

    if args.impl == "reference":
        if rank != 0:
            return
        ds, meta = make_dataset(WORKLOAD["seed"])
        ref = run_reference_sample(ds, args.steps, max(0, args.warmup))
        v = ref["overlaps_per_s"]
        print(json.dumps({
            "impl": "reference", "metric": "overlaps/s graph build", "value": v, "unit": "overlaps/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * (ref["load_s"] + ref["paths_s"]), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config,
            "paths_per_s": None if not ref["paths_s"] else (len(WORKLOAD["metrics"]) * 10 * REF_SAMPLE["rb_att"]) / ref["paths_s"],
            "cpu_baseline": {"value": v, "unit": "overlaps/s", "cores": 1, "kind": ref["kind"], "sample": ref["sample"]},
            "e2e": {"value": v, "unit": "overlaps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import numpy as np
    import torch
    from telomeri_b200 import build as b, capi
    if rank == 0:
        b.build_library()
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(local_rank)
    bind_to_gpu_cpus(local_rank)
    if rank == 0:
        make_dataset(WORKLOAD["seed"], world)
    if world > 1:
        dist.barrier()
    ds, meta = make_dataset(WORKLOAD["seed"], world)
    h = capi.Hera(local_rank)
    if world > 1:
        # the library's own NCCL communicator (collectives run on its stream); torch.distributed only hands out the id
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(capi.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        h.comm_init_nccl(rank, world, bytes(idt.cpu().numpy()))
        rr_s = Slice(ds / "rr_overlaps.paf", rank, world, dev, capi, torch, np)
        cr_s = Slice(ds / "cr_overlaps.paf", rank, world, dev, capi, torch, np)
        rr_n, cr_n = rr_s.avail, cr_s.avail
        rr_d, cr_d = rr_s.dev, cr_s.dev
    else:
        def pinned(path):
            n = path.stat().st_size
            t = torch.empty(max(n, 1), dtype=torch.uint8, pin_memory=True)
            t.numpy()[:n] = np.fromfile(path, dtype=np.uint8)
            return t, n
        rr_h, rr_n = pinned(ds / "rr_overlaps.paf")
        cr_h, cr_n = pinned(ds / "cr_overlaps.paf")
        rr_d, cr_d = rr_h.to(dev), cr_h.to(dev)
    torch.cuda.synchronize(dev)
    pp = capi.default_path_params(rb_att=WORKLOAD["rb_att"])
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def sources(device_resident):
        if world > 1:
            return rr_s, cr_s
        return ((rr_d.data_ptr(), rr_n), (cr_d.data_ptr(), cr_n)) if device_resident else ((rr_h.data_ptr(), rr_n), (cr_h.data_ptr(), cr_n))

    def timed(device_resident, read_back, steps):
        per = []
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            if device_resident:
                flush.zero_()            # L2 flush between timed iterations (outside the event-timed stages)
                torch.cuda.synchronize(dev)
            src = sources(device_resident)
            per.append(hot_path_step(h, capi, pp, src[0], src[1], device_resident, read_back))
        barrier()
        return per, time.perf_counter() - t0

    for _ in range(max(3, args.warmup)):
        hot_path_step(h, capi, pp, *sources(True), True, False)
    with ClockSampler(local_rank) as clocks:
        dev_steps, _ = timed(True, False, args.steps)
        e2e_steps, e2e_wall = timed(False, True, args.steps)
    # measured ceiling for the walks: dependent list gathers on the finalized graph of the last step, same resident warps
    h.probe_list_gather(64, 64)
    gather_gbs, gather_ms = h.probe_list_gather(64, 256)
    # roofline leg: the dominant streaming kernel (paf_parse_kernel on the rr text) timed alone by its own CUDA events, L2 flushed
    h.set_option("flush_l2", 1)
    parse_ms = []
    for _ in range(args.steps + 1):
        h.graph_reset()
        leg = h.load_device_view_ptr(rr_d.data_ptr(), rr_n, False)   # not collective: this rank's rr text as a stand-alone file
        parse_ms.append(leg.parse_ms)
    h.set_option("flush_l2", 0)
    parse_ms = statistics.mean(parse_ms[1:])

    def agg(steps, key):
        return sum(s[key] for s in steps)
    sums = torch.tensor([agg(dev_steps, "graph_ms"), agg(dev_steps, "paths_ms"), agg(dev_steps, "groups_ms"), e2e_wall * 1e3,
                         agg(e2e_steps, "graph_ms")], dtype=torch.float64, device=dev)
    # records / walks / paths reported by the collective calls are already global; launches and copied bytes are per rank
    counts = torch.tensor([agg(dev_steps, "launches"), rr_n + cr_n, e2e_steps[-1]["d2h"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.MAX)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    graph_ms, paths_ms, groups_ms, e2e_ms, e2e_graph_ms = sums.tolist()
    launches, h2d_bytes, d2h_bytes = counts.tolist()
    records, walks, e2e_records, uniq = agg(dev_steps, "records"), agg(dev_steps, "walks"), agg(e2e_steps, "records"), agg(dev_steps, "unique_paths")
    if rank == 0:
        s0 = dev_steps[-1]
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except OSError:
            pass
        peak = peaks.get("hbm_gbs", 6650.0)
        algo_bytes = leg.bytes + 73 * leg.kept   # every PAF byte read once + 73 B written per kept overlap
        achieved = algo_bytes / (parse_ms * 1e-3) / 1e9
        line = {
            "metric": "overlaps/s graph build", "value": records / (graph_ms * 1e-3), "unit": "overlaps/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": (graph_ms + paths_ms + groups_ms) / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
            "paths_per_s": walks / (paths_ms * 1e-3), "accepted_paths_per_s": agg(dev_steps, "found") / (paths_ms * 1e-3),
            "unique_paths_per_step": uniq / args.steps,
            "stages_ms_per_step": {k: statistics.mean(s[k] for s in dev_steps) for k in
                                   ("graph_ms", "parse_ms", "exchange_ms", "intern_ms", "csr_ms", "rank_ms", "paths_ms", "groups_ms")},
            "per_step": {"records": s0["records"], "kept_overlaps": s0["kept"], "nodes": s0["nodes"], "walks": s0["walks"],
                         "adjacency_entries_examined": s0["walk_entries"], "anchor_pairs": s0["pairs"]},
            "gpu_launches": int(launches),
            "e2e": {"value": e2e_records / (e2e_graph_ms * 1e-3), "unit": "overlaps/s", "h2d_bytes_per_step": int(h2d_bytes),
                    "d2h_bytes_per_step": int(d2h_bytes), "whole_step_ms": e2e_ms / args.steps,
                    "note": "graph build from pinned host buffers (H2D inside the timed region); whole_step_ms adds walks, groups and the D2H read of the result"},
            "roofline": {"bound": "hbm", "kernel": "paf_parse_kernel (rr text)", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s", "algorithmic_bytes": algo_bytes,
                         "kernel_ms": parse_ms, "traffic": parse_traffic()},
            # north_star: the pointer-chasing walks are reported against the measured random-gather ceiling
            "roofline_walks": {"bound": "dependent list gather (L2/HBM latency)", "kernel": "walk_mc_kernel x2 + walk_det_kernel x2 in flight together",
                               "achieved": s0["walk_entries"] * 16 / (statistics.mean(s["paths_ms"] for s in dev_steps) * 1e-3) / 1e9 / world,
                               "peak": gather_gbs, "unit": "GB/s per GPU", "frac": None,
                               "peak_source": f"hera_probe_list_gather(64 warps/SM, 256 hops) on this graph: {gather_ms:.3f} ms",
                               "algorithmic_bytes": int(s0["walk_entries"]) * 16},
            "clocks": clocks.summary(),
        }
        # % of the HBM roofline for every streaming stage, from its algorithmic bytes (E kept overlaps,
        # 2E slots, N nodes, P paths) and the stage's device time
        E_, N_, M_ = s0["kept"], s0["nodes"], len(WORKLOAD["metrics"])
        st = line["stages_ms_per_step"]
        stage_bytes = {"K3 interning (intern_ms)": (2 * E_ * 10 + 2 * E_ * 4, st["intern_ms"]),
                       "K4 CSR + K4b ranks (csr_ms + rank_ms)": (2 * E_ * 80 + (N_ + 1) * 4 + 2 * E_ * 4 * (M_ + 1), st["csr_ms"] + st["rank_ms"])}
        line["roofline_stages"] = [{"stage": k, "algorithmic_bytes": int(b), "ms": ms, "achieved": b / (ms * 1e-3) / 1e9, "peak": peak,
                                    "unit": "GB/s per GPU (replicated stages: every GPU handles all kept overlaps)", "frac": b / (ms * 1e-3) / 1e9 / peak}
                                   for k, (b, ms) in stage_bytes.items()]
        rw = line["roofline_walks"]
        rw["frac"] = rw["achieved"] / rw["peak"] if rw["peak"] else None
        if world == 1 and not args.no_cpu_baseline:
            ref = run_reference_sample(ds, 1, 0)
            line["cpu_baseline"] = {"value": ref["overlaps_per_s"], "unit": "overlaps/s", "cores": 1, "kind": ref["kind"], "sample": ref["sample"],
                                    "paths_per_s": (len(WORKLOAD["metrics"]) * 10 * REF_SAMPLE["rb_att"]) / ref["paths_s"] if ref["paths_s"] else None}
        print(json.dumps(line))
    h.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
