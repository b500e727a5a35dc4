"""Host logic of the sharded load with world_size 2 and 3 over gloo (CPU): the byte ranges handed out by
hera_shard_range / shard.read_slice, together with the exchange protocol the CUDA tokenizer follows
(token counts all-gathered, record k = global tokens [16k, 16k+16), a record belongs to the rank that owns
its first token and is finished from the look-ahead), partition the file's records exactly."""
import os
import random
import socket
import sys
from pathlib import Path

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

SPACES = b" \t\n\r\x0b\x0c"


def tokens_with_offsets(buf):
    out, i, n = [], 0, len(buf)
    while i < n:
        while i < n and buf[i] in SPACES:
            i += 1
        if i >= n:
            break
        j = i
        while j < n and buf[j] not in SPACES:
            j += 1
        out.append((i, bytes(buf[i:j])))
        i = j
    return out


def whole_file_records(text):
    toks = [t for _, t in tokens_with_offsets(text)]
    recs = [tuple(toks[k:k + 12]) for k in range(0, len(toks) - 11, 16)]
    return recs


def worker(rank, world, port, path, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from telomeri_b200 import shard
    sl = shard.read_slice(path, rank, world)
    buf = sl.data.tobytes()
    toks = tokens_with_offsets(buf)
    prev_nonws = sl.prev_byte >= 0 and bytes([sl.prev_byte]) not in [bytes([c]) for c in SPACES]
    if prev_nonws and toks and toks[0][0] == 0:
        toks = toks[1:]           # the first "token" continues one that started in the previous slice
    owned = [(o, t) for o, t in toks if o < sl.owned]
    counts = [None] * world
    dist.all_gather_object(counts, len(owned))
    base = sum(counts[:rank])
    total = sum(counts)
    recs = []
    for k, (o, _) in enumerate(owned):
        g = base + k
        if g % 16 == 0 and g + 12 <= total:
            i = toks.index(owned[k])
            assert i + 12 <= len(toks), "look-ahead too short"
            recs.append((g // 16, tuple(t for _, t in toks[i:i + 12])))
    allrecs = [None] * world
    dist.all_gather_object(allrecs, recs)
    if rank == 0:
        q.put((counts, [r for part in allrecs for r in part]))
    dist.barrier()
    dist.destroy_process_group()


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def run_world(world, path):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=worker, args=(r, world, port, str(path), q)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return out


def make_text(seed, nrec, ragged):
    rnd = random.Random(seed)
    lines = []
    for i in range(nrec):
        f = [f"read{rnd.randrange(10**6):07d}", "15000", str(rnd.randrange(5000)), "14990", rnd.choice("+-"), f"read{rnd.randrange(10**6):07d}", "16000", "0",
             str(rnd.randrange(9000, 15000)), "9000", "14000", "0", "tp:A:S", "cm:i:1", "s1:i:9", "dv:f:0.1"]
        sep = (lambda: rnd.choice([" ", "\t", "  ", "\r\n", "\n"])) if ragged else (lambda: "\t")
        lines.append("".join(t + sep() for t in f[:-1]) + f[-1] + "\n")
    text = "".join(lines)
    if ragged:
        text = "  \n" + text + "tail1 15000 3 4 + tail2 16000 0 1 2 3 0"   # 12-token last record, no newline
    return text.encode()


def test_slices_partition_the_records(tmp_path):
    for seed, nrec, ragged in ((1, 400, False), (2, 333, True), (3, 3, True), (4, 0, False)):
        text = make_text(seed, nrec, ragged)
        path = tmp_path / f"t{seed}.paf"
        path.write_bytes(text)
        want = whole_file_records(text)
        for world in (2, 3):
            counts, recs = run_world(world, path)
            assert sum(counts) == len(tokens_with_offsets(text))
            assert [k for k, _ in recs] == list(range(len(want))), (seed, world)
            assert [r for _, r in recs] == want, (seed, world)


def test_shard_range_properties():
    from telomeri_b200 import capi
    for n in (0, 1, 15, 16, 17, 1000, 10**6 + 3, 5 * 10**9 + 11):
        for world in (1, 2, 3, 8):
            cuts = [capi.shard_range(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            for r in range(world):
                b, e, a = cuts[r]
                assert b <= e <= a <= n and b % 16 == 0
                if r + 1 < world:
                    assert cuts[r + 1][0] == e
                assert a == min(n, e + (1 << 20))
