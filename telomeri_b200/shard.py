"""Host side of the sharded load: which bytes of a PAF file a rank reads and what it
tells hera_graph_load_shard about them.  Pure plumbing -- the tokenizer itself is CUDA."""
from collections import namedtuple
from pathlib import Path

import numpy as np

from . import capi

FileSlice = namedtuple("FileSlice", "data owned avail prev_byte at_eof begin file_bytes")


def read_slice(path, rank, world):
    """This rank's bytes of `path`: it owns the tokens that start in [begin, begin + owned) and also
    holds the look-ahead up to begin + avail so that a record starting in its range can be finished;
    prev_byte is the file byte before `begin` (-1 at the start of the file)."""
    path = Path(path)
    n = path.stat().st_size
    b, e, a = capi.shard_range(n, rank, world)
    prev = -1
    with open(path, "rb") as f:
        if b:
            f.seek(b - 1)
            prev = f.read(1)[0]
        f.seek(b)
        data = np.frombuffer(f.read(a - b), dtype=np.uint8)
    return FileSlice(data, e - b, a - b, prev, a == n, b, n)


def load_slice(hera, sl, anchors, test_load_num=0, device_ptr=None):
    """Collective hera_graph_load_shard call for a FileSlice (host bytes, or device_ptr if the slice is resident)."""
    ptr = device_ptr if device_ptr is not None else (sl.data.ctypes.data if sl.avail else 0)
    return hera.load_shard_ptr(ptr, device_ptr is not None, sl.owned, sl.avail, sl.prev_byte, sl.at_eof, anchors, test_load_num)
