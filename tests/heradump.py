"""Reader for the heradump binary format (oracle/hera_dump.hpp): name -> numpy array."""
import struct

import numpy as np

_DTYPES = {0: np.uint8, 1: np.uint32, 2: np.int32, 3: np.uint64, 4: np.int64, 5: np.float32, 6: np.float64}


def read_dump(path):
    out = {}
    with open(path, "rb") as f:
        data = f.read()
    pos = 0
    while pos < len(data):
        (nl,) = struct.unpack_from("<I", data, pos)
        pos += 4
        name = data[pos:pos + nl].decode()
        pos += nl
        dt, count = struct.unpack_from("<IQ", data, pos)
        pos += 12
        dtype = np.dtype(_DTYPES[dt])
        nbytes = count * dtype.itemsize
        out[name] = np.frombuffer(data, dtype=dtype, count=count, offset=pos).copy()
        pos += nbytes
    return out


def text(arr):
    return arr.tobytes().decode()
