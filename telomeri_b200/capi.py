"""ctypes binding of libhera_b200.so (include/hera_b200.h).  No compute happens in Python: every
method is one C-ABI call; results come back as numpy views copied out of the library's pinned
mirrors.  Importing this module without the built library raises -- there is no CPU fallback."""
import ctypes as C
from pathlib import Path

import numpy as np

PKG = Path(__file__).resolve().parent
import os as _os

# HERA_B200_LIB: a differently built copy of the same library (kernel tuning experiments)
LIB_PATH = Path(_os.environ.get("HERA_B200_LIB") or PKG / "libhera_b200.so")

MODE = {"min": 0, "max": 1, "avg": 2, "sum": 3}
METRIC = {"es": 0, "os": 1, "es_sqrt": 2, "os_sqrt": 3}

u8p, u32p, u64p, i64p, i32p, f32p = (C.POINTER(t) for t in (C.c_uint8, C.c_uint32, C.c_uint64, C.c_int64, C.c_int32, C.c_float))


class FilterParams(C.Structure):
    _fields_ = [("mode", C.c_int32), ("min_overlap_length", C.c_uint64), ("min_overlap_percentage", C.c_float),
                ("max_overhang_length", C.c_uint64), ("max_overhang_percentage", C.c_float)]


class PathParams(C.Structure):
    _fields_ = [("rebuild_attempts", C.c_int32), ("backtrack_attempts", C.c_int32), ("len_threshold", C.c_int64),
                ("window_size", C.c_uint64), ("ratio_threshold", C.c_float)]


class LoadResult(C.Structure):
    _fields_ = [("records", C.c_uint64), ("kept", C.c_uint64), ("kept_total", C.c_uint64), ("bytes", C.c_uint64),
                ("parse_ms", C.c_float), ("h2d_ms", C.c_float), ("exchange_ms", C.c_float)]


class GraphInfo(C.Structure):
    _fields_ = [("nodes", C.c_uint64), ("anchors", C.c_uint64), ("edges", C.c_uint64), ("slots", C.c_uint64),
                ("intern_ms", C.c_float), ("csr_ms", C.c_float), ("rank_ms", C.c_float)]


class GraphView(C.Structure):
    _fields_ = [("nodes", C.c_uint64), ("edges", C.c_uint64), ("slots", C.c_uint64),
                ("node_len", u32p), ("node_anchor", u8p), ("node_name_off", u64p), ("node_names", C.POINTER(C.c_char)),
                ("e_q", u32p), ("e_t", u32p), ("e_qs", u32p), ("e_qe", u32p), ("e_ts", u32p), ("e_te", u32p),
                ("e_os", f32p), ("e_si", f32p), ("e_es", f32p), ("e_strand", u8p),
                ("adj_off", u64p), ("adj_rec", u32p), ("adj_side", u8p),
                ("rank_es", u32p), ("rank_os", u32p), ("rank_es_sqrt", u32p), ("rank_os_sqrt", u32p), ("canon", u32p)]


class PathsView(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("total_nodes", C.c_uint64), ("off", u64p), ("nodes", u32p), ("slots", u32p), ("len", i64p),
                ("calls", C.c_uint64), ("call_kind", u32p), ("call_metric", u32p), ("call_found", u64p),
                ("mc_walks", C.c_uint64), ("mc_draws_per_walk", u32p)]


class GroupsView(C.Structure):
    _fields_ = [("pairs", C.c_uint64), ("groups", C.c_uint64), ("sorted_total", C.c_uint64), ("pig_total", C.c_uint64),
                ("pair_a1", u32p), ("pair_a2", u32p), ("pair_off", u64p), ("pair_sorted", u32p),
                ("grp_pair", u32p), ("grp_size0", u32p), ("grp_pig_off", u64p), ("grp_pig", u32p),
                ("grp_consensus", i64p), ("grp_valid", i32p), ("pair_consensus", i64p), ("ms", C.c_float)]


class ScaffoldView(C.Structure):
    _fields_ = [("nodes", C.c_uint64), ("node", u32p), ("slot", u32p), ("logged", C.c_uint64), ("logged_index", u32p),
                ("fasta_bytes", C.c_uint64), ("sequence_length", C.c_uint64), ("fasta", C.POINTER(C.c_char)), ("ms", C.c_float)]


class KeptColumns(C.Structure):
    _fields_ = [(k, u32p) for k in ("qs", "qe", "ts", "te", "qlen", "tlen", "rec")] + [(k, f32p) for k in ("os", "si", "es")] + [("flags", u8p)] + \
               [(k, u64p) for k in ("qhash", "thash", "qname", "tname")]


NOALLOC_MAX_TEXT_BYTES = 16384 * 30720 - 30720


class WalkStats(C.Structure):
    _fields_ = [("walks", C.c_uint64), ("found", C.c_uint64), ("steps", C.c_uint64), ("ms", C.c_float), ("launches", C.c_uint32)]


# name -> (restype, argtypes); must list every symbol of include/hera_b200.h
SIGNATURES = {
    "hera_create": (C.c_void_p, [C.c_int]),
    "hera_destroy": (None, [C.c_void_p]),
    "hera_last_error": (C.c_char_p, [C.c_void_p]),
    "hera_kernel_launches": (C.c_uint64, [C.c_void_p]),
    "hera_timer_start": (C.c_int, [C.c_void_p]),
    "hera_timer_stop": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "hera_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "hera_probe_list_gather": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_double), C.POINTER(C.c_float)]),
    "hera_digest": (C.c_int, [C.c_void_p, u64p]),
    "hera_paf_parse_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_uint64, C.c_int, C.POINTER(FilterParams), C.c_uint32, C.POINTER(KeptColumns), C.c_uint64,
                                         C.c_void_p, C.c_void_p, C.c_char_p, C.c_uint64]),
    "hera_graph_set_filter": (C.c_int, [C.c_void_p, C.POINTER(FilterParams)]),
    "hera_graph_load_file": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int, C.c_uint32, C.POINTER(LoadResult)]),
    "hera_graph_load_bytes": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_uint32, C.POINTER(LoadResult)]),
    "hera_graph_load_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_uint32, C.POINTER(LoadResult)]),
    "hera_graph_load_device_view": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_uint32, C.POINTER(LoadResult)]),
    "hera_nccl_unique_id": (C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64]),
    "hera_comm_init_nccl": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "hera_comm_local_create": (C.c_void_p, [C.c_int]),
    "hera_comm_local_destroy": (None, [C.c_void_p]),
    "hera_comm_attach_local": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "hera_comm_rank": (C.c_int, [C.c_void_p]),
    "hera_comm_barrier": (C.c_int, [C.c_void_p]),
    "hera_comm_size": (C.c_int, [C.c_void_p]),
    "hera_shard_range": (None, [C.c_uint64, C.c_int, C.c_int, u64p, u64p, u64p]),
    "hera_graph_load_shard": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_uint32, C.POINTER(LoadResult)]),
    "hera_graph_reset": (C.c_int, [C.c_void_p]),
    "hera_graph_finalize": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(GraphInfo)]),
    "hera_graph_export": (C.c_int, [C.c_void_p, C.POINTER(GraphView)]),
    "hera_graph_gather_rows": (C.c_int, [C.c_void_p]),
    "hera_graph_save": (C.c_int, [C.c_void_p, C.c_char_p]),
    "hera_graph_load_cache": (C.c_int, [C.c_void_p, C.c_char_p]),
    "hera_paths_monte_carlo": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(PathParams), C.c_uint64, C.POINTER(WalkStats)]),
    "hera_paths_deterministic": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(WalkStats)]),
    "hera_paths_build": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(PathParams), C.c_uint64, C.POINTER(WalkStats)]),
    "hera_paths_unique": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "hera_paths_reset": (C.c_int, [C.c_void_p]),
    "hera_paths_export": (C.c_int, [C.c_void_p, C.POINTER(PathsView)]),
    "hera_groups": (C.c_int, [C.c_void_p, C.POINTER(PathParams), C.POINTER(GroupsView)]),
    "hera_scaffold": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.c_uint64, C.POINTER(ScaffoldView)]),
}

DIGEST_NAMES = ["node_len", "node_anchor", "e_q", "e_t", "adj_off", "adj_pay", "walk_view", "canon", "rank_0", "rank_1", "rank_2", "rank_3",
                "u_off", "u_nodes", "u_slots", "u_len", "pair_a1", "pair_a2", "pair_sorted", "grp_size0", "grp_pig", "grp_consensus", "grp_valid",
                "pair_consensus"]
DIGEST_WORDS = len(DIGEST_NAMES)


def digest_words(a):
    """numpy restatement of the digest of include/hera_b200.h (hera_digest) for one array."""
    a = np.ascontiguousarray(a)
    w = a.astype(np.uint64) if a.dtype.itemsize == 1 else a.view(np.uint32).astype(np.uint64)
    i = np.arange(1, w.size + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = (i * np.uint64(0x9E3779B97F4A7C15)) ^ ((w + np.uint64(1)) * np.uint64(0xC2B2AE3D27D4EB4F))
        x ^= x >> np.uint64(32)
        x *= np.uint64(0xD6E8FEB86659FD93)
        x ^= x >> np.uint64(32)
        return int(x.sum(dtype=np.uint64))


_lib = None


def load_library():
    """dlopen the in-tree CUDA library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -m telomeri_b200.build` (nvcc, sm_100a). There is no CPU fallback.")
        lib = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class HeraError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"hera error {code}: {msg}")
        self.code = code


def _arr(ptr, n, dtype, copy=True):
    """copy=False returns a view into the library's pinned mirror (valid until the next call that mutates that object)."""
    if n == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(int(n),)).astype(dtype, copy=copy)


def default_filter(mode="avg", min_oll=0, min_olp=0.0, max_ohl=2 ** 64 - 1, max_ohp=1.0):
    return FilterParams(MODE[mode], min_oll, min_olp, max_ohl, max_ohp)


def default_path_params(rb_att=2000, bt_att=60, len_thr=4700000, w_size=1000, r_thr=0.9):
    return PathParams(rb_att, bt_att, len_thr, w_size, r_thr)


class Hera:
    """Thin object wrapper: one hera_ctx."""

    def __init__(self, device=0):
        self.lib = load_library()
        self.h = self.lib.hera_create(device)
        if not self.h:
            raise HeraError(-1, self.lib.hera_last_error(None).decode())

    def close(self):
        if self.h:
            self.lib.hera_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc):
        if rc != 0:
            raise HeraError(rc, self.lib.hera_last_error(self.h).decode())

    @property
    def launches(self):
        return int(self.lib.hera_kernel_launches(self.h))

    def timer_start(self):
        self._check(self.lib.hera_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float()
        self._check(self.lib.hera_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def comm_barrier(self):
        self._check(self.lib.hera_comm_barrier(self.h))

    def probe_list_gather(self, warps_per_sm=64, hops=256):
        """Measured list-walking ceiling on the finalized graph: (GB/s of adjacency entries, ms)."""
        gbs, ms = C.c_double(), C.c_float()
        self._check(self.lib.hera_probe_list_gather(self.h, int(warps_per_sm), int(hops), C.byref(gbs), C.byref(ms)))
        return gbs.value, ms.value

    def set_option(self, key, value):
        self._check(self.lib.hera_set_option(self.h, key.encode(), int(value)))

    def digest(self):
        """hera_digest: {table name: 64-bit digest of the device-resident table}."""
        out = (C.c_uint64 * DIGEST_WORDS)()
        self._check(self.lib.hera_digest(self.h, out))
        return dict(zip(DIGEST_NAMES, (int(v) for v in out)))

    def set_filter(self, fp):
        self._check(self.lib.hera_graph_set_filter(self.h, C.byref(fp)))

    def load_file(self, path, anchors, test_load_num=0):
        r = LoadResult()
        self._check(self.lib.hera_graph_load_file(self.h, str(path).encode(), int(anchors), test_load_num, C.byref(r)))
        return r

    def load_bytes(self, data, anchors, test_load_num=0):
        r = LoadResult()
        buf = (C.c_char * len(data)).from_buffer_copy(data) if len(data) else None
        self._check(self.lib.hera_graph_load_bytes(self.h, buf, len(data), int(anchors), test_load_num, C.byref(r)))
        return r

    def load_host_ptr(self, ptr, n, anchors, test_load_num=0):
        r = LoadResult()
        self._check(self.lib.hera_graph_load_bytes(self.h, C.c_void_p(ptr), n, int(anchors), test_load_num, C.byref(r)))
        return r

    def load_device_ptr(self, ptr, n, anchors, test_load_num=0):
        r = LoadResult()
        self._check(self.lib.hera_graph_load_device(self.h, C.c_void_p(ptr), n, int(anchors), test_load_num, C.byref(r)))
        return r

    def load_device_view_ptr(self, ptr, n, anchors, test_load_num=0):
        """hera_graph_load_device_view: parsed in place, the caller keeps the buffer alive until graph_reset / close."""
        r = LoadResult()
        self._check(self.lib.hera_graph_load_device_view(self.h, C.c_void_p(ptr), n, int(anchors), test_load_num, C.byref(r)))
        return r

    # ---- sharded runs (one rank per GPU) -------------------------------------------------------
    @property
    def rank(self):
        return int(self.lib.hera_comm_rank(self.h))

    @property
    def size(self):
        return int(self.lib.hera_comm_size(self.h))

    def comm_init_nccl(self, rank, nranks, unique_id):
        """unique_id: the 128 bytes rank 0 got from nccl_unique_id(), handed to every rank by the launcher."""
        buf = (C.c_char * 128).from_buffer_copy(bytes(unique_id))
        self._check(self.lib.hera_comm_init_nccl(self.h, rank, nranks, buf))

    def comm_attach_local(self, gang, rank):
        self._check(self.lib.hera_comm_attach_local(self.h, gang, rank))

    def load_shard_ptr(self, ptr, on_device, owned, avail, prev_byte, at_eof, anchors, test_load_num=0):
        """Collective.  ptr = this rank's slice (see shard_range); returns global counts."""
        r = LoadResult()
        self._check(self.lib.hera_graph_load_shard(self.h, C.c_void_p(ptr), int(on_device), owned, avail, int(prev_byte), int(at_eof), int(anchors),
                                                   test_load_num, C.byref(r)))
        return r

    def graph_reset(self):
        self._check(self.lib.hera_graph_reset(self.h))

    def finalize(self, metrics=("es", "os")):
        mask = 0
        for m in metrics:
            mask |= 1 << (METRIC[m] if isinstance(m, str) else int(m))
        info = GraphInfo()
        self._check(self.lib.hera_graph_finalize(self.h, mask, C.byref(info)))
        return info

    def gather_rows(self):
        """Sharded contexts: the collective part of export_graph / save_graph / scaffold / digest, for ranks that skip those calls."""
        self._check(self.lib.hera_graph_gather_rows(self.h))

    def save_graph(self, path):
        self._check(self.lib.hera_graph_save(self.h, str(path).encode()))

    def load_graph_cache(self, path):
        self._check(self.lib.hera_graph_load_cache(self.h, str(path).encode()))

    def export_graph(self):
        v = GraphView()
        self._check(self.lib.hera_graph_export(self.h, C.byref(v)))
        N, E, S = v.nodes, v.edges, v.slots
        out = {
            "node_len": _arr(v.node_len, N, np.uint32), "node_anchor": _arr(v.node_anchor, N, np.uint8),
            "node_name_off": _arr(v.node_name_off, N + 1, np.uint64),
            "adj_off": _arr(v.adj_off, N + 1, np.uint64), "adj_rec": _arr(v.adj_rec, S, np.uint32), "adj_side": _arr(v.adj_side, S, np.uint8),
            "canon": _arr(v.canon, S, np.uint32),
        }
        nbytes = int(out["node_name_off"][-1]) if N else 0
        out["node_names"] = np.frombuffer(C.string_at(v.node_names, nbytes), dtype=np.uint8).copy() if nbytes else np.zeros(0, np.uint8)
        for k in ("e_q", "e_t", "e_qs", "e_qe", "e_ts", "e_te"):
            out[k] = _arr(getattr(v, k), E, np.uint32)
        for k in ("e_os", "e_si", "e_es"):
            out[k] = _arr(getattr(v, k), E, np.float32)
        out["e_strand"] = _arr(v.e_strand, E, np.uint8)
        for i, k in enumerate(("rank_es", "rank_os", "rank_es_sqrt", "rank_os_sqrt")):
            if getattr(v, k):
                out[f"rank_{i}"] = _arr(getattr(v, k), S, np.uint32)
        return out

    def monte_carlo(self, metric, pp, seed=0):
        s = WalkStats()
        self._check(self.lib.hera_paths_monte_carlo(self.h, METRIC[metric] if isinstance(metric, str) else metric, C.byref(pp), seed, C.byref(s)))
        return s

    def deterministic(self, metric):
        s = WalkStats()
        self._check(self.lib.hera_paths_deterministic(self.h, METRIC[metric] if isinstance(metric, str) else metric, C.byref(s)))
        return s

    def build_paths(self, calls, pp, seed=0):
        """calls: [("mc" | "det", metric), ...] -- all walk kernels in flight together, paths appended in call order."""
        n = len(calls)
        kinds = (C.c_int * n)(*[0 if k == "mc" else 1 for k, _ in calls])
        metrics = (C.c_int * n)(*[METRIC[m] if isinstance(m, str) else int(m) for _, m in calls])
        stats = (WalkStats * n)()
        self._check(self.lib.hera_paths_build(self.h, n, kinds, metrics, C.byref(pp), seed, stats))
        return list(stats)

    def unique(self):
        n = C.c_uint64()
        self._check(self.lib.hera_paths_unique(self.h, C.byref(n)))
        return n.value

    def paths_reset(self):
        self._check(self.lib.hera_paths_reset(self.h))

    def export_paths(self, prefix="", copy=True):
        v = PathsView()
        self._check(self.lib.hera_paths_export(self.h, C.byref(v)))
        P = v.paths
        return {
            prefix + "off": _arr(v.off, P + 1, np.uint64, copy), prefix + "nodes": _arr(v.nodes, v.total_nodes, np.uint32, copy),
            prefix + "slots": _arr(v.slots, v.total_nodes - P, np.uint32, copy), prefix + "len": _arr(v.len, P, np.int64, copy),
            "call_kind": _arr(v.call_kind, v.calls, np.uint32), "call_metric": _arr(v.call_metric, v.calls, np.uint32),
            "call_found": _arr(v.call_found, v.calls, np.uint64), "mc_draws_per_walk": _arr(v.mc_draws_per_walk, v.mc_walks, np.uint32),
        }

    def scaffold(self, reads_fasta, contigs_fasta, min_path_num=10):
        """Chain of consensus paths + Scaffolder; returns the scaffold path and the bytes of HERA_result.fasta."""
        v = ScaffoldView()
        self._check(self.lib.hera_scaffold(self.h, str(reads_fasta).encode(), str(contigs_fasta).encode(), min_path_num, C.byref(v)))
        return {"scaffold_nodes": _arr(v.node, v.nodes, np.uint32), "scaffold_slots": _arr(v.slot, v.nodes - 1, np.uint32),
                "logged": _arr(v.logged_index, v.logged, np.uint32), "fasta": C.string_at(v.fasta, v.fasta_bytes),
                "sequence_length": int(v.sequence_length), "ms": float(v.ms)}

    def groups(self, pp, copy=True):
        """copy=False: views into the library's host tables (valid until the next groups call)."""
        v = GroupsView()
        self._check(self.lib.hera_groups(self.h, C.byref(pp), C.byref(v)))
        K, G = v.pairs, v.groups
        return {
            "pair_a1": _arr(v.pair_a1, K, np.uint32, copy), "pair_a2": _arr(v.pair_a2, K, np.uint32, copy), "pair_off": _arr(v.pair_off, K + 1, np.uint64, copy),
            "pair_sorted": _arr(v.pair_sorted, v.sorted_total, np.uint32, copy), "grp_pair": _arr(v.grp_pair, G, np.uint32, copy),
            "grp_size0": _arr(v.grp_size0, G, np.uint32, copy), "grp_pig_off": _arr(v.grp_pig_off, G + 1, np.uint64, copy),
            "grp_pig": _arr(v.grp_pig, v.pig_total, np.uint32, copy), "grp_consensus": _arr(v.grp_consensus, G, np.int64, copy),
            "grp_valid": _arr(v.grp_valid, G, np.int32, copy), "pair_consensus": _arr(v.pair_consensus, K, np.int64, copy), "ms": float(v.ms),
        }


def paf_parse_noalloc(device, text_ptr, n, anchors, fp, columns, capacity, result_ptr, stream=0, test_load_num=0):
    """hera_paf_parse_noalloc: columns = {name: device pointer} for the 15 KeptColumns fields, result_ptr = 4 u64 in device
    memory.  Queues the work on `stream` (a cudaStream_t handle, 0 = default stream); raises HeraError on rejection."""
    lib = load_library()
    cols = KeptColumns()
    for name, ctype in KeptColumns._fields_:
        setattr(cols, name, C.cast(C.c_void_p(int(columns[name])), ctype))
    err = C.create_string_buffer(512)
    rc = lib.hera_paf_parse_noalloc(int(device), C.c_void_p(int(text_ptr)), int(n), int(anchors), C.byref(fp), int(test_load_num), C.byref(cols), int(capacity),
                                    C.c_void_p(int(result_ptr)), C.c_void_p(int(stream)), err, 512)
    if rc != 0:
        raise HeraError(rc, err.value.decode())


def nccl_unique_id():
    """128-byte NCCL id (rank 0 calls this, every rank passes it to Hera.comm_init_nccl)."""
    lib = load_library()
    out = (C.c_char * 128)()
    err = C.create_string_buffer(512)
    rc = lib.hera_nccl_unique_id(out, err, 512)
    if rc != 0:
        raise HeraError(rc, err.value.decode())
    return bytes(out)


def shard_range(file_bytes, rank, nranks):
    """(begin, end, avail_end) of rank's slice of a file_bytes long PAF file (hera_shard_range)."""
    lib = load_library()
    b, e, a = C.c_uint64(), C.c_uint64(), C.c_uint64()
    lib.hera_shard_range(file_bytes, rank, nranks, C.byref(b), C.byref(e), C.byref(a))
    return b.value, e.value, a.value


class LocalGang:
    """Several contexts of ONE process that run the sharded path together (local transport): one
    host thread per rank; every method call is made on all ranks at once.  Used for single-process
    multi-GPU runs and to exercise the sharded path on a single GPU."""

    def __init__(self, nranks, devices=None):
        import concurrent.futures
        self.lib = load_library()
        self.n = nranks
        self.gang = self.lib.hera_comm_local_create(nranks)
        devices = devices or [0] * nranks
        self.ranks = [Hera(d) for d in devices]
        for r, h in enumerate(self.ranks):
            h.comm_attach_local(self.gang, r)
        self.pool = concurrent.futures.ThreadPoolExecutor(max_workers=nranks)

    def each(self, fn):
        """fn(rank, hera) on every rank concurrently (ctypes calls release the GIL); returns the list of results."""
        futs = [self.pool.submit(fn, r, h) for r, h in enumerate(self.ranks)]
        return [f.result() for f in futs]

    def close(self):
        for h in self.ranks:
            h.close()
        self.ranks = []
        if self.gang:
            self.lib.hera_comm_local_destroy(self.gang)
            self.gang = None
        self.pool.shutdown()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
