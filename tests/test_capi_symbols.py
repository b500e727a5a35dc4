"""CPU-side checks of the boundary: the library builds, loads, and exports every symbol that
include/hera_b200.h declares (no compute calls -- there is no GPU here)."""
import ctypes
import re

import pytest

import common


@pytest.fixture(scope="module")
def lib_path():
    from telomeri_b200 import build
    return build.build_library()


def declared_symbols():
    text = (common.ROOT / "include" / "hera_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hera_[a-z_]+)\s*\(", text)))


def test_header_symbols_are_exported(lib_path):
    lib = ctypes.CDLL(str(lib_path))
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/hera_b200.h but not exported"


def test_binding_covers_header(lib_path):
    from telomeri_b200 import capi
    assert sorted(capi.SIGNATURES) == declared_symbols()
    capi.load_library()


def test_struct_layouts_match_header():
    """ctypes mirrors must have the C sizes (x86-64 SysV)."""
    from telomeri_b200 import capi
    assert ctypes.sizeof(capi.FilterParams) == 40
    assert ctypes.sizeof(capi.PathParams) == 32
    assert ctypes.sizeof(capi.LoadResult) == 48   # 4 x u64 + parse_ms, h2d_ms, exchange_ms + padding
    assert ctypes.sizeof(capi.GraphInfo) == 48
    assert ctypes.sizeof(capi.WalkStats) == 32
    assert ctypes.sizeof(capi.GraphView) == 24 + 22 * 8
    assert ctypes.sizeof(capi.PathsView) == 12 * 8
    assert ctypes.sizeof(capi.GroupsView) == 4 * 8 + 11 * 8 + 8


def test_create_without_gpu_fails_loudly(lib_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from telomeri_b200 import capi
    with pytest.raises(capi.HeraError):
        capi.Hera(0)


def test_bench_digest_golden_covers_the_bench_workloads():
    """bench.py proves parity of its own run (any number of GPUs) by comparing hera_digest with these oracle-verified words."""
    import json
    import bench
    from telomeri_b200 import capi
    golden = json.loads(bench.DIGEST_GOLDEN.read_text())
    for key in bench.WORKLOADS:
        assert sorted(golden[key]["digests"]) == sorted(capi.DIGEST_NAMES), key
        assert "bit-exact against oracle" in golden[key]["verified"] and int(golden[key]["digests"]["u_len"]) != 0
