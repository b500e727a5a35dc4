"""The built library is sm_100a code and uses the Blackwell paths DESIGN.md claims (no GPU needed: cuobjdump of the
in-tree .so).  Guards the claims of profiles/r03_sass_counts.txt against a silent rebuild with other flags."""
import re
import shutil
import subprocess

import pytest

import common


@pytest.fixture(scope="module")
def sass():
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    from telomeri_b200 import build
    lib = build.LIB
    assert lib.exists(), "library not built (python -m telomeri_b200.build)"
    return subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True, check=True).stdout


def kernels(sass):
    out, cur = {}, None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            out[cur] = []
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            out[cur].append(m.group(1))
    return out


def test_every_kernel_is_sm100a_code(sass):
    archs = set(re.findall(r"arch = (\S+)", sass))
    assert "sm_100a" in archs
    # (an sm_52 entry is the empty stub of the `nvcc -shared` link step: it must hold no function)
    for block in sass.split("Fatbin elf code:")[1:]:
        if "arch = sm_100a" not in block:
            assert "Function :" not in block, "a kernel was compiled for another architecture"


def test_parse_kernel_stages_tiles_with_bulk_copies_and_mbarriers(sass):
    k = kernels(sass)
    parse = [ops for name, ops in k.items() if "paf_parse_kernel" in name]
    assert len(parse) == 1
    ops = [o.split(".")[0] for o in parse[0]]
    assert ops.count("UBLKCP") >= 1, "cp.async.bulk staging is gone"
    assert ops.count("SYNCS") >= 2, "mbarrier arrive / wait is gone"
    assert ops.count("LDGSTS") <= 4, "per-thread cp.async is back on the main path"


def test_default_walk_kernels_do_not_spill(sass):
    from telomeri_b200 import build
    res = subprocess.run(["cuobjdump", "--dump-resource-usage", str(build.LIB)], capture_output=True, text=True, check=True).stdout
    cur = None
    seen = 0
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
        elif cur and "REG:" in line and ("walk_mc_kernelILb0ELi4ELi8E" in cur or "walk_mc_kernelILb1ELi4ELi8E" in cur or "walk_det_kernel" in cur):   # the default variants
            regs = int(re.search(r"REG:(\d+)", line).group(1))
            stack = int(re.search(r"STACK:(\d+)", line).group(1))
            assert stack == 0 and regs <= 64, (cur, line)   # 64 registers: 32 resident warps per SM (hera_set_option walk_warps_per_sm)
            seen += 1
            cur = None
    assert seen >= 3
