"""In-tree build of the native pieces: libhera_b200.so (CUDA, sm_100a) and the host tools.

    python -m telomeri_b200.build            # build everything that is out of date
"""
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
LIB = PKG / "libhera_b200.so"
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
              "-Xcompiler", "-Wall", "--fmad=false", "-Xcudafe", "--diag_suppress=20011,--diag_suppress=20014"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found: the CUDA library cannot be built (there is no CPU fallback)")


def _newer(target, sources):
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(s).stat().st_mtime > t for s in sources)


def build_library(force=False, verbose=False):
    sources = sorted(CSRC.glob("*.cu"))
    deps = sources + sorted(CSRC.glob("*.cuh")) + [ROOT / "include" / "hera_b200.h"]
    objdir = PKG / "build"
    objdir.mkdir(exist_ok=True)
    nvcc = _nvcc()
    objs = []
    procs = []
    for src in sources:
        obj = objdir / (src.stem + ".o")
        objs.append(obj)
        if force or _newer(obj, deps):
            cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", str(src), "-o", str(obj)]
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            print(f"--- {src.name}\n{out}", file=sys.stderr)
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    if force or procs or not LIB.exists():
        subprocess.run([nvcc, "-shared", "-o", str(LIB)] + [str(o) for o in objs] + ["-lcudart"], check=True)
    return LIB


def build_tools():
    """Dataset generator, oracle binary and (when /root/reference exists) oracle/_ref."""
    tools_bin = ROOT / "tools" / "bin"
    tools_bin.mkdir(parents=True, exist_ok=True)
    gen = tools_bin / "gen_dataset"
    if _newer(gen, [ROOT / "tools" / "gen_dataset.cpp"]):
        subprocess.run(["g++", "-O2", "-std=c++17", "-o", str(gen), str(ROOT / "tools" / "gen_dataset.cpp")], check=True)
    subprocess.run(["make", "-s", "-C", str(ROOT / "oracle"), "bin/hera_oracle", "ref"], check=True, stdout=sys.stderr)


def build_host():
    """Drop-in command-line binary: reference class surface (telomeri_b200/host) over the C ABI.
    bin/Telomeri uses this repo's Main.cpp; bin/Telomeri_refmain is the REFERENCE's own, unmodified
    src/Main.cpp compiled against our headers (only where /root/reference exists; output is
    git-ignored and travels to the GPU box like the other built files)."""
    host = PKG / "host"
    out_dir = PKG / "bin"
    out_dir.mkdir(exist_ok=True)
    srcs = sorted((host / "src").glob("*.cpp"))
    deps = srcs + sorted((host / "include").glob("*.hpp")) + [ROOT / "include" / "hera_b200.h", LIB]
    common = ["g++", "-O2", "-std=c++17", f"-I{host / 'include'}", f"-I{ROOT / 'include'}"]
    link = [f"-L{PKG}", "-lhera_b200", "-Wl,-rpath,$ORIGIN/.."]
    exe = out_dir / "Telomeri"
    if _newer(exe, deps):
        subprocess.run(common + ["-Wall", "-o", str(exe)] + [str(s) for s in srcs] + link, check=True)
    ref_main = Path("/root/reference/src/Main.cpp")
    if ref_main.exists():
        exe2 = out_dir / "Telomeri_refmain"
        if _newer(exe2, deps + [ref_main]):
            subprocess.run(common + ["-w", "-o", str(exe2), str(ref_main)] + [str(s) for s in srcs if s.name != "Main.cpp"] + link, check=True)
    return exe


if __name__ == "__main__":
    build_library(force="--force" in sys.argv, verbose="-v" in sys.argv)
    build_tools()
    build_host()
    print(LIB)
