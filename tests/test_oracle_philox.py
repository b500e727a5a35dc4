"""Known-answer vectors for Philox4x32-10 (Random123 kat_vectors) through oracle/philox.hpp."""
import subprocess

import common

SRC = r'''
#include <cstdio>
#include "philox.hpp"
int main() {
    using hera_oracle::philox4x32_10;
    auto p = [](hera_oracle::Philox4 r) { std::printf("%08x %08x %08x %08x\n", r.v[0], r.v[1], r.v[2], r.v[3]); };
    p(philox4x32_10(0, 0, 0, 0, 0, 0));
    p(philox4x32_10(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu));
    p(philox4x32_10(0x243f6a88u, 0x85a308d3u, 0x13198a2eu, 0x03707344u, 0xa4093822u, 0x299f31d0u));
    std::printf("%.17g\n", hera_oracle::mc_draw(42, 1, 7, 3, 0));
    return 0;
}
'''

KAT = ["6627e8d5 e169c58d bc57ac4c 9b00dbd8", "408f276d 41c83b0e a20bc7c6 6d5451fd", "d16cfe09 94fdcceb 5001e420 24126ea1"]


def test_philox_known_answers(tmp_path):
    src = tmp_path / "kat.cpp"
    src.write_text(SRC)
    exe = tmp_path / "kat"
    subprocess.run(["g++", "-O2", "-std=c++17", f"-I{common.ORACLE_DIR}", "-o", str(exe), str(src)], check=True)
    lines = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split("\n")
    assert lines[:3] == KAT
    assert 0.0 <= float(lines[3]) < 1.0
