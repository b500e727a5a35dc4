"""telomeri_b200/csrc/introsort.cuh is __host__ __device__ code: compile it for the host and check
that it moves elements exactly like libstdc++'s std::sort (the GPU tests then check the device side
through group parity)."""
import subprocess

import common

SRC = r'''
#define __host__
#define __device__
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <vector>
#include "introsort.cuh"
struct Rng { uint64_t s; uint64_t next() { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); } };
int main() {
    Rng rng{777};
    for (int round = 0; round < 3000; round++) {
        size_t n = round < 100 ? round : rng.next() % 4000;
        uint64_t distinct = 1 + rng.next() % (round % 5 == 0 ? 4 : n + 1);
        std::vector<long> key(n);
        for (size_t i = 0; i < n; i++) key[i] = round % 3 == 0 ? long(rng.next() % distinct) : round % 3 == 1 ? long((n - i) % distinct) : long(i % 2 ? i / 2 : n - i / 2);
        if (round % 7 == 3 && n > 64) { size_t k = n / 2; for (size_t i = 0; i < k; i++) key[i] = (i & 1) ? long(k + i) : long(i); for (size_t i = k; i < n; i++) key[i] = long(2 * (i - k)); }
        std::vector<uint32_t> a(n), b(n);
        for (size_t i = 0; i < n; i++) a[i] = b[i] = i;
        std::sort(a.begin(), a.end(), [&](uint32_t x, uint32_t y) { return key[x] < key[y]; });
        hera::libstdcxx_sort(b.data(), long(n), [&](uint32_t x, uint32_t y) { return key[x] < key[y]; });
        if (a != b) { std::printf("MISMATCH %d %zu\n", round, n); return 1; }
    }
    std::printf("OK\n");
    return 0;
}
'''


def test_device_introsort_header_equals_std_sort(tmp_path):
    src = tmp_path / "t.cpp"
    src.write_text(SRC)
    exe = tmp_path / "t"
    subprocess.run(["g++", "-O2", "-std=c++17", f"-I{common.ROOT / 'telomeri_b200' / 'csrc'}", "-o", str(exe), str(src)], check=True)
    assert subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.startswith("OK")
