"""The introsort restatement (oracle/hera_oracle_sort.hpp) must produce the SAME permutation as
libstdc++'s std::sort for the comparator the reference uses (ties everywhere)."""
import subprocess

import common

SRC = r'''
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <vector>
#include "hera_oracle_sort.hpp"
struct Rng { uint64_t s; uint64_t next() { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); } };
int main() {
    Rng rng{12345};
    long cases = 0;
    for (int round = 0; round < 4000; round++) {
        size_t n = round < 200 ? round : (rng.next() % 3000);
        uint64_t distinct = 1 + rng.next() % (round % 7 == 0 ? 3 : (n + 1));
        std::vector<long> key(n);
        int shape = round % 6;
        for (size_t i = 0; i < n; i++) {
            switch (shape) {
                case 0: key[i] = rng.next() % distinct; break;
                case 1: key[i] = i % distinct; break;                       // sawtooth
                case 2: key[i] = (n - i) / (1 + distinct % 5); break;       // descending with ties
                case 3: key[i] = i < n / 2 ? i : n - i; break;              // organ pipe
                case 4: key[i] = (i & 1) ? i / 2 : n - i / 2; break;        // interleaved (deep recursion)
                default: key[i] = 7; break;                                 // all equal
            }
        }
        if (shape == 4 && n > 64) {  // median-of-3 killer style input to reach the heap-sort fallback
            size_t k = n / 2;
            for (size_t i = 0; i < k; i++) { key[i] = (i & 1) ? long(k + i) : long(i); }
            for (size_t i = k; i < n; i++) key[i] = long(2 * (i - k));
        }
        std::vector<uint32_t> a(n), b(n);
        for (size_t i = 0; i < n; i++) a[i] = b[i] = i;
        std::sort(a.begin(), a.end(), [&](uint32_t x, uint32_t y) { return key[x] < key[y]; });
        hera_oracle::libstdcxx_sort(b.data(), b.data() + n, [&](uint32_t x, uint32_t y) { return key[x] < key[y]; });
        if (a != b) { std::printf("MISMATCH round %d n %zu shape %d\n", round, n, shape); return 1; }
        // descending comparator as in src/Main.cpp:312-315
        for (size_t i = 0; i < n; i++) a[i] = b[i] = i;
        std::sort(a.begin(), a.end(), [&](uint32_t x, uint32_t y) { return key[x] > key[y]; });
        hera_oracle::libstdcxx_sort(b.data(), b.data() + n, [&](uint32_t x, uint32_t y) { return key[x] > key[y]; });
        if (a != b) { std::printf("MISMATCH desc round %d n %zu shape %d\n", round, n, shape); return 1; }
        cases += 2;
    }
    std::printf("OK %ld\n", cases);
    return 0;
}
'''


def test_introsort_restatement_equals_std_sort(tmp_path):
    src = tmp_path / "sort_check.cpp"
    src.write_text(SRC)
    exe = tmp_path / "sort_check"
    subprocess.run(["g++", "-O2", "-std=c++17", f"-I{common.ORACLE_DIR}", "-o", str(exe), str(src)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    assert out.startswith("OK"), out
