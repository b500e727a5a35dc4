// Test infrastructure: Philox4x32-10 (Salmon et al., SC'11 "Parallel random numbers: as easy
// as 1, 2, 3") restated for the oracle.  The CUDA walks carry their own copy in
// telomeri_b200/csrc; the two are compared through the known-answer vectors in
// tests/test_oracle_philox.py and through walk parity.
#ifndef HERA_ORACLE_PHILOX_HPP
#define HERA_ORACLE_PHILOX_HPP

#include <cstdint>

namespace hera_oracle {

struct Philox4 { uint32_t v[4]; };

inline Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int round = 0; round < 10; round++) {
        uint64_t p0 = static_cast<uint64_t>(M0) * c0;
        uint64_t p1 = static_cast<uint64_t>(M1) * c2;
        uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = static_cast<uint32_t>(p1);
        uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = static_cast<uint32_t>(p0);
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    return Philox4{{c0, c1, c2, c3}};
}

// One Monte-Carlo draw.  counter = (anchor node index, attempt, draw number within the walk,
// metric id), key = seed.  The 64 random bits are turned into [0,1) the way libstdc++'s
// generate_canonical<double,53> does for a 32-bit engine (the reference's dis(gen),
// src/PathManager.cpp:13-14,104): u = rn(x0 + x1*2^32) / 2^64, clamped below 1.
inline double mc_draw(uint64_t seed, uint32_t metric, uint32_t anchor, uint32_t attempt, uint32_t draw) {
    Philox4 r = philox4x32_10(anchor, attempt, draw, metric, static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    double sum = static_cast<double>(r.v[0]) + static_cast<double>(r.v[1]) * 4294967296.0;
    double u = sum / 18446744073709551616.0;
    if (u >= 1.0) u = 0.99999999999999988897769753748434595763683319091796875;  // nextafter(1, 0)
    return u;
}

}  // namespace hera_oracle
#endif
