// Test infrastructure: recorded-draw queue used by the replay build of the reference harness.
// The reference's Monte-Carlo walk draws u = dis(gen) once per random step
// (src/PathManager.cpp:104).  In the replay build that expression returns the next value of a
// pre-recorded stream (the Philox draws the oracle / CUDA walks consumed, concatenated in the
// reference's own walk order: anchors in node order x attempts), so an unmodified
// buildMonteCarlo reproduces the counter-RNG walks exactly -- or visibly desynchronises.
#ifndef HERA_REPLAY_STATE_HPP
#define HERA_REPLAY_STATE_HPP

#include <cstdint>
#include <cstdio>
#include <vector>

namespace hera_replay {

struct State {
    std::vector<double> draws;
    uint64_t next = 0;
    uint64_t underflow = 0;
};

inline State &state() {
    static State s;
    return s;
}

inline bool load(const char *path) {
    State &s = state();
    std::FILE *f = std::fopen(path, "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    long bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    s.draws.resize(static_cast<size_t>(bytes) / sizeof(double));
    size_t got = s.draws.empty() ? 0 : std::fread(s.draws.data(), sizeof(double), s.draws.size(), f);
    std::fclose(f);
    s.next = 0;
    return got == s.draws.size();
}

inline double next_draw() {
    State &s = state();
    if (s.next < s.draws.size()) return s.draws[s.next++];
    s.underflow++;
    return 0.5;
}

inline uint64_t consumed() { return state().next; }
inline uint64_t available() { return state().draws.size(); }
inline uint64_t underflow() { return state().underflow; }

}  // namespace hera_replay
#endif
