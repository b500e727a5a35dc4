// Test infrastructure: force-included (-include) ONLY when compiling the reference's
// src/PathManager.cpp for the replay harness.  It swaps the two std names that
// buildMonteCarlo instantiates (src/PathManager.cpp:13-14) for stand-ins whose dis(gen)
// pops the recorded-draw queue.  The reference source file itself is compiled unmodified,
// straight from /root/reference.
#ifndef HERA_REPLAY_SHIM_HPP
#define HERA_REPLAY_SHIM_HPP

// Pull in everything PathManager.cpp includes first, so the macros below only touch its body.
#include <algorithm>
#include <bitset>
#include <chrono>
#include <climits>
#include <cmath>
#include <iomanip>
#include <iostream>
#include <map>
#include <random>
#include <set>
#include <sstream>
#include <string>
#include <tuple>
#include <vector>

#include "replay_state.hpp"

namespace std {
struct hera_replay_engine {};
template <typename T = double>
struct hera_replay_dist {
    hera_replay_dist(T, T) {}
    T operator()(hera_replay_engine &) { return static_cast<T>(hera_replay::next_draw()); }
};
}  // namespace std

#define mt19937 hera_replay_engine
#define uniform_real_distribution hera_replay_dist
#endif
