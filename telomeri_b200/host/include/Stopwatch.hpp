// Host mirror of the reference's Stopwatch (reference include/Stopwatch.hpp:6-21).
#ifndef STOPWATCH_HPP
#define STOPWATCH_HPP

#include <chrono>

class Stopwatch {
    using clock = std::chrono::system_clock;
    clock::time_point start_, lapStart_;
public:
    void start() { start_ = lapStart_ = clock::now(); }
    double lap() {
        const auto now = clock::now();
        const std::chrono::duration<double> d = now - lapStart_;
        lapStart_ = now;
        return d.count();
    }
    double stop() { return std::chrono::duration<double>(clock::now() - start_).count(); }
};

#endif
