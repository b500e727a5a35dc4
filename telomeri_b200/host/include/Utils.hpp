// Host mirror of the helpers main() and the classes use (reference include/Utils.hpp:12-76).
#ifndef TELOMERI_UTILS_H
#define TELOMERI_UTILS_H

#include <algorithm>
#include <cctype>
#include <chrono>
#include <cmath>
#include <sstream>
#include <string>
#include <vector>

#include <OverlapGraph.hpp>

namespace Utils {

inline bool startsWithInsensitive(const std::string &str, const std::string &temp) {
    if (str.size() < temp.size()) return false;
    return std::equal(temp.begin(), temp.end(), str.begin(),
                      [](char a, char b) { return std::tolower(static_cast<unsigned char>(a)) == std::tolower(static_cast<unsigned char>(b)); });
}

template<typename T>
inline bool contains(const std::vector<T> &v, const T &e) { return std::find(v.begin(), v.end(), e) != v.end(); }

enum class Metrics { EXTENSION_SCORE, OVERLAP_SCORE, EXTENSION_SCORE_SQRT, OVERLAP_SCORE_SQRT };  // values = hera_b200.h metric ids

inline float getMetric(const OverlapGraph::Edge &e, const Utils::Metrics &metric) {
    const bool ext = metric == Metrics::EXTENSION_SCORE || metric == Metrics::EXTENSION_SCORE_SQRT;
    const bool root = metric == Metrics::EXTENSION_SCORE_SQRT || metric == Metrics::OVERLAP_SCORE_SQRT;
    const float v = ext ? e.extension_score : e.overlap_score;
    return root ? std::sqrt(v) : v;
}

inline char *getMetricName(const Utils::Metrics &metric) {
    static const char *names[] = {"EXTENSION_SCORE", "OVERLAP_SCORE", "EXTENSION_SCORE_SQRT", "OVERLAP_SCORE_SQRT"};
    const int i = static_cast<int>(metric);
    return const_cast<char *>(i >= 0 && i < 4 ? names[i] : "");
}

}  // namespace Utils

#endif
