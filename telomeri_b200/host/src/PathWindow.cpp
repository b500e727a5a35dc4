// PathWindow (reference src/PathWindow.cpp): kept for source compatibility / printing; the windows
// that decide the groups are built on the device inside group_pairs_kernel.
#include <PathWindow.hpp>

#include <ostream>

PathWindow::PathWindow(ulong l, ulong u, const std::vector<const Path *> &sp) : sum_frqs(0) {
    for (const Path *p : sp) {
        const ulong len = static_cast<ulong>(p->length());
        if (len >= u) break;
        if (len < l) continue;
        piw_.push_back(p);
        frqs[len]++;
        sum_frqs++;
    }
}

std::pair<ulong, int> PathWindow::getLowestFrequencyEntry() const {
    auto best = frqs.begin();
    for (auto it = frqs.begin(); it != frqs.end(); ++it) if (it->second < best->second) best = it;
    return *best;
}

std::pair<ulong, int> PathWindow::getHighestFrequencyEntry() const {
    auto best = frqs.begin();
    for (auto it = frqs.begin(); it != frqs.end(); ++it) if (best->second < it->second) best = it;
    return *best;
}

std::ostream &operator<<(std::ostream &s, const PathWindow &pw) {
    for (const Path *p : pw.piw_) s << '[' << p->length() << "] ";
    return s;
}
