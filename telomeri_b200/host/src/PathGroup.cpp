// PathGroup (reference src/PathGroup.cpp).  The three steps main() runs per group publish what
// group_pairs_kernel computed for this group; a group that was not created by
// PathManager::constructGroups has no device results and refuses to guess (no CPU fallback).
#include <PathGroup.hpp>

#include <iostream>
#include <stdexcept>
#include <string>

PathGroup::PathGroup(std::vector<const Path *>::const_iterator begin, std::vector<const Path *>::const_iterator end)
        : pig_(begin, end), consensus(nullptr), valid_path_number(0) {
    for (const Path *p : pig_) frqs[static_cast<ulong>(p->length())]++;
}

void PathGroup::require_device_results(const char *what) const {
    if (!has_device_results_)
        throw std::logic_error(std::string("PathGroup::") + what + ": group was not produced by PathManager::constructGroups, no device results attached");
}

void PathGroup::discardNotFrequent() {
    require_device_results("discardNotFrequent");
    pig_ = device_pig_;
    std::map<ulong, int> kept;
    for (const Path *p : pig_) kept[static_cast<ulong>(p->length())]++;
    frqs.swap(kept);
}

void PathGroup::calculateConsensusPath() {
    require_device_results("calculateConsensusPath");
    consensus = device_consensus_;
}

void PathGroup::calculateValidPathNumber() {
    require_device_results("calculateValidPathNumber");
    valid_path_number = consensus ? device_valid_ : 0;
}

std::ostream &operator<<(std::ostream &s, const PathGroup &pg) {
    int i = 0;
    s << '\t';
    for (const Path *p : pg.pig_) {
        s << '[' << p->length() << "] ";
        if (++i % 5 == 0) s << "\n\t";
    }
    return s;
}
