// PathManager over the hera_b200 C ABI: walks (walk_mc_kernel / walk_det_kernel), the unique
// filter and grouping (group_pairs_kernel) run on the GPU; this file keeps the host-side paths_
// vector and the pointer maps of the reference interface consistent with the device tables.
#include <PathManager.hpp>
#include <Scaffolder.hpp>

#include <cstdlib>
#include <iostream>
#include <stdexcept>

#include "hera_b200.h"

PathManager *PathManager::grouping_owner_ = nullptr;

namespace {
[[noreturn]] void die(hera_ctx *ctx, const char *what) {
    std::cerr << what << ": " << hera_last_error(ctx) << std::endl;
    std::exit(1);
}
hera_path_params to_c(const PathManager::Parameters &p) {
    return hera_path_params{p.rebuild_attempts, p.backtrack_attempts, p.len_threshold, p.window_size, p.ratio_threshold};
}
// One hera_groups call evaluates every anchor pair; main() then asks for the pairs one by one.
struct GroupCache {
    const PathManager *owner = nullptr;
    bool valid = false;
    PathManager::Parameters params{};
    hera_groups_view view{};
} g_groups;
bool same_params(const PathManager::Parameters &a, const PathManager::Parameters &b) {
    return a.rebuild_attempts == b.rebuild_attempts && a.backtrack_attempts == b.backtrack_attempts && a.len_threshold == b.len_threshold &&
           a.window_size == b.window_size && a.ratio_threshold == b.ratio_threshold;
}
}  // namespace

unsigned long long PathManager::monte_carlo_seed() {
    const char *s = std::getenv("HERA_SEED");
    return s ? std::strtoull(s, nullptr, 10) : 0ull;
}

void PathManager::pull_paths() {
    hera_paths_view v{};
    if (hera_paths_export(graph_->device(), &v)) die(graph_->device(), "hera_paths_export");
    g_groups.valid = false;
    paths_.clear();
    paths_.resize(v.paths);
    for (uint64_t i = 0; i < v.paths; i++) {
        Path &p = paths_[i];
        const uint64_t b = v.off[i], n = v.off[i + 1] - b;
        p.nodes_.reserve(n);
        p.edges_.reserve(n - 1);
        for (uint64_t k = 0; k < n; k++) p.nodes_.push_back(&graph_->nodes_[v.nodes[b + k]]);
        for (uint64_t k = 0; k + 1 < n; k++) p.edges_.push_back(graph_->edge_of_slot(v.slots[b - i + k]));
        p.length_ = v.len[i];
        p.device_id_ = static_cast<long>(i);
    }
}

void PathManager::buildMonteCarlo(const OverlapGraph &g, const Utils::Metrics &metric) {
    graph_ = &g;
    std::cout << "> Monte Carlo heuristic: " << Utils::getMetricName(metric) << std::endl;
    hera_path_params pp = to_c(params_);
    hera_walk_stats st{};
    if (hera_paths_monte_carlo(g.device(), static_cast<int>(metric), &pp, monte_carlo_seed(), &st)) die(g.device(), "hera_paths_monte_carlo");
    pull_paths();
    std::cout << "Found " << st.found << " paths." << std::endl;
}

void PathManager::buildDeterministic(const OverlapGraph &g, const Utils::Metrics &metric) {
    graph_ = &g;
    std::cout << "> Deterministic heuristic: " << Utils::getMetricName(metric) << std::endl;
    hera_walk_stats st{};
    if (hera_paths_deterministic(g.device(), static_cast<int>(metric), &st)) die(g.device(), "hera_paths_deterministic");
    pull_paths();
    std::cout << "Found " << st.found << " paths." << std::endl;
}

void PathManager::filterUnique() {
    if (!graph_) return;  // no build call yet: nothing to filter
    uint64_t n = 0;
    if (hera_paths_unique(graph_->device(), &n)) die(graph_->device(), "hera_paths_unique");
    pull_paths();
}

std::tuple<ulong, ulong, ulong> PathManager::getMinMaxSumPathLength() {
    ulong min_len = ULONG_MAX, max_len = 0, sum_len = 0;
    for (const Path &p : paths_) {
        const ulong l = static_cast<ulong>(p.length());
        sum_len += l;
        min_len = std::min(min_len, l);
        max_len = std::max(max_len, l);
    }
    return std::make_tuple(min_len, max_len, sum_len);
}

std::string PathManager::stats() {  // text of src/PathManager.cpp:440-465
    std::stringstream str;
    ulong min_len, max_len, sum_len;
    std::tie(min_len, max_len, sum_len) = getMinMaxSumPathLength();
    ulong negatives = 0;
    for (const Path &p : paths_) negatives += p.length() < 0;
    str << "Paths\n- total_num: " << paths_.size() << "\n-   min_len: " << min_len << "\n-   max_len: " << max_len
        << "\n-   avg_len: " << (paths_.size() > 0 ? sum_len / paths_.size() : 0) << std::endl;
    if (negatives > 0) str << "- negatives: " << negatives << std::endl;
    return str.str();
}

std::pair<ulong, ulong> PathManager::getMinMaxPathLength(std::vector<const Path *> &v) {
    ulong min_len = ULONG_MAX, max_len = 0;
    for (const Path *p : v) {
        min_len = std::min(min_len, static_cast<ulong>(p->length()));
        max_len = std::max(max_len, static_cast<ulong>(p->length()));
    }
    return {min_len, max_len};
}

// Pointer bookkeeping only: key = (first node, last node), values in paths_ order
// (src/PathManager.cpp:626-645).  The device does its own keying inside hera_groups.
std::map<std::pair<const OverlapGraph::Node *, const OverlapGraph::Node *>, std::vector<const Path *>> PathManager::getPathsBetweenAnchors() {
    std::map<std::pair<const OverlapGraph::Node *, const OverlapGraph::Node *>, std::vector<const Path *>> out;
    for (const Path &p : paths_) out[{p.nodes_.front(), p.nodes_.back()}].push_back(&p);
    grouping_owner_ = this;
    return out;
}

// Static in the reference, so the instance is found through the paths themselves: v holds pointers
// into the paths_ of the PathManager that last handed out getPathsBetweenAnchors().  The device
// evaluates all anchor pairs in one hera_groups call; this function returns the slice for v's pair,
// puts v into the device's sorted order and attaches the per-group results.
std::vector<PathGroup> PathManager::constructGroups(std::vector<const Path *> &v, PathManager::Parameters params) {
    PathManager *pm = grouping_owner_;
    if (!pm || !pm->graph_ || v.empty() || v.front()->device_id_ < 0)
        throw std::logic_error("PathManager::constructGroups: paths are not backed by the device path store (no CPU fallback)");
    hera_ctx *ctx = pm->graph_->device();
    hera_path_params pp = to_c(params);
    if (!g_groups.valid || g_groups.owner != pm || !same_params(g_groups.params, params)) {
        if (hera_groups(ctx, &pp, &g_groups.view)) die(ctx, "hera_groups");
        g_groups.owner = pm;
        g_groups.params = params;
        g_groups.valid = true;
    }
    const hera_groups_view &gv = g_groups.view;
    const uint32_t a1 = v.front()->nodes_.front()->index, a2 = v.front()->nodes_.back()->index;
    uint64_t k = 0;
    while (k < gv.pairs && !(gv.pair_a1[k] == a1 && gv.pair_a2[k] == a2)) k++;
    if (k == gv.pairs || gv.pair_off[k + 1] - gv.pair_off[k] != v.size())
        throw std::logic_error("PathManager::constructGroups: vector does not match a device anchor pair");
    const Path *base = pm->paths_.data();
    for (uint64_t i = 0; i < v.size(); i++) v[i] = base + gv.pair_sorted[gv.pair_off[k] + i];  // the reference sorts v in place (:510-513)
    std::vector<PathGroup> groups;
    uint64_t begin = 0;
    for (uint64_t g = 0; g < gv.groups; g++) {
        if (gv.grp_pair[g] != k) continue;
        groups.emplace_back(v.begin() + static_cast<long>(begin), v.begin() + static_cast<long>(begin + gv.grp_size0[g]));
        begin += gv.grp_size0[g];
        PathGroup &pg = groups.back();
        pg.has_device_results_ = true;
        for (uint64_t i = gv.grp_pig_off[g]; i < gv.grp_pig_off[g + 1]; i++) pg.device_pig_.push_back(base + gv.grp_pig[i]);
        pg.device_consensus_ = gv.grp_consensus[g] >= 0 ? base + gv.grp_consensus[g] : nullptr;
        pg.device_valid_ = gv.grp_valid[g];
    }
    return groups;
}

// Scaffold chaining (src/PathManager.cpp:664-767): O(#anchor pairs) host bookkeeping on the device's
// consensus choices; kept on the host.
Path PathManager::constructConsensusPath(
        const std::map<std::pair<const OverlapGraph::Node *, const OverlapGraph::Node *>, std::vector<const Path *>> &between,
        std::map<std::pair<const OverlapGraph::Node *, const OverlapGraph::Node *>, const Path *> &consensus, ulong min_path_num) {
    using Key = std::pair<const OverlapGraph::Node *, const OverlapGraph::Node *>;
    Scaffolder::min_path_num_ = min_path_num;   // Scaffolder::write hands the same threshold to hera_scaffold
    std::map<Key, std::pair<const Path *, ulong>> pool;
    std::vector<const OverlapGraph::Node *> anchors;
    ulong best = 0;
    Key best_key{nullptr, nullptr};
    for (const auto &kv : between) {
        const Path *c = consensus.at(kv.first);
        const ulong num = kv.second.size();
        if (!c || num < min_path_num) continue;
        for (const OverlapGraph::Node *a : {kv.first.first, kv.first.second})
            if (!Utils::contains(anchors, a)) anchors.push_back(a);
        pool[kv.first] = {c, num};
        if (best < num) { best = num; best_key = kv.first; }
    }
    Path scaffold = *consensus.at(best_key);  // throws std::out_of_range when no pair qualifies, like the reference
    const OverlapGraph::Node *left = best_key.first, *right = best_key.second;
    const std::vector<const OverlapGraph::Node *> seen = {left, right};  // never grows in the reference (:708-710)
    while (!pool.empty()) {
        bool grow_left = false, any = false;
        ulong top = 0;
        Key pick = best_key;
        for (const OverlapGraph::Node *n : anchors) {
            if (Utils::contains(seen, n)) continue;
            auto it = pool.find(Key{n, left});  // cursors move inside the scan (:722-738)
            if (it != pool.end() && top < it->second.second) { top = it->second.second; pick = it->first; grow_left = true; any = true; left = n; }
            it = pool.find(Key{right, n});
            if (it != pool.end() && top < it->second.second) { top = it->second.second; pick = it->first; grow_left = false; any = true; right = n; }
        }
        if (!any) break;
        const Path *p = pool[pick].first;
        if (!grow_left) {
            scaffold.nodes_.insert(scaffold.nodes_.end(), p->nodes_.begin() + 1, p->nodes_.end());
            scaffold.edges_.insert(scaffold.edges_.end(), p->edges_.begin(), p->edges_.end());
        } else {
            scaffold.nodes_.insert(scaffold.nodes_.begin(), p->nodes_.begin(), p->nodes_.begin() + static_cast<long>(p->edges_.size()));
            scaffold.edges_.insert(scaffold.edges_.begin(), p->edges_.begin(), p->edges_.end());
        }
        pool.erase(pick);
    }
    scaffold.device_id_ = -1;
    return scaffold;
}
