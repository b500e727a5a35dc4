// Host mirror of the reference's PathManager (reference include/PathManager.hpp:12-65) over the
// hera_b200 C ABI: walks, unique filter and grouping run on the GPU, this class keeps paths_ (and the
// pointer-based views main() expects) in step with the device path store.
#ifndef TELOMERI_PATHMANAGER_HPP
#define TELOMERI_PATHMANAGER_HPP

#include <map>
#include <sstream>
#include <tuple>

#include <OverlapGraph.hpp>
#include <Path.hpp>
#include <PathGroup.hpp>
#include <Utils.hpp>

class PathManager {
private:
    std::vector<Path> paths_;
    const OverlapGraph *graph_ = nullptr;
    void pull_paths();                       // paths_ <- device path store
    static PathManager *grouping_owner_;     // instance whose paths the static constructGroups serves
public:
    struct Parameters {
        int rebuild_attempts;
        int backtrack_attempts;
        long len_threshold;
        ulong window_size;
        float ratio_threshold;
    };
    Parameters params_;

    /** Seed of the counter RNG that replaces the reference's shared std::mt19937 (environment
     *  variable HERA_SEED, default 0). */
    static unsigned long long monte_carlo_seed();

    void buildMonteCarlo(const OverlapGraph &g, const Utils::Metrics &metric);
    void buildDeterministic(const OverlapGraph &g, const Utils::Metrics &metric);
    void filterUnique();

    std::tuple<ulong, ulong, ulong> getMinMaxSumPathLength();
    std::string stats();

    static std::vector<PathGroup> constructGroups(std::vector<const Path *> &v, PathManager::Parameters params);
    static std::pair<ulong, ulong> getMinMaxPathLength(std::vector<const Path *> &v);

    std::map<std::pair<const OverlapGraph::Node *, const OverlapGraph::Node *>, std::vector<const Path *>> getPathsBetweenAnchors();

    Path constructConsensusPath(
            const std::map<std::pair<const OverlapGraph::Node *, const OverlapGraph::Node *>, std::vector<const Path *>> &,
            std::map<std::pair<const OverlapGraph::Node *, const OverlapGraph::Node *>, const Path *> &, ulong min_path_num);
};

#endif
