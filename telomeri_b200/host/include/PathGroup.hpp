// Host mirror of the reference's PathGroup (reference include/PathGroup.hpp:8-56).  Groups come
// out of PathManager::constructGroups with the device results attached; the three methods main()
// calls (src/Main.cpp:290-293) publish those results step by step.
#ifndef PATHGROUP_HPP
#define PATHGROUP_HPP

#include <Path.hpp>

#include <map>

class PathGroup {
public:
    std::vector<const Path *> pig_;
    std::map<ulong, int> frqs;
    const Path *consensus;
    int valid_path_number;

    PathGroup(std::vector<const Path *>::const_iterator begin, std::vector<const Path *>::const_iterator end);

    void discardNotFrequent();
    void calculateConsensusPath();
    void calculateValidPathNumber();

private:
    friend class PathManager;
    friend std::ostream &operator<<(std::ostream &s, const PathGroup &pg);
    // device results (group_pairs_kernel), attached by constructGroups
    bool has_device_results_ = false;
    std::vector<const Path *> device_pig_;
    const Path *device_consensus_ = nullptr;
    int device_valid_ = 0;
    void require_device_results(const char *what) const;
    static constexpr ulong CONSENSUS_THRESHOLD = 100000;
};

#endif
