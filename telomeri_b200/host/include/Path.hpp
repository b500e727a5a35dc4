// Host mirror of the reference's Path (reference include/Path.hpp:7-21).  Paths are produced on the
// device; nodes_/edges_ point into the OverlapGraph mirror, length_ is the device-computed value.
#ifndef PATH_HPP
#define PATH_HPP

#include <OverlapGraph.hpp>
#include <climits>

class Path {
private:
    long length_ = 0;
    long device_id_ = -1;   // index in the device path store this object mirrors (-1: built on the host)
    friend class PathManager;
    friend class PathGroup;
public:
    std::vector<const OverlapGraph::Node *> nodes_;
    std::vector<const OverlapGraph::Edge *> edges_;

    /** src/Path.cpp:6-16 (used for host-built paths such as the scaffold). */
    void updateLength();
    long length() const;
    friend std::ostream &operator<<(std::ostream &s, const Path &p);
    bool operator==(const Path &other) const;
};

#endif
