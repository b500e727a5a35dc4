// Path (reference src/Path.cpp).  Device-produced paths carry the device-computed length; this
// formula is only evaluated for paths assembled on the host (the scaffold).
#include <Path.hpp>

#include <algorithm>
#include <iostream>

void Path::updateLength() {
    length_ = 0;
    if (nodes_.empty()) return;
    length_ = nodes_[0]->length;
    for (size_t i = 0; i + 1 < edges_.size(); i++) length_ += static_cast<long>(edges_[i]->q_start) - static_cast<long>(edges_[i]->t_start);
}

long Path::length() const { return length_; }

std::ostream &operator<<(std::ostream &s, const Path &p) {
    for (size_t i = 0; i < p.nodes_.size(); i++) {
        if (i) s << "-";
        if (p.nodes_[i]->anchor) s << '*';
        s << 'n' << p.nodes_[i]->index;
    }
    if (p.edges_.size() >= p.nodes_.size() && !p.nodes_.empty()) s << "-";
    return s;
}

bool Path::operator==(const Path &other) const {
    return nodes_.size() == other.nodes_.size() && std::equal(nodes_.begin(), nodes_.end(), other.nodes_.begin());
}
