// OverlapGraph over the hera_b200 C ABI.  All graph construction happens on the GPU
// (paf_parse_kernel, interning, CSR build); this file only marshals the exported tables into the
// pointer-friendly containers the reference's interface promises.
#include <OverlapGraph.hpp>

#include <algorithm>
#include <cfloat>
#include <climits>
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <stdexcept>

#include "hera_b200.h"

namespace { hera_ctx *g_current = nullptr; }

namespace {
[[noreturn]] void die(hera_ctx *ctx, const char *what) {
    std::cerr << what << ": " << hera_last_error(ctx) << std::endl;
    std::exit(1);
}
}  // namespace

OverlapGraph::OverlapGraph() : filter_params_{AVG, 0, 0.0f, ULONG_MAX, 1.0f}, ctx_(nullptr) {
    const char *dev = std::getenv("HERA_DEVICE");
    ctx_ = hera_create(dev ? std::atoi(dev) : 0);
    g_current = ctx_;
    if (!ctx_) {
        // no CPU fallback: without the CUDA device the graph cannot be built at all
        std::cerr << "hera_b200: " << hera_last_error(nullptr) << std::endl;
        std::exit(1);
    }
}

hera_ctx *OverlapGraph::current_device() { return g_current; }

OverlapGraph::~OverlapGraph() {
    if (g_current == ctx_) g_current = nullptr;
    hera_destroy(ctx_);
}

bool OverlapGraph::load(char *filepath, bool anchors) {
    hera_filter_params fp{static_cast<int32_t>(filter_params_.mode), filter_params_.min_overlap_length, filter_params_.min_overlap_percentage,
                          filter_params_.max_overhang_length, filter_params_.max_overhang_percentage};
    if (hera_graph_set_filter(ctx_, &fp)) die(ctx_, "hera_graph_set_filter");
    hera_load_result r{};
    const int rc = hera_graph_load_file(ctx_, filepath, anchors ? 1 : 0, test_load_num_, &r);
    if (rc == 1) {  // same message and result as src/OverlapGraph.cpp:100-103
        std::cerr << "Cannot open file: " << filepath << std::endl;
        return false;
    }
    if (rc) die(ctx_, "hera_graph_load_file");
    hera_graph_info info{};
    if (hera_graph_finalize(ctx_, 0xF, &info)) die(ctx_, "hera_graph_finalize");
    mirror_from_device();
    std::cout << "Loaded " << edges_.size() << '/' << r.records << " edges from " << filepath << std::endl;
    return true;
}

void OverlapGraph::mirror_from_device() {
    hera_graph_view v{};
    if (hera_graph_export(ctx_, &v)) die(ctx_, "hera_graph_export");
    nodes_.clear();
    edges_.clear();
    nodes_.reserve(v.nodes);
    edges_.reserve(v.edges);
    for (uint64_t i = 0; i < v.nodes; i++) {
        nodes_.emplace_back(v.node_anchor[i] != 0, static_cast<uint>(i), v.node_len[i],
                            std::string(v.node_names + v.node_name_off[i], v.node_names + v.node_name_off[i + 1]));
    }
    for (uint64_t e = 0; e < v.edges; e++) {
        edges_.emplace_back(v.e_q[e], v.e_t[e], v.e_qs[e], v.e_qe[e], v.e_ts[e], v.e_te[e], v.e_os[e], v.e_si[e], v.e_es[e], v.e_strand[e] != 0);
    }
    adj_off_.assign(v.adj_off, v.adj_off + v.nodes + 1);
    for (uint64_t x = 0; x < v.nodes; x++) {
        std::vector<Edge> &list = nodes_[x].edges;
        list.reserve(v.adj_off[x + 1] - v.adj_off[x]);
        for (uint64_t s = v.adj_off[x]; s < v.adj_off[x + 1]; s++) {
            const uint32_t e = v.adj_rec[s];
            if (v.adj_side[s] == 0)  // stored as read: owner is the PAF target
                list.emplace_back(v.e_q[e], v.e_t[e], v.e_qs[e], v.e_qe[e], v.e_ts[e], v.e_te[e], v.e_os[e], v.e_si[e], v.e_es[e], v.e_strand[e] != 0);
            else                     // q/t swapped copy: owner is the PAF query (src/OverlapGraph.cpp:245-248)
                list.emplace_back(v.e_t[e], v.e_q[e], v.e_ts[e], v.e_te[e], v.e_qs[e], v.e_qe[e], v.e_os[e], v.e_si[e], v.e_es[e], v.e_strand[e] != 0);
        }
    }
}

const OverlapGraph::Edge *OverlapGraph::edge_of_slot(uint32_t slot) const {
    const auto it = std::upper_bound(adj_off_.begin(), adj_off_.end(), static_cast<uint64_t>(slot));
    const size_t owner = static_cast<size_t>(it - adj_off_.begin()) - 1;
    return &nodes_[owner].edges[slot - adj_off_[owner]];
}

uint32_t OverlapGraph::slot_of_edge(const Edge *e) const {
    const Node &owner = nodes_[e->t_index];
    return static_cast<uint32_t>(adj_off_[owner.index] + static_cast<uint64_t>(e - owner.edges.data()));
}

// Same report text as src/OverlapGraph.cpp:24-95 (host-side formatting of exported tables).
std::string OverlapGraph::stats() {
    std::stringstream str;
    ulong min_len = ULONG_MAX, max_len = 0, min_con = ULONG_MAX, max_con = 0, anchors = 0, reads = 0;
    for (const Node &n : nodes_) {
        n.anchor ? anchors++ : reads++;
        min_len = std::min<ulong>(min_len, n.length);
        max_len = std::max<ulong>(max_len, n.length);
        min_con = std::min<ulong>(min_con, n.edges.size());
        max_con = std::max<ulong>(max_con, n.edges.size());
    }
    ulong min_OL = ULONG_MAX, max_OL = 0;
    float min_OS = FLT_MAX, max_OS = 0, min_ES = FLT_MAX, max_ES = 0, min_SI = FLT_MAX, max_SI = 0;
    for (const Edge &e : edges_) {
        min_OS = std::min(min_OS, e.overlap_score); max_OS = std::max(max_OS, e.overlap_score);
        min_ES = std::min(min_ES, e.extension_score); max_ES = std::max(max_ES, e.extension_score);
        min_SI = std::min(min_SI, e.sequence_identity); max_SI = std::max(max_SI, e.sequence_identity);
        const int ol = static_cast<int>(std::max(e.q_end - e.q_start, e.t_end - e.t_start));
        const ulong olu = static_cast<ulong>(static_cast<long>(ol));
        min_OL = std::min(min_OL, olu);
        max_OL = std::max(max_OL, olu);
    }
    str << "Nodes\n-  anchor: " << anchors << "\n-    read: " << reads << "\n-   total: " << nodes_.size() << "\n- min_len: " << min_len
        << "\n- max_len: " << max_len << "\n- min_con: " << min_con << "\n- max_con: " << max_con << "\nEdges\n-   total: " << edges_.size()
        << "\n-  min_OL: " << min_OL << "\n-  max_OL: " << max_OL << "\n-  min_OS: " << min_OS << "\n-  max_OS: " << max_OS
        << "\n-  min_ES: " << min_ES << "\n-  max_ES: " << max_ES << "\n-  min_SI: " << min_SI << "\n-  max_SI: " << max_SI << '\n' << std::endl;
    return str.str();
}

OverlapGraph::Node::Node(bool anchor, uint index, uint length, const std::string &name) : anchor(anchor), index(index), length(length), name(name) {}

bool OverlapGraph::Node::operator==(const Node &rhs) const { return name == rhs.name; }

void OverlapGraph::Node::to_stream(std::ostream &s) const {
    s << (anchor ? '*' : ' ') << 'n' << index << "  len: " << length << "  name: " << name << "  edges: " << edges.size() << std::endl;
}

OverlapGraph::Edge::Edge(uint q_index, uint t_index, uint q_start, uint q_end, uint t_start, uint t_end, float overlap_score,
                         float sequence_identity, float extension_score, bool relative_strand)
        : q_index(q_index), t_index(t_index), q_start(q_start), q_end(q_end), t_start(t_start), t_end(t_end), overlap_score(overlap_score),
          sequence_identity(sequence_identity), extension_score(extension_score), relative_strand(relative_strand) {}

OverlapGraph::Edge &OverlapGraph::Edge::operator=(OverlapGraph::Edge &&e) noexcept {
    q_index = e.q_index; t_index = e.t_index; q_start = e.q_start; q_end = e.q_end; t_start = e.t_start; t_end = e.t_end;
    overlap_score = e.overlap_score; sequence_identity = e.sequence_identity; extension_score = e.extension_score;
    relative_strand = e.relative_strand;
    return *this;
}

void OverlapGraph::Edge::to_stream(std::ostream &s) const {
    s << q_index << "-" << t_index << "  OS: " << overlap_score << "  ES: " << extension_score << "  SI: " << sequence_identity << std::endl;
}
