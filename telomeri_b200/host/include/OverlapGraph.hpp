// Host mirror of the reference's OverlapGraph class surface (reference include/OverlapGraph.hpp:9-130)
// over the hera_b200 C ABI.  load() runs the CUDA tokenizer / filter / scoring / interning / CSR
// build and then materialises nodes_ and edges_ from one bulk export, so code written against the
// reference header (its main(), src/Main.cpp:211-226) compiles and behaves the same.
#ifndef OVERLAP_GRAPH_HPP
#define OVERLAP_GRAPH_HPP

#include <sys/types.h>

#include <cstdint>
#include <ostream>
#include <string>
#include <vector>

struct hera_ctx;
enum class ContigPosition : char;

class OverlapGraph {
public:
    struct Edge;

    struct Node {
        bool anchor;
        uint index;
        std::vector<Edge> edges;   // this node's adjacency list, reference insertion order
        uint length;
        std::string name;

        Node(bool anchor, uint index, uint length, const std::string &name);
        bool operator==(const Node &) const;
        void to_stream(std::ostream &) const;
    };

    struct Edge {
        uint q_index, t_index;                 // in a node's list: t_index = owner, q_index = neighbour
        uint q_start, q_end, t_start, t_end;
        float overlap_score, sequence_identity, extension_score;
        bool relative_strand;

        Edge(uint q_index, uint t_index, uint q_start, uint q_end, uint t_start, uint t_end, float overlap_score,
             float sequence_identity, float extension_score, bool relative_strand);
        constexpr Edge(const Edge &) = default;
        Edge &operator=(Edge &&) noexcept;
        void to_stream(std::ostream &) const;
    };

    struct PAFOverlap {
        std::string query_name;
        int query_len, query_start, query_end;
        char relative_strand;
        std::string target_name;
        int target_len, target_start, target_end;
        int residue_matches_num, alignment_block_len, mapping_quality;
    };

    enum LengthCalculationMode { MIN, MAX, AVG, SUM };

    struct FilterParameters {
        LengthCalculationMode mode;
        ulong min_overlap_length;
        float min_overlap_percentage;
        ulong max_overhang_length;
        float max_overhang_percentage;
    };

    std::vector<Node> nodes_;
    std::vector<Edge> edges_;
    FilterParameters filter_params_;
    uint test_load_num_ = 0;

    OverlapGraph();
    ~OverlapGraph();
    OverlapGraph(const OverlapGraph &) = delete;
    OverlapGraph &operator=(const OverlapGraph &) = delete;

    /** Appends the overlaps of a PAF file (CUDA path); false + message on std::cerr if it cannot be read. */
    bool load(char *filepath, bool anchors);
    std::string stats();

    // ---- not in the reference: access for the other host classes
    hera_ctx *device() const { return ctx_; }
    /** the context of the live OverlapGraph (one per process in main()), for Scaffolder */
    static hera_ctx *current_device();
    /** slot index (adj_off[owner] + k) -> pointer to nodes_[owner].edges[k] */
    const Edge *edge_of_slot(uint32_t slot) const;
    uint32_t slot_of_edge(const Edge *e) const;

private:
    hera_ctx *ctx_;
    std::vector<uint64_t> adj_off_;
    void mirror_from_device();
};

enum class ContigPosition : char { NONE, QUERY, TARGET };

#endif
