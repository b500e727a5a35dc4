// TEST INFRASTRUCTURE -- not product code.  CPU restatement of the reference's repeat-resolution
// hot path (lirfu/Telomeri @ eb7ab23), written from the behaviour of the reference sources
// cited per function in hera_oracle.cpp.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may build or execute it; the product (telomeri_b200/)
// never links or calls it.
//
// Parity status: PINNED.  tests/test_oracle_vs_reference.py compares every array this code
// produces with the unmodified reference classes (oracle/_ref/ref_dump, built from
// /root/reference by oracle/Makefile) on the bundled test/res/EColi_CR_overlaps.paf under
// four filter modes and on synthetic datasets, including Monte-Carlo with the reference's own
// default-seeded std::mt19937 stream; tests/golden/ holds fixtures generated that way.
#ifndef HERA_ORACLE_HPP
#define HERA_ORACLE_HPP

#include <cstdint>
#include <string>
#include <vector>

namespace hera_oracle {

enum LengthMode { MODE_MIN = 0, MODE_MAX = 1, MODE_AVG = 2, MODE_SUM = 3 };  // include/OverlapGraph.hpp:56-65
enum Metric { EXTENSION_SCORE = 0, OVERLAP_SCORE = 1, EXTENSION_SCORE_SQRT = 2, OVERLAP_SCORE_SQRT = 3 };  // include/Utils.hpp:43-45

struct FilterParams {  // include/OverlapGraph.hpp:67-78, defaults src/Main.cpp:56-62
    int mode = MODE_AVG;
    uint64_t min_overlap_length = 0;
    float min_overlap_percentage = 0.0f;
    uint64_t max_overhang_length = UINT64_MAX;
    float max_overhang_percentage = 1.0f;
};

struct PathParams {  // include/PathManager.hpp:27-40, defaults src/Main.cpp:63-69
    int rebuild_attempts = 2000;
    int backtrack_attempts = 60;
    int64_t len_threshold = 4700000;
    uint64_t window_size = 1000;
    float ratio_threshold = 0.9f;
};

struct PafRecord {  // include/OverlapGraph.hpp:45-54
    std::string qname, tname;
    int32_t qlen, qs, qe, tlen, ts, te, matches, block, mapq;
    char strand;
};

// Graph in flat form.  Kept overlaps in file order (rr file first), CSR adjacency in the
// reference's per-node insertion order.
struct Graph {
    // nodes
    std::vector<uint32_t> node_len;
    std::vector<uint8_t> node_anchor;
    std::vector<std::string> node_name;
    // kept overlaps ("edges_")
    std::vector<uint32_t> e_q, e_t, e_qs, e_qe, e_ts, e_te;
    std::vector<float> e_os, e_si, e_es;
    std::vector<uint8_t> e_strand;
    // adjacency: slot s in [adj_off[x], adj_off[x+1]) belongs to node x.
    std::vector<uint64_t> adj_off;
    std::vector<uint32_t> adj_rec;   // kept-overlap index the slot was made from
    std::vector<uint8_t> adj_side;   // 0 = stored as read (owner is the PAF target), 1 = q/t swapped (owner is the PAF query)
    // bookkeeping per load() call
    std::vector<uint64_t> load_ctr, load_cum_edges;

    size_t nodes() const { return node_len.size(); }
    size_t edges() const { return e_q.size(); }
    size_t slots() const { return adj_rec.size(); }
    uint32_t slot_neighbour(uint64_t s) const { return adj_side[s] ? e_t[adj_rec[s]] : e_q[adj_rec[s]]; }
    uint32_t slot_owner(uint64_t s) const { return adj_side[s] ? e_q[adj_rec[s]] : e_t[adj_rec[s]]; }
    // Edge fields as the owner's list stores them (q_* = neighbour's coordinates, t_* = owner's).
    uint32_t slot_qs(uint64_t s) const { return adj_side[s] ? e_ts[adj_rec[s]] : e_qs[adj_rec[s]]; }
    uint32_t slot_qe(uint64_t s) const { return adj_side[s] ? e_te[adj_rec[s]] : e_qe[adj_rec[s]]; }
    uint32_t slot_ts(uint64_t s) const { return adj_side[s] ? e_qs[adj_rec[s]] : e_ts[adj_rec[s]]; }
    uint32_t slot_te(uint64_t s) const { return adj_side[s] ? e_qe[adj_rec[s]] : e_te[adj_rec[s]]; }
    int64_t slot_delta(uint64_t s) const { return static_cast<int64_t>(slot_qs(s)) - static_cast<int64_t>(slot_ts(s)); }
    float slot_metric(uint64_t s, int metric) const;
};

// Incremental loader: call load_text twice (rr then cr) like src/Main.cpp:223, then finalize().
class GraphBuilder {
public:
    explicit GraphBuilder(const FilterParams &fp, uint32_t test_load_num = 0) : fp_(fp), test_load_num_(test_load_num) {}
    void load_text(const std::string &text, bool anchors);
    bool load_file(const char *path, bool anchors);
    Graph finalize();

private:
    struct Kept { PafRecord r; bool is_cr; };
    FilterParams fp_;
    uint32_t test_load_num_;
    std::vector<Kept> kept_;
    std::vector<uint64_t> ctr_, cum_;
};

bool filter_record(const PafRecord &r, const FilterParams &fp);
void score_record(const PafRecord &r, float &si, float &os, float &es);
// Faithful char-stream record reader (what `fs >> string >> int ... ` accepts); returns the
// number of records consumed and appends them to out.  stop_after: 0 = whole stream.
uint64_t read_records(const std::string &text, uint64_t stop_after, std::vector<PafRecord> &out);

// Stable descending rank permutation of every node's list for one metric
// (src/PathManager.cpp:289-296): rank_slot[adj_off[x] + k] = absolute slot of x's k-th best edge.
std::vector<uint32_t> rank_permutation(const Graph &g, int metric);
// canon[s] = first slot in the owner's list with the same neighbour (src/PathManager.cpp:340-347).
std::vector<uint32_t> canonical_slots(const Graph &g);

struct PathSet {
    std::vector<uint64_t> off{0};      // node offsets
    std::vector<uint32_t> nodes;
    std::vector<uint32_t> slots;       // one fewer than nodes per path
    std::vector<int64_t> len;
    size_t size() const { return len.size(); }
    void append(const std::vector<uint32_t> &n, const std::vector<uint32_t> &s, int64_t l);
    bool same_nodes(size_t a, size_t b) const;
};

int64_t path_length(const Graph &g, const std::vector<uint32_t> &nodes, const std::vector<uint32_t> &slots);

enum RngMode { RNG_MT19937 = 0, RNG_PHILOX = 1 };
struct McStats {
    std::vector<uint32_t> draws_per_walk;  // canonical walk order: anchors in node order x attempts
    std::vector<double> draws;             // the values drawn, concatenated in that order
};
// src/PathManager.cpp:12-177.  Returns the number of paths found (appended to ps).
uint64_t walk_monte_carlo(const Graph &g, int metric, const PathParams &pp, RngMode mode, uint64_t seed, PathSet &ps, McStats *stats);
// src/PathManager.cpp:179-430.
uint64_t walk_deterministic(const Graph &g, int metric, PathSet &ps);
// src/PathManager.cpp:432-438 (std::unique: consecutive duplicates only).
PathSet filter_unique(const PathSet &ps);

struct Grouping {
    std::vector<uint32_t> pair_a1, pair_a2;
    std::vector<uint64_t> pair_off{0};
    std::vector<uint32_t> pair_sorted;   // path ids per pair after the (unstable) sort by length
    std::vector<uint32_t> grp_pair, grp_size0;
    std::vector<uint64_t> grp_pig_off{0};
    std::vector<uint32_t> grp_pig;       // members after discardNotFrequent
    std::vector<int64_t> grp_consensus;  // path id or -1
    std::vector<int32_t> grp_valid;
    std::vector<int64_t> pair_consensus; // inter-group pick, path id or -1
};
// src/PathManager.cpp:508-645, src/PathWindow.cpp, src/PathGroup.cpp, src/Main.cpp:277-342.
Grouping group_paths(const PathSet &ps, const PathParams &pp);

struct Scaffold {
    bool ok = false;
    std::vector<uint32_t> nodes, slots;
};
// src/PathManager.cpp:664-767 with min_path_num (src/Main.cpp:352 passes 10).
Scaffold chain_scaffold(const PathSet &ps, const Grouping &gr, uint64_t min_path_num);

// src/Scaffolder.cpp:12-147.  names/seqs are both FASTA files concatenated (reads then contigs).
bool load_fasta(const char *path, std::vector<std::string> &names, std::vector<std::string> &seqs);
bool stitch_scaffold(const Graph &g, const Scaffold &sc, const std::vector<std::string> &names,
                     const std::vector<std::string> &seqs, std::string &fasta_text, std::string &log_line);

// libstdc++ (GCC 13) std::sort restated on an index array so the tie order of the reference's
// two unstable sorts (src/PathManager.cpp:510-513, src/Main.cpp:312-315) is reproduced.
// less(a, b) compares the VALUES stored at two positions.
template <typename T, typename Less>
void libstdcxx_sort(T *first, T *last, Less less);

std::string graph_stats(const Graph &g);          // src/OverlapGraph.cpp:24-95
std::string path_stats(const PathSet &ps);        // src/PathManager.cpp:440-465

}  // namespace hera_oracle

#include "hera_oracle_sort.hpp"
#endif
