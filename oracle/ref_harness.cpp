// Test infrastructure: drives the UNMODIFIED reference classes (compiled from
// /root/reference/src by oracle/Makefile into oracle/_ref/) and dumps every intermediate
// result of the hot path in the heradump format, so that the CPU restatement
// (oracle/hera_oracle.cpp) and the CUDA path can be compared with it bit for bit.
//
// Nothing here is product code.  The orchestration mirrors the call order of the reference's
// main() (src/Main.cpp:223-352): load rr, load cr, 2x buildMonteCarlo, 2x buildDeterministic,
// filterUnique, getPathsBetweenAnchors, constructGroups, PathGroup steps, inter-group pick,
// constructConsensusPath, Scaffolder.
//
// Usage: ref_dump [options] <rr.paf> <cr.paf> <dump.bin> [<reads.fa> <contigs.fa> <out.fa>]
//   --mode min|max|avg|sum  --min-oll N --min-olp F --max-ohl N --max-ohp F
//   --rb-att N --bt-att N --len-thr N --w-size N --r-thr F --test-load-num N
//   --metrics es,os,es_sqrt,os_sqrt   (which metrics to run, reference default es,os)
//   --draws <file>   (replay build only: raw little-endian doubles consumed by dis(gen))
#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <tuple>
#include <vector>

// Standard headers are in above; only the reference's own classes see this.
#define private public
#include <PathManager.hpp>
#undef private
#include <PathWindow.hpp>
#include <Scaffolder.hpp>

#include "hera_dump.hpp"

#ifdef HERA_REPLAY
#include "replay_state.hpp"
#endif

using Node = OverlapGraph::Node;
using Edge = OverlapGraph::Edge;
using AnchorPair = std::pair<const Node *, const Node *>;

static std::vector<uint64_t> adjacency_offsets(const OverlapGraph &g) {
    std::vector<uint64_t> off(g.nodes_.size() + 1, 0);
    for (size_t i = 0; i < g.nodes_.size(); i++) off[i + 1] = off[i] + g.nodes_[i].edges.size();
    return off;
}

static void dump_edge_list(heradump::Writer &w, const std::string &prefix, const std::vector<const Edge *> &es) {
    std::vector<uint32_t> q, t, qs, qe, ts, te;
    std::vector<float> os, si, esc;
    std::vector<uint8_t> strand;
    for (const Edge *e : es) {
        q.push_back(e->q_index); t.push_back(e->t_index);
        qs.push_back(e->q_start); qe.push_back(e->q_end);
        ts.push_back(e->t_start); te.push_back(e->t_end);
        os.push_back(e->overlap_score); si.push_back(e->sequence_identity);
        esc.push_back(e->extension_score); strand.push_back(e->relative_strand ? 1 : 0);
    }
    w.put(prefix + "q", q); w.put(prefix + "t", t);
    w.put(prefix + "qs", qs); w.put(prefix + "qe", qe);
    w.put(prefix + "ts", ts); w.put(prefix + "te", te);
    w.put(prefix + "os", os); w.put(prefix + "si", si); w.put(prefix + "es", esc);
    w.put(prefix + "strand", strand);
}

static void dump_paths(heradump::Writer &w, const std::string &prefix, const OverlapGraph &g,
                       const std::vector<uint64_t> &adj_off, const std::vector<Path> &paths) {
    std::vector<uint64_t> off(1, 0);
    std::vector<uint32_t> nodes, slots;
    std::vector<int64_t> len;
    for (const Path &p : paths) {
        for (const Node *n : p.nodes_) nodes.push_back(n->index);
        for (const Edge *e : p.edges_) {
            const Node &owner = g.nodes_[e->t_index];
            slots.push_back(static_cast<uint32_t>(adj_off[owner.index] + (e - owner.edges.data())));
        }
        off.push_back(nodes.size());
        len.push_back(p.length());
    }
    w.put(prefix + "off", off);      // node offsets; path i has off[i+1]-off[i] nodes and one edge fewer
    w.put(prefix + "nodes", nodes);
    w.put(prefix + "slots", slots);  // absolute adjacency slot (adj_off[t_index] + position in that list)
    w.put(prefix + "len", len);
}

int main(int argc, char **argv) {
    OverlapGraph graph;
    graph.filter_params_ = {OverlapGraph::AVG, 0, 0.0f, ULONG_MAX, 1.0f};
    PathManager pm;
    pm.params_ = {2000, 60, 4700000l, 1000ul, 0.9f};
    std::vector<Utils::Metrics> metrics = {Utils::Metrics::EXTENSION_SCORE, Utils::Metrics::OVERLAP_SCORE};
    const char *draws_file = nullptr;
    std::vector<char *> pos;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char * { if (i + 1 >= argc) { std::cerr << "missing value for " << a << "\n"; exit(2); } return argv[++i]; };
        if (a == "--mode") {
            std::string m = next();
            graph.filter_params_.mode = m == "min" ? OverlapGraph::MIN : m == "max" ? OverlapGraph::MAX
                                      : m == "sum" ? OverlapGraph::SUM : OverlapGraph::AVG;
        } else if (a == "--min-oll") graph.filter_params_.min_overlap_length = std::stoul(next());
        else if (a == "--min-olp") graph.filter_params_.min_overlap_percentage = std::stof(next());
        else if (a == "--max-ohl") graph.filter_params_.max_overhang_length = std::stoul(next());
        else if (a == "--max-ohp") graph.filter_params_.max_overhang_percentage = std::stof(next());
        else if (a == "--rb-att") pm.params_.rebuild_attempts = std::stoi(next());
        else if (a == "--bt-att") pm.params_.backtrack_attempts = std::stoi(next());
        else if (a == "--len-thr") pm.params_.len_threshold = std::stol(next());
        else if (a == "--w-size") pm.params_.window_size = std::stoul(next());
        else if (a == "--r-thr") pm.params_.ratio_threshold = std::stof(next());
        else if (a == "--test-load-num") graph.test_load_num_ = static_cast<uint>(std::stoul(next()));
        else if (a == "--draws") draws_file = next();
        else if (a == "--metrics") {
            metrics.clear();
            std::stringstream ss(next());
            std::string m;
            while (std::getline(ss, m, ',')) {
                metrics.push_back(m == "es" ? Utils::Metrics::EXTENSION_SCORE : m == "os" ? Utils::Metrics::OVERLAP_SCORE
                                  : m == "es_sqrt" ? Utils::Metrics::EXTENSION_SCORE_SQRT : Utils::Metrics::OVERLAP_SCORE_SQRT);
            }
        } else pos.push_back(argv[i]);
    }
    if (pos.size() != 3 && pos.size() != 6) {
        std::cerr << "usage: ref_dump [options] rr.paf cr.paf dump.bin [reads.fa contigs.fa out.fa]\n";
        return 2;
    }
#ifdef HERA_REPLAY
    if (draws_file) hera_replay::load(draws_file);
#else
    (void) draws_file;
#endif

    heradump::Writer w(pos[2]);
    if (!w.ok()) { std::cerr << "cannot write " << pos[2] << "\n"; return 2; }

    // ---- graph (src/Main.cpp:223).  "Loaded a/b" goes to stdout only; capture it for ctr.
    std::vector<uint64_t> ctr, cum_edges;
    for (int f = 0; f < 2; f++) {
        std::ostringstream cap;
        std::streambuf *old = std::cout.rdbuf(cap.rdbuf());
        bool ok = graph.load(pos[f], f == 1);
        std::cout.rdbuf(old);
        std::cout << cap.str();
        if (!ok) return 1;
        unsigned long a = 0, b = 0;
        std::sscanf(cap.str().c_str(), "Loaded %lu/%lu", &a, &b);
        cum_edges.push_back(a);
        ctr.push_back(b);
    }
    w.put("load_ctr", ctr);
    w.put("load_cum_edges", cum_edges);
    std::string stats = graph.stats();
    w.text("graph_stats", stats);

    const size_t N = graph.nodes_.size();
    std::vector<uint32_t> nlen(N);
    std::vector<uint8_t> nanchor(N);
    std::vector<uint64_t> name_off(1, 0);
    std::string names;
    for (size_t i = 0; i < N; i++) {
        nlen[i] = graph.nodes_[i].length;
        nanchor[i] = graph.nodes_[i].anchor;
        names += graph.nodes_[i].name;
        name_off.push_back(names.size());
    }
    w.put("node_len", nlen);
    w.put("node_anchor", nanchor);
    w.put("node_name_off", name_off);
    w.text("node_names", names);

    std::vector<const Edge *> ev;
    for (const Edge &e : graph.edges_) ev.push_back(&e);
    dump_edge_list(w, "e_", ev);

    std::vector<uint64_t> adj_off = adjacency_offsets(graph);
    w.put("adj_off", adj_off);
    ev.clear();
    for (const Node &n : graph.nodes_) for (const Edge &e : n.edges) ev.push_back(&e);
    dump_edge_list(w, "a_", ev);

    // ---- walks (src/Main.cpp:237-247): all MC calls first, then all deterministic calls.
    std::vector<uint64_t> found;
    std::vector<uint32_t> call_kind, call_metric;
    for (auto m : metrics) {
        size_t before = pm.paths_.size();
        pm.buildMonteCarlo(graph, m);
        found.push_back(pm.paths_.size() - before);
        call_kind.push_back(0); call_metric.push_back(static_cast<uint32_t>(m));
    }
    for (auto m : metrics) {
        size_t before = pm.paths_.size();
        pm.buildDeterministic(graph, m);
        found.push_back(pm.paths_.size() - before);
        call_kind.push_back(1); call_metric.push_back(static_cast<uint32_t>(m));
    }
    w.put("call_kind", call_kind);      // 0 = Monte Carlo, 1 = deterministic
    w.put("call_metric", call_metric);  // Utils::Metrics enum value
    w.put("call_found", found);
#ifdef HERA_REPLAY
    w.scalar<uint64_t>("replay_consumed", hera_replay::consumed());
    w.scalar<uint64_t>("replay_available", hera_replay::available());
    w.scalar<uint64_t>("replay_underflow", hera_replay::underflow());
#endif
    dump_paths(w, "p_", graph, adj_off, pm.paths_);
    pm.filterUnique();
    dump_paths(w, "u_", graph, adj_off, pm.paths_);
    w.text("path_stats", pm.stats());

    // ---- grouping (src/Main.cpp:251-342)
    auto pba = pm.getPathsBetweenAnchors();
    std::map<AnchorPair, std::vector<PathGroup>> groups_for_anchors;
    std::vector<uint32_t> pair_a1, pair_a2, sorted_paths;
    std::vector<uint64_t> pair_off(1, 0);
    const Path *base = pm.paths_.data();
    for (auto &kv : pba) {
        pair_a1.push_back(kv.first.first->index);
        pair_a2.push_back(kv.first.second->index);
        groups_for_anchors[kv.first] = PathManager::constructGroups(kv.second, pm.params_);
        for (const Path *p : kv.second) sorted_paths.push_back(static_cast<uint32_t>(p - base));
        pair_off.push_back(sorted_paths.size());
    }
    w.put("pair_a1", pair_a1);
    w.put("pair_a2", pair_a2);
    w.put("pair_off", pair_off);
    w.put("pair_sorted", sorted_paths);  // unique-path ids per pair, in the order constructGroups left them

    std::vector<uint32_t> grp_pair, grp_size0, pig;
    std::vector<uint64_t> pig_off(1, 0);
    std::vector<int64_t> grp_consensus, pair_consensus;
    std::vector<int32_t> grp_valid;
    std::map<AnchorPair, const Path *> consensus_for_anchors;
    long consensus_num = 0;
    uint32_t pair_idx = 0;
    for (auto &kv : groups_for_anchors) {
        std::vector<PathGroup *> with_consensus;
        for (PathGroup &pg : kv.second) {
            grp_pair.push_back(pair_idx);
            grp_size0.push_back(static_cast<uint32_t>(pg.pig_.size()));
            pg.discardNotFrequent();
            pg.calculateConsensusPath();
            pg.calculateValidPathNumber();
            for (const Path *p : pg.pig_) pig.push_back(static_cast<uint32_t>(p - base));
            pig_off.push_back(pig.size());
            grp_consensus.push_back(pg.consensus ? pg.consensus - base : -1);
            grp_valid.push_back(pg.valid_path_number);
            if (pg.consensus) with_consensus.push_back(&pg);
        }
        // Inter-group pick, same decisions as src/Main.cpp:306-332.
        const Path *chosen = nullptr;
        if (with_consensus.size() == 1) {
            chosen = with_consensus[0]->consensus;
        } else if (with_consensus.size() >= 2) {
            std::sort(with_consensus.begin(), with_consensus.end(), [](const PathGroup *a, const PathGroup *b) {
                return a->consensus->length() > b->consensus->length();
            });
            const PathGroup *keep = with_consensus[0];
            if (with_consensus.size() > 2) {
                for (size_t i = 1; i < with_consensus.size(); i++) {
                    if (2 * keep->valid_path_number <= with_consensus[i]->valid_path_number) keep = with_consensus[i];
                }
            }
            chosen = keep->consensus;
        }
        consensus_for_anchors[kv.first] = chosen;
        pair_consensus.push_back(chosen ? chosen - base : -1);
        if (chosen) consensus_num++;
        pair_idx++;
    }
    w.put("grp_pair", grp_pair);
    w.put("grp_size0", grp_size0);  // group size before discardNotFrequent (consecutive ranges of pair_sorted)
    w.put("grp_pig_off", pig_off);
    w.put("grp_pig", pig);          // pig_ after discardNotFrequent
    w.put("grp_consensus", grp_consensus);
    w.put("grp_valid", grp_valid);
    w.put("pair_consensus", pair_consensus);

    // ---- scaffold (src/Main.cpp:344-388)
    std::vector<uint32_t> sc_nodes, sc_slots;
    uint8_t have_scaffold = 0;
    Path scaffold;
    if (consensus_num > 0) {
        try {
            scaffold = pm.constructConsensusPath(pba, consensus_for_anchors, 10);
            have_scaffold = 1;
        } catch (const std::exception &e) {
            std::cerr << "constructConsensusPath threw: " << e.what() << "\n";
        }
    }
    if (have_scaffold) {
        for (const Node *n : scaffold.nodes_) sc_nodes.push_back(n->index);
        for (const Edge *e : scaffold.edges_) {
            const Node &owner = graph.nodes_[e->t_index];
            sc_slots.push_back(static_cast<uint32_t>(adj_off[owner.index] + (e - owner.edges.data())));
        }
    }
    w.scalar<uint8_t>("have_scaffold", have_scaffold);
    w.put("scaffold_nodes", sc_nodes);
    w.put("scaffold_slots", sc_slots);

    uint8_t fasta_ok = 0;
    if (have_scaffold && pos.size() == 6) {
        Scaffolder scaff(scaffold);
        if (!scaff.load(pos[3]) || !scaff.load(pos[4])) return 1;
        try {  // the reference's substr calls throw on some scaffolds (src/Scaffolder.cpp:62,111)
            fasta_ok = scaff.write(pos[5]) ? 1 : 0;
        } catch (const std::exception &e) {
            std::cerr << "Scaffolder::write threw: " << e.what() << "\n";
            std::remove(pos[5]);
        }
    }
    w.scalar<uint8_t>("fasta_ok", fasta_ok);
    return 0;
}
