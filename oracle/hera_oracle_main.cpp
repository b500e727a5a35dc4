// TEST INFRASTRUCTURE: command-line front end of the CPU restatement.  Writes the same heradump
// sections as oracle/ref_harness.cpp (so dumps can be compared array by array) plus the
// derived arrays the CUDA path is checked against (adjacency provenance, rank permutations,
// canonical slots, Monte-Carlo draw counts).
//
// Usage: hera_oracle [options] <rr.paf> <cr.paf> <dump.bin> [<reads.fa> <contigs.fa> <out.fa>]
//   options of ref_dump, plus  --rng mt19937|philox  --seed N  --write-draws FILE
//   --stage graph|paths|all (stop early)   --timing (print stage wall times as JSON on stdout)
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>

#include "hera_dump.hpp"
#include "hera_oracle.hpp"

using namespace hera_oracle;

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static void dump_paths(heradump::Writer &w, const std::string &prefix, const PathSet &ps) {
    w.put(prefix + "off", ps.off);
    w.put(prefix + "nodes", ps.nodes);
    w.put(prefix + "slots", ps.slots);
    w.put(prefix + "len", ps.len);
}

int main(int argc, char **argv) {
    FilterParams fp;
    PathParams pp;
    uint32_t test_load_num = 0;
    std::vector<int> metrics = {EXTENSION_SCORE, OVERLAP_SCORE};
    RngMode rng = RNG_PHILOX;
    uint64_t seed = 0;
    const char *draws_out = nullptr;
    std::string stage = "all";
    bool timing = false;
    std::vector<const char *> pos;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char * { if (i + 1 >= argc) { std::cerr << "missing value for " << a << "\n"; std::exit(2); } return argv[++i]; };
        if (a == "--mode") {
            std::string m = next();
            fp.mode = m == "min" ? MODE_MIN : m == "max" ? MODE_MAX : m == "sum" ? MODE_SUM : MODE_AVG;
        } else if (a == "--min-oll") fp.min_overlap_length = std::stoul(next());
        else if (a == "--min-olp") fp.min_overlap_percentage = std::stof(next());
        else if (a == "--max-ohl") fp.max_overhang_length = std::stoul(next());
        else if (a == "--max-ohp") fp.max_overhang_percentage = std::stof(next());
        else if (a == "--rb-att") pp.rebuild_attempts = std::stoi(next());
        else if (a == "--bt-att") pp.backtrack_attempts = std::stoi(next());
        else if (a == "--len-thr") pp.len_threshold = std::stol(next());
        else if (a == "--w-size") pp.window_size = std::stoul(next());
        else if (a == "--r-thr") pp.ratio_threshold = std::stof(next());
        else if (a == "--test-load-num") test_load_num = static_cast<uint32_t>(std::stoul(next()));
        else if (a == "--rng") rng = std::string(next()) == "mt19937" ? RNG_MT19937 : RNG_PHILOX;
        else if (a == "--seed") seed = std::strtoull(next(), nullptr, 10);
        else if (a == "--write-draws") draws_out = next();
        else if (a == "--draws") next();  // accepted for command-line symmetry with ref_dump_replay
        else if (a == "--stage") stage = next();
        else if (a == "--timing") timing = true;
        else if (a == "--metrics") {
            metrics.clear();
            std::stringstream ss(next());
            std::string m;
            while (std::getline(ss, m, ',')) metrics.push_back(m == "es" ? EXTENSION_SCORE : m == "os" ? OVERLAP_SCORE : m == "es_sqrt" ? EXTENSION_SCORE_SQRT : OVERLAP_SCORE_SQRT);
        } else pos.push_back(argv[i]);
    }
    if (pos.size() != 3 && pos.size() != 6) {
        std::cerr << "usage: hera_oracle [options] rr.paf cr.paf dump.bin [reads.fa contigs.fa out.fa]\n";
        return 2;
    }
    heradump::Writer w(pos[2]);
    if (!w.ok()) { std::cerr << "cannot write " << pos[2] << "\n"; return 2; }

    double t0 = now_s();
    GraphBuilder gb(fp, test_load_num);
    for (int f = 0; f < 2; f++) {
        if (!gb.load_file(pos[f], f == 1)) { std::cerr << "Cannot open file: " << pos[f] << std::endl; return 1; }
    }
    Graph g = gb.finalize();
    double t_graph = now_s() - t0;

    w.put("load_ctr", g.load_ctr);
    w.put("load_cum_edges", g.load_cum_edges);
    w.text("graph_stats", graph_stats(g));
    std::vector<uint64_t> name_off(1, 0);
    std::string names;
    for (const std::string &n : g.node_name) { names += n; name_off.push_back(names.size()); }
    w.put("node_len", g.node_len);
    w.put("node_anchor", g.node_anchor);
    w.put("node_name_off", name_off);
    w.text("node_names", names);
    w.put("e_q", g.e_q); w.put("e_t", g.e_t); w.put("e_qs", g.e_qs); w.put("e_qe", g.e_qe);
    w.put("e_ts", g.e_ts); w.put("e_te", g.e_te); w.put("e_os", g.e_os); w.put("e_si", g.e_si);
    w.put("e_es", g.e_es); w.put("e_strand", g.e_strand);
    w.put("adj_off", g.adj_off);
    {
        const size_t S = g.slots();
        std::vector<uint32_t> q(S), t(S), qs(S), qe(S), ts(S), te(S);
        std::vector<float> os(S), si(S), es(S);
        std::vector<uint8_t> strand(S);
        for (size_t s = 0; s < S; s++) {
            q[s] = g.slot_neighbour(s); t[s] = g.slot_owner(s);
            qs[s] = g.slot_qs(s); qe[s] = g.slot_qe(s); ts[s] = g.slot_ts(s); te[s] = g.slot_te(s);
            uint32_t e = g.adj_rec[s];
            os[s] = g.e_os[e]; si[s] = g.e_si[e]; es[s] = g.e_es[e]; strand[s] = g.e_strand[e];
        }
        w.put("a_q", q); w.put("a_t", t); w.put("a_qs", qs); w.put("a_qe", qe); w.put("a_ts", ts); w.put("a_te", te);
        w.put("a_os", os); w.put("a_si", si); w.put("a_es", es); w.put("a_strand", strand);
    }
    w.put("adj_rec", g.adj_rec);
    w.put("adj_side", g.adj_side);
    for (int m : metrics) w.put("rank_" + std::to_string(m), rank_permutation(g, m));
    w.put("canon", canonical_slots(g));
    if (stage == "graph") {
        if (timing) std::printf("{\"graph_s\": %.6f, \"records\": %llu}\n", t_graph, static_cast<unsigned long long>(g.load_ctr[0] + g.load_ctr[1]));
        return 0;
    }

    double t1 = now_s();
    PathSet all;
    std::vector<uint64_t> found;
    std::vector<uint32_t> call_kind, call_metric;
    McStats mc;
    uint64_t walks = 0;
    for (int m : metrics) {
        size_t before = mc.draws_per_walk.size();
        found.push_back(walk_monte_carlo(g, m, pp, rng, seed, all, &mc));
        walks += mc.draws_per_walk.size() - before;
        call_kind.push_back(0); call_metric.push_back(static_cast<uint32_t>(m));
    }
    for (int m : metrics) {
        found.push_back(walk_deterministic(g, m, all));
        call_kind.push_back(1); call_metric.push_back(static_cast<uint32_t>(m));
    }
    PathSet uniq = filter_unique(all);
    double t_paths = now_s() - t1;
    w.put("call_kind", call_kind);
    w.put("call_metric", call_metric);
    w.put("call_found", found);
    w.put("mc_draws_per_walk", mc.draws_per_walk);
    if (draws_out) {
        std::ofstream f(draws_out, std::ios::binary);
        f.write(reinterpret_cast<const char *>(mc.draws.data()), static_cast<std::streamsize>(mc.draws.size() * sizeof(double)));
    }
    dump_paths(w, "p_", all);
    dump_paths(w, "u_", uniq);
    w.text("path_stats", path_stats(uniq));
    if (stage == "paths") {
        if (timing) std::printf("{\"graph_s\": %.6f, \"paths_s\": %.6f, \"mc_walks\": %llu}\n", t_graph, t_paths, static_cast<unsigned long long>(walks));
        return 0;
    }

    double t2 = now_s();
    Grouping gr = group_paths(uniq, pp);
    double t_groups = now_s() - t2;
    w.put("pair_a1", gr.pair_a1); w.put("pair_a2", gr.pair_a2); w.put("pair_off", gr.pair_off);
    w.put("pair_sorted", gr.pair_sorted);
    w.put("grp_pair", gr.grp_pair); w.put("grp_size0", gr.grp_size0);
    w.put("grp_pig_off", gr.grp_pig_off); w.put("grp_pig", gr.grp_pig);
    w.put("grp_consensus", gr.grp_consensus); w.put("grp_valid", gr.grp_valid);
    w.put("pair_consensus", gr.pair_consensus);

    bool any_consensus = false;
    for (int64_t c : gr.pair_consensus) any_consensus |= c >= 0;
    Scaffold sc;
    if (any_consensus) sc = chain_scaffold(uniq, gr, 10);
    w.scalar<uint8_t>("have_scaffold", sc.ok ? 1 : 0);
    w.put("scaffold_nodes", sc.nodes);
    w.put("scaffold_slots", sc.slots);
    uint8_t fasta_ok = 0;
    if (sc.ok && pos.size() == 6) {
        std::vector<std::string> names_v, seqs;
        if (!load_fasta(pos[3], names_v, seqs) || !load_fasta(pos[4], names_v, seqs)) return 1;
        std::string fasta, log_line;
        if (stitch_scaffold(g, sc, names_v, seqs, fasta, log_line)) {
            std::ofstream f(pos[5], std::ios::binary);
            f << fasta;
            fasta_ok = 1;
        } else {
            std::remove(pos[5]);
        }
    }
    w.scalar<uint8_t>("fasta_ok", fasta_ok);
    if (timing) {
        std::printf("{\"graph_s\": %.6f, \"paths_s\": %.6f, \"groups_s\": %.6f, \"mc_walks\": %llu, \"records\": %llu}\n", t_graph, t_paths, t_groups,
                    static_cast<unsigned long long>(walks), static_cast<unsigned long long>(g.load_ctr[0] + g.load_ctr[1]));
    }
    return 0;
}
