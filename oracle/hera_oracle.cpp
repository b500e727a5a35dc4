// TEST INFRASTRUCTURE -- see hera_oracle.hpp.  CPU restatement of the reference hot path on
// flat arrays (kept-overlap table + CSR adjacency) instead of the reference's pointer graph.
// Every function cites the reference lines whose behaviour it restates.
#include "hera_oracle.hpp"

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <thread>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstring>
#include <fstream>
#include <map>
#include <random>
#include <set>
#include <sstream>
#include <unordered_map>

#include "philox.hpp"

namespace hera_oracle {

// ------------------------------------------------------------------------------------------
// PAF reader.  src/OverlapGraph.cpp:109-128 reads with formatted stream extraction, so the
// input is a character stream, not lines: string = maximal run of non-space characters,
// int = optional sign + decimal digits (stops at the first other character, fails when there
// is no digit or the value leaves int range), char = one non-space character.  Twelve typed
// fields, then four tokens are skipped.
// ------------------------------------------------------------------------------------------
namespace {

inline bool is_space(unsigned char c) { return c == ' ' || (c >= '\t' && c <= '\r'); }

struct Cursor {
    const std::string &s;
    size_t p = 0;
    bool good = true;
    explicit Cursor(const std::string &text) : s(text) {}

    void skip_space() { while (p < s.size() && is_space(static_cast<unsigned char>(s[p]))) p++; }

    bool word(std::string &out) {
        if (!good) return false;
        skip_space();
        if (p >= s.size()) return good = false;
        size_t b = p;
        while (p < s.size() && !is_space(static_cast<unsigned char>(s[p]))) p++;
        out.assign(s, b, p - b);
        return true;
    }

    bool integer(int32_t &out) {
        if (!good) return false;
        skip_space();
        if (p >= s.size()) return good = false;
        bool neg = false;
        if (s[p] == '+' || s[p] == '-') { neg = s[p] == '-'; p++; }
        int64_t v = 0;
        size_t digits = 0;
        bool overflow = false;
        while (p < s.size() && s[p] >= '0' && s[p] <= '9') {
            if (!overflow) {
                v = v * 10 + (s[p] - '0');
                if (v > (neg ? 2147483648LL : 2147483647LL)) overflow = true;
            }
            digits++;
            p++;
        }
        if (digits == 0 || overflow) return good = false;
        out = static_cast<int32_t>(neg ? -v : v);
        return true;
    }

    bool single(char &out) {
        if (!good) return false;
        skip_space();
        if (p >= s.size()) return good = false;
        out = s[p++];
        return true;
    }
};

}  // namespace

// The read loop of src/OverlapGraph.cpp:105-148; `sink` sees every record in file order (so a caller can filter on
// the fly instead of holding all records of a 42 GB file).
template <typename Sink>
static uint64_t read_records_each(const std::string &text, uint64_t stop_after, Sink sink) {
    Cursor c(text);
    uint64_t ctr = 0;
    std::string skip;
    for (;;) {
        PafRecord r;
        bool ok = c.word(r.qname) && c.integer(r.qlen) && c.integer(r.qs) && c.integer(r.qe) && c.single(r.strand) &&
                  c.word(r.tname) && c.integer(r.tlen) && c.integer(r.ts) && c.integer(r.te) && c.integer(r.matches) &&
                  c.integer(r.block) && c.integer(r.mapq);
        if (!ok) break;
        for (int k = 0; k < 4; k++) c.word(skip);  // src/OverlapGraph.cpp:124-127
        ctr++;
        sink(std::move(r));
        if (stop_after > 0 && stop_after < ctr) break;  // test_load_num_, src/OverlapGraph.cpp:145-147
    }
    return ctr;
}

uint64_t read_records(const std::string &text, uint64_t stop_after, std::vector<PafRecord> &out) {
    return read_records_each(text, stop_after, [&](PafRecord &&r) { out.push_back(std::move(r)); });
}

// src/OverlapGraph.cpp:156-202.  All comparisons are fp32; the integer thresholds are
// converted to float first.
bool filter_record(const PafRecord &r, const FilterParams &fp) {
    if (r.qname == r.tname) return false;
    const int32_t q_ovl = r.qe - r.qs, t_ovl = r.te - r.ts;     // :297-303
    const int32_t q_ohg = r.qlen - r.qe, t_ohg = r.ts;          // :305-311
    float ovl, ohg, tot;
    switch (fp.mode) {
        case MODE_MIN:
            ovl = static_cast<float>(std::min(q_ovl, t_ovl));
            ohg = static_cast<float>(std::min(q_ohg, t_ohg));
            tot = static_cast<float>(std::min(r.qlen, r.tlen));
            break;
        case MODE_MAX:
            ovl = static_cast<float>(std::max(q_ovl, t_ovl));
            ohg = static_cast<float>(std::max(q_ohg, t_ohg));
            tot = static_cast<float>(std::max(r.qlen, r.tlen));
            break;
        case MODE_AVG:
            ovl = static_cast<float>(q_ovl + t_ovl) / 2.0f;
            ohg = static_cast<float>(q_ohg + t_ohg) / 2.0f;
            tot = static_cast<float>(r.qlen + r.tlen) / 2.0f;
            break;
        case MODE_SUM:
            ovl = static_cast<float>(q_ovl + t_ovl);
            ohg = static_cast<float>(q_ohg + t_ohg);
            tot = static_cast<float>(r.qlen + r.tlen);
            break;
        default:
            return false;
    }
    const float min_len = static_cast<float>(fp.min_overlap_length);
    const float max_ohl = static_cast<float>(fp.max_overhang_length);
    if (ovl < min_len || (ovl / tot) < fp.min_overlap_percentage) return false;
    return ohg <= max_ohl && (ohg / ovl) <= fp.max_overhang_percentage;
}

// src/OverlapGraph.cpp:221-231, 321-324.  Left to right, every operation rounded to fp32.
void score_record(const PafRecord &r, float &si, float &os, float &es) {
    const int32_t q_ovl = r.qe - r.qs, t_ovl = r.te - r.ts;
    const int32_t q_ohg = r.qlen - r.qe, t_ohg = r.ts;
    si = static_cast<float>(r.matches) / static_cast<float>(std::min(r.qlen, r.tlen));
    const float half_ovl = static_cast<float>(q_ovl + t_ovl) / 2.0f;
    os = half_ovl * si;
    const float ext = static_cast<float>(r.qs) / 2.0f;
    const float ohg = static_cast<float>(q_ohg + t_ohg) / 2.0f;
    const float a = os + ext;
    es = std::fabs(a - ohg);
}

static bool starts_with_ctg(const std::string &name) {  // include/Utils.hpp:17-27 with "ctg"
    return name.size() >= 3 && std::tolower(static_cast<unsigned char>(name[0])) == 'c' &&
           std::tolower(static_cast<unsigned char>(name[1])) == 't' && std::tolower(static_cast<unsigned char>(name[2])) == 'g';
}

void GraphBuilder::load_text(const std::string &text, bool anchors) {
    uint64_t ctr = read_records_each(text, test_load_num_, [&](PafRecord &&r) {
        if (filter_record(r, fp_)) kept_.push_back(Kept{std::move(r), anchors});
    });
    ctr_.push_back(ctr);
    cum_.push_back(kept_.size());
}

bool GraphBuilder::load_file(const char *path, bool anchors) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) return false;
    std::string text(static_cast<size_t>(f.tellg()), '\0');   // one copy of the file, not two
    f.seekg(0);
    f.read(text.data(), static_cast<std::streamsize>(text.size()));
    load_text(text, anchors);
    return true;
}

// Node numbering = first appearance among kept records, query before target
// (src/OverlapGraph.cpp:208-218); adjacency = two directed copies per kept record
// (src/OverlapGraph.cpp:241-248), each node's list in record order.
Graph GraphBuilder::finalize() {
    Graph g;
    g.load_ctr = ctr_;
    g.load_cum_edges = cum_;
    std::unordered_map<std::string, uint32_t> ids;
    auto intern = [&](const std::string &name, int32_t len, bool anchor) -> uint32_t {
        auto it = ids.find(name);
        if (it != ids.end()) return it->second;
        uint32_t id = static_cast<uint32_t>(g.node_len.size());
        ids.emplace(name, id);
        g.node_len.push_back(static_cast<uint32_t>(len));
        g.node_anchor.push_back(anchor ? 1 : 0);
        g.node_name.push_back(name);
        return id;
    };
    for (const Kept &k : kept_) {
        const PafRecord &r = k.r;
        // ContigPosition, src/OverlapGraph.cpp:130-138
        const bool contig_is_query = k.is_cr && starts_with_ctg(r.qname);
        const bool contig_is_target = k.is_cr && !contig_is_query;
        uint32_t q = intern(r.qname, r.qlen, contig_is_query);
        uint32_t t = intern(r.tname, r.tlen, contig_is_target);
        float si, os, es;
        score_record(r, si, os, es);
        g.e_q.push_back(q); g.e_t.push_back(t);
        g.e_qs.push_back(static_cast<uint32_t>(r.qs)); g.e_qe.push_back(static_cast<uint32_t>(r.qe));
        g.e_ts.push_back(static_cast<uint32_t>(r.ts)); g.e_te.push_back(static_cast<uint32_t>(r.te));
        g.e_os.push_back(os); g.e_si.push_back(si); g.e_es.push_back(es);
        g.e_strand.push_back(r.strand == '-' ? 1 : 0);
    }
    const size_t N = g.nodes(), E = g.edges();
    g.adj_off.assign(N + 1, 0);
    for (size_t e = 0; e < E; e++) { g.adj_off[g.e_t[e] + 1]++; g.adj_off[g.e_q[e] + 1]++; }
    for (size_t i = 0; i < N; i++) g.adj_off[i + 1] += g.adj_off[i];
    g.adj_rec.resize(2 * E);
    g.adj_side.resize(2 * E);
    std::vector<uint64_t> cur(g.adj_off.begin(), g.adj_off.end() - 1);
    for (size_t e = 0; e < E; e++) {
        uint64_t st = cur[g.e_t[e]]++;
        g.adj_rec[st] = static_cast<uint32_t>(e); g.adj_side[st] = 0;
        uint64_t sq = cur[g.e_q[e]]++;
        g.adj_rec[sq] = static_cast<uint32_t>(e); g.adj_side[sq] = 1;
    }
    return g;
}

float Graph::slot_metric(uint64_t s, int metric) const {  // include/Utils.hpp:47-60
    const uint32_t e = adj_rec[s];
    switch (metric) {
        case EXTENSION_SCORE: return e_es[e];
        case EXTENSION_SCORE_SQRT: return std::sqrt(e_es[e]);
        case OVERLAP_SCORE: return e_os[e];
        case OVERLAP_SCORE_SQRT: return std::sqrt(e_os[e]);
        default: return 0.0f;
    }
}

std::vector<uint32_t> rank_permutation(const Graph &g, int metric) {
    std::vector<uint32_t> rank(g.slots());
    for (size_t i = 0; i < rank.size(); i++) rank[i] = static_cast<uint32_t>(i);
    for (size_t x = 0; x < g.nodes(); x++) {
        std::stable_sort(rank.begin() + static_cast<ptrdiff_t>(g.adj_off[x]), rank.begin() + static_cast<ptrdiff_t>(g.adj_off[x + 1]),
                         [&](uint32_t a, uint32_t b) { return g.slot_metric(a, metric) > g.slot_metric(b, metric); });
    }
    return rank;
}

std::vector<uint32_t> canonical_slots(const Graph &g) {
    std::vector<uint32_t> canon(g.slots());
    for (size_t x = 0; x < g.nodes(); x++) {
        std::unordered_map<uint32_t, uint32_t> first;
        for (uint64_t s = g.adj_off[x]; s < g.adj_off[x + 1]; s++) {
            auto it = first.emplace(g.slot_neighbour(s), static_cast<uint32_t>(s)).first;
            canon[s] = it->second;
        }
    }
    return canon;
}

// ------------------------------------------------------------------------------------------
// Paths
// ------------------------------------------------------------------------------------------
void PathSet::append(const std::vector<uint32_t> &n, const std::vector<uint32_t> &s, int64_t l) {
    nodes.insert(nodes.end(), n.begin(), n.end());
    slots.insert(slots.end(), s.begin(), s.end());
    off.push_back(nodes.size());
    len.push_back(l);
}

bool PathSet::same_nodes(size_t a, size_t b) const {  // src/Path.cpp:46-49
    uint64_t na = off[a + 1] - off[a], nb = off[b + 1] - off[b];
    return na == nb && std::equal(nodes.begin() + static_cast<ptrdiff_t>(off[a]), nodes.begin() + static_cast<ptrdiff_t>(off[a + 1]),
                                  nodes.begin() + static_cast<ptrdiff_t>(off[b]));
}

// src/Path.cpp:6-16: first node's length plus the (neighbour start - own start) of every edge
// except the last one.
int64_t path_length(const Graph &g, const std::vector<uint32_t> &nodes, const std::vector<uint32_t> &slots) {
    if (nodes.empty()) return 0;
    int64_t len = g.node_len[nodes[0]];
    for (size_t i = 0; i + 1 < slots.size(); i++) len += g.slot_delta(slots[i]);
    return len;
}

// ---- running the anchors of a build call on several host threads (test infrastructure for the full-size configurations:
// HERA_ORACLE_THREADS, default 1).  The walks of different anchors are independent in the reference except for the shared
// std::mt19937 stream, so the mt19937 mode always runs on one thread; results are appended in anchor order either way.
namespace {
struct AnchorOut {
    PathSet ps;
    McStats st;
    uint64_t found = 0;
};

int oracle_threads() {
    const char *v = std::getenv("HERA_ORACLE_THREADS");
    const int n = v ? std::atoi(v) : 1;
    return n < 1 ? 1 : n;
}

void append_set(PathSet &dst, const PathSet &src) {
    const uint64_t base = dst.nodes.size();
    dst.nodes.insert(dst.nodes.end(), src.nodes.begin(), src.nodes.end());
    dst.slots.insert(dst.slots.end(), src.slots.begin(), src.slots.end());
    for (size_t i = 1; i < src.off.size(); i++) dst.off.push_back(base + src.off[i]);
    dst.len.insert(dst.len.end(), src.len.begin(), src.len.end());
}

template <typename Body>
uint64_t for_each_anchor(const Graph &g, int threads, PathSet &ps, McStats *stats, Body body) {
    std::vector<uint32_t> anchors;
    for (uint32_t x = 0; x < g.nodes(); x++) if (g.node_anchor[x]) anchors.push_back(x);
    std::vector<AnchorOut> out(anchors.size());
    std::atomic<size_t> next{0};
    auto work = [&]() {
        std::vector<uint8_t> visited(g.nodes(), 0);
        for (size_t i = next.fetch_add(1); i < anchors.size(); i = next.fetch_add(1)) body(anchors[i], out[i], visited);
    };
    if (threads <= 1) {
        work();
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; t++) pool.emplace_back(work);
        for (auto &t : pool) t.join();
    }
    uint64_t found = 0;
    for (AnchorOut &o : out) {
        append_set(ps, o.ps);
        found += o.found;
        if (stats) {
            stats->draws_per_walk.insert(stats->draws_per_walk.end(), o.st.draws_per_walk.begin(), o.st.draws_per_walk.end());
            stats->draws.insert(stats->draws.end(), o.st.draws.begin(), o.st.draws.end());
        }
    }
    return found;
}
}  // namespace

uint64_t walk_monte_carlo(const Graph &g, int metric, const PathParams &pp, RngMode mode, uint64_t seed, PathSet &ps, McStats *stats) {
    std::mt19937 engine;  // default seed 5489, one per call (src/PathManager.cpp:13)
    std::uniform_real_distribution<> unit(0., 1.);
    const int threads = mode == RNG_MT19937 ? 1 : oracle_threads();
    return for_each_anchor(g, threads, ps, stats, [&](uint32_t anchor, AnchorOut &o, std::vector<uint8_t> &visited) {
        std::vector<uint32_t> nodes, slots, cand, touched;
        for (int attempt = 0; attempt < pp.rebuild_attempts; attempt++) {
            for (uint32_t x : touched) visited[x] = 0;   // (the reference allocates a fresh vector<bool>(N) per walk, src/PathManager.cpp:32)
            touched.assign(1, anchor);
            visited[anchor] = 1;
            nodes.assign(1, anchor);
            slots.clear();
            bool accepted = false;
            int backtracks = 0;
            uint32_t cur = anchor, draw_no = 0;
            for (;;) {
                double sum = 0;
                int64_t forced = -1;
                cand.clear();
                for (uint64_t s = g.adj_off[cur]; s < g.adj_off[cur + 1]; s++) {  // :53-67
                    uint32_t nb = g.slot_neighbour(s);
                    if (g.node_anchor[nb] && nb != anchor) { forced = static_cast<int64_t>(s); break; }
                    if (!visited[nb]) { sum += g.slot_metric(s, metric); cand.push_back(static_cast<uint32_t>(s)); }
                }
                int64_t pick = -1;
                if (forced >= 0) {
                    pick = forced;
                } else if (cand.empty()) {  // dead end, :72-101
                    if (backtracks < pp.backtrack_attempts && slots.size() > 1) {
                        bool open = false;
                        do {
                            nodes.pop_back();
                            slots.pop_back();
                            uint32_t back = nodes.back();
                            for (uint64_t s = g.adj_off[back]; s < g.adj_off[back + 1]; s++) {
                                if (!visited[g.slot_neighbour(s)]) { open = true; break; }
                            }
                        } while (!slots.empty() && !open);
                        if (slots.empty()) break;
                        backtracks++;
                        cur = nodes.back();
                        continue;
                    }
                    break;
                } else {  // roulette, :102-115
                    double u = mode == RNG_MT19937 ? unit(engine) : mc_draw(seed, static_cast<uint32_t>(metric), anchor, static_cast<uint32_t>(attempt), draw_no);
                    draw_no++;
                    if (stats) o.st.draws.push_back(u);
                    double target = u * sum;
                    double acc = 0;
                    for (uint32_t s : cand) {
                        acc += g.slot_metric(s, metric);
                        if (acc >= target) { pick = s; break; }
                    }
                    if (pick < 0) break;  // reference would dereference null here (NaN metrics only)
                }
                cur = g.slot_neighbour(static_cast<uint64_t>(pick));
                nodes.push_back(cur);
                slots.push_back(static_cast<uint32_t>(pick));
                visited[cur] = 1;
                touched.push_back(cur);
                int64_t len = path_length(g, nodes, slots);
                if (len <= 0) break;                        // :126
                if (g.node_anchor[cur]) { accepted = true; break; }  // :131
                if (len >= pp.len_threshold) break;         // :137
            }
            if (stats) o.st.draws_per_walk.push_back(draw_no);
            if (accepted) {
                int64_t len = path_length(g, nodes, slots);
                if (len < 0) continue;
                o.ps.append(nodes, slots, len);
                o.found++;
            }
        }
        for (uint32_t x : touched) visited[x] = 0;
    });
}

uint64_t walk_deterministic(const Graph &g, int metric, PathSet &ps) {
    const std::vector<uint32_t> rank = rank_permutation(g, metric);
    const std::vector<uint32_t> canon = canonical_slots(g);
    return for_each_anchor(g, oracle_threads(), ps, nullptr, [&](uint32_t anchor, AnchorOut &o, std::vector<uint8_t> &visited) {
        std::vector<uint32_t> nodes, slots, touched;
        for (uint64_t fs = g.adj_off[anchor]; fs < g.adj_off[anchor + 1]; fs++) {  // :204
            for (uint32_t x : touched) visited[x] = 0;   // (a fresh vector<bool>(N) per walk in the reference, src/PathManager.cpp:212)
            visited[anchor] = 1;
            uint32_t cur = g.slot_neighbour(fs);
            visited[cur] = 1;
            touched = {anchor, cur};
            nodes = {anchor, cur};
            slots = {static_cast<uint32_t>(fs)};
            if (g.node_anchor[cur]) {  // :224-229 -- and no further first edges for this anchor
                o.ps.append(nodes, slots, path_length(g, nodes, slots));
                o.found++;
                break;
            }
            bool reached = false, went_back = false;
            int step = 0;
            uint64_t skip = 0;
            for (;;) {
                const uint64_t base = g.adj_off[cur], deg = g.adj_off[cur + 1] - base;
                int64_t chosen = -1;
                if (deg != 0 && skip < deg - 1) {  // :245,:267
                    for (uint64_t k = skip; k < deg - 1; k++) {  // worst-ranked edge is never tried, :302
                        uint32_t s = rank[base + k];
                        if (!visited[g.slot_neighbour(s)]) { chosen = s; break; }
                    }
                }
                if (chosen < 0) {  // dead end: one step back, once per walk (:246-264, :268-285, :318-336)
                    if (step == 0 || went_back) break;
                    went_back = true;
                    step--;
                    skip++;  // never reset afterwards
                    nodes.pop_back();
                    slots.pop_back();
                    cur = nodes.back();
                    visited[cur] = 0;
                    continue;
                }
                slots.push_back(canon[static_cast<size_t>(chosen)]);  // first parallel edge, :340-347
                visited[cur] = 1;
                touched.push_back(cur);
                cur = g.slot_neighbour(static_cast<uint64_t>(chosen));
                nodes.push_back(cur);
                if (g.node_anchor[cur]) { reached = true; break; }
                step++;
            }
            if (reached) {
                int64_t len = path_length(g, nodes, slots);
                if (len < 0) continue;
                o.ps.append(nodes, slots, len);
                o.found++;
            }
        }
        for (uint32_t x : touched) visited[x] = 0;
    });
}

PathSet filter_unique(const PathSet &ps) {
    PathSet out;
    std::vector<uint32_t> n, s;
    size_t last_kept = 0;
    for (size_t i = 0; i < ps.size(); i++) {
        if (i > 0 && ps.same_nodes(last_kept, i)) continue;
        last_kept = i;
        n.assign(ps.nodes.begin() + static_cast<ptrdiff_t>(ps.off[i]), ps.nodes.begin() + static_cast<ptrdiff_t>(ps.off[i + 1]));
        // path i owns slots [off[i]-i, off[i+1]-(i+1))
        s.assign(ps.slots.begin() + static_cast<ptrdiff_t>(ps.off[i] - i), ps.slots.begin() + static_cast<ptrdiff_t>(ps.off[i + 1] - (i + 1)));
        out.append(n, s, ps.len[i]);
    }
    return out;
}

// ------------------------------------------------------------------------------------------
// Windows and groups
// ------------------------------------------------------------------------------------------
namespace {

struct Window {  // src/PathWindow.cpp:6-39
    std::map<uint64_t, int> freq;
    int total = 0;
    std::pair<uint64_t, int> lowest() const {
        auto best = freq.begin();
        for (auto it = freq.begin(); it != freq.end(); ++it) if (it->second < best->second) best = it;
        return *best;
    }
    std::pair<uint64_t, int> highest() const {
        auto best = freq.begin();
        for (auto it = freq.begin(); it != freq.end(); ++it) if (best->second < it->second) best = it;
        return *best;
    }
};

// src/PathManager.cpp:508-624.  v = path ids sorted by length; returns group sizes (consecutive ranges).
std::vector<uint64_t> split_into_groups(const std::vector<uint32_t> &v, const std::vector<int64_t> &len, const PathParams &pp) {
    const uint64_t min_len = static_cast<uint64_t>(len[v.front()]), max_len = static_cast<uint64_t>(len[v.back()]);
    if (max_len - min_len < static_cast<uint64_t>(pp.len_threshold)) return {v.size()};
    std::vector<Window> wins;
    for (uint64_t lower = min_len, upper = lower + pp.window_size; lower <= max_len; lower = upper, upper += pp.window_size) {
        Window w;
        for (uint32_t id : v) {
            uint64_t l = static_cast<uint64_t>(len[id]);
            if (l >= upper) break;
            if (l >= lower) { w.freq[l]++; w.total++; }
        }
        if (w.total != 0) wins.push_back(std::move(w));
    }
    std::set<uint64_t> borders;
    for (size_t i = 1; i + 1 < wins.size(); i++) {
        const Window *trio[3] = {&wins[i - 1], &wins[i], &wins[i + 1]};
        const Window *valley = trio[0], *peak = trio[0];
        for (int k = 1; k < 3; k++) {
            if (trio[k]->total < valley->total) valley = trio[k];  // first minimum
            if (peak->total < trio[k]->total) peak = trio[k];      // first maximum
        }
        std::pair<uint64_t, int> lo = valley->lowest(), hi = peak->highest();
        if (static_cast<float>(lo.second) < pp.ratio_threshold * static_cast<float>(hi.second)) borders.insert(lo.first);
    }
    if (borders.empty()) return {v.size()};
    std::vector<uint64_t> sizes;
    size_t begin = 0;
    for (uint64_t border : borders) {
        size_t end = begin + 1;
        while (end < v.size() && static_cast<uint64_t>(len[v[end]]) < border) end++;  // reference scans unbounded (:570)
        sizes.push_back(end - begin);
        begin = end;
    }
    sizes.push_back(v.size() - begin);
    return sizes;
}

}  // namespace

Grouping group_paths(const PathSet &ps, const PathParams &pp) {
    Grouping out;
    // src/PathManager.cpp:626-645: ordered map keyed by (first node, last node); node pointers
    // order like node indices because nodes_ is one contiguous vector.
    std::map<std::pair<uint32_t, uint32_t>, std::vector<uint32_t>> by_pair;
    for (size_t i = 0; i < ps.size(); i++) {
        by_pair[{ps.nodes[ps.off[i]], ps.nodes[ps.off[i + 1] - 1]}].push_back(static_cast<uint32_t>(i));
    }
    uint32_t pair_idx = 0;
    for (auto &kv : by_pair) {
        std::vector<uint32_t> &v = kv.second;
        out.pair_a1.push_back(kv.first.first);
        out.pair_a2.push_back(kv.first.second);
        libstdcxx_sort(v.data(), v.data() + v.size(), [&](uint32_t a, uint32_t b) { return ps.len[a] < ps.len[b]; });  // :510-513
        out.pair_sorted.insert(out.pair_sorted.end(), v.begin(), v.end());
        out.pair_off.push_back(out.pair_sorted.size());

        struct GroupResult { int64_t consensus; int32_t valid; };
        std::vector<GroupResult> with_consensus;
        size_t begin = 0;
        for (uint64_t size0 : split_into_groups(v, ps.len, pp)) {
            std::vector<uint32_t> pig(v.begin() + static_cast<ptrdiff_t>(begin), v.begin() + static_cast<ptrdiff_t>(begin + size0));
            begin += size0;
            out.grp_pair.push_back(pair_idx);
            out.grp_size0.push_back(static_cast<uint32_t>(size0));
            // discardNotFrequent, src/PathGroup.cpp:40-55,113-125
            std::map<uint64_t, int> freq;
            for (uint32_t id : pig) freq[static_cast<uint64_t>(ps.len[id])]++;
            int top = 0;
            for (auto &f : freq) top = std::max(top, f.second);
            int thr = top / 2;
            if (thr != 0) {
                std::vector<uint32_t> keep;
                for (uint32_t id : pig) if (freq[static_cast<uint64_t>(ps.len[id])] >= thr) keep.push_back(id);
                pig.swap(keep);
            }
            // calculateConsensusPath, src/PathGroup.cpp:58-79
            int64_t consensus = -1;
            uint64_t span = static_cast<uint64_t>(ps.len[pig.back()] - ps.len[pig.front()]);
            if (!(span > 100000ull)) {  // include/PathGroup.hpp:55
                int32_t acc = 0;  // std::accumulate with an int init value: the running sum is cut to 32 bits every step
                for (uint32_t id : pig) acc = static_cast<int32_t>(static_cast<uint64_t>(static_cast<int64_t>(acc)) + static_cast<uint64_t>(ps.len[id]));
                uint64_t avg = static_cast<uint64_t>(static_cast<int64_t>(acc)) / pig.size();
                consensus = pig.back();
                for (uint32_t id : pig) if (static_cast<uint64_t>(ps.len[id]) >= avg) { consensus = id; break; }
            }
            // calculateValidPathNumber, src/PathGroup.cpp:83-93
            int32_t valid = 0;
            if (consensus >= 0) for (uint32_t id : pig) if (ps.same_nodes(static_cast<size_t>(consensus), id)) valid++;
            out.grp_pig.insert(out.grp_pig.end(), pig.begin(), pig.end());
            out.grp_pig_off.push_back(out.grp_pig.size());
            out.grp_consensus.push_back(consensus);
            out.grp_valid.push_back(valid);
            if (consensus >= 0) with_consensus.push_back({consensus, valid});
        }
        // inter-group pick, src/Main.cpp:306-332
        int64_t chosen = -1;
        if (with_consensus.size() == 1) {
            chosen = with_consensus[0].consensus;
        } else if (with_consensus.size() >= 2) {
            libstdcxx_sort(with_consensus.data(), with_consensus.data() + with_consensus.size(),
                           [&](const GroupResult &a, const GroupResult &b) { return ps.len[a.consensus] > ps.len[b.consensus]; });
            size_t keep = 0;
            if (with_consensus.size() > 2) {
                for (size_t i = 1; i < with_consensus.size(); i++) {
                    if (2 * with_consensus[keep].valid <= with_consensus[i].valid) keep = i;
                }
            }
            chosen = with_consensus[keep].consensus;
        }
        out.pair_consensus.push_back(chosen);
        pair_idx++;
    }
    return out;
}

// src/PathManager.cpp:664-767.
Scaffold chain_scaffold(const PathSet &ps, const Grouping &gr, uint64_t min_path_num) {
    Scaffold sc;
    using Key = std::pair<uint32_t, uint32_t>;
    std::map<Key, std::pair<int64_t, uint64_t>> pool;  // pair -> (consensus path, number of paths)
    std::vector<uint32_t> anchors;
    uint64_t best_num = 0;
    Key best_key{UINT32_MAX, UINT32_MAX};
    for (size_t p = 0; p < gr.pair_a1.size(); p++) {
        if (gr.pair_consensus[p] < 0) continue;
        uint64_t num = gr.pair_off[p + 1] - gr.pair_off[p];
        if (num < min_path_num) continue;
        for (uint32_t a : {gr.pair_a1[p], gr.pair_a2[p]}) {
            if (std::find(anchors.begin(), anchors.end(), a) == anchors.end()) anchors.push_back(a);
        }
        Key k{gr.pair_a1[p], gr.pair_a2[p]};
        pool[k] = {gr.pair_consensus[p], num};
        if (best_num < num) { best_num = num; best_key = k; }
    }
    if (!pool.count(best_key)) return sc;  // reference: map::at throws here (no pair with enough paths)

    auto path_nodes = [&](int64_t id, size_t i) { return ps.nodes[ps.off[static_cast<size_t>(id)] + i]; };
    auto path_slots = [&](int64_t id, size_t i) { return ps.slots[ps.off[static_cast<size_t>(id)] - static_cast<size_t>(id) + i]; };
    auto path_edges = [&](int64_t id) { return static_cast<size_t>(ps.off[static_cast<size_t>(id) + 1] - ps.off[static_cast<size_t>(id)] - 1); };

    int64_t seed_path = pool[best_key].first;
    for (size_t i = 0; i <= path_edges(seed_path); i++) sc.nodes.push_back(path_nodes(seed_path, i));
    for (size_t i = 0; i < path_edges(seed_path); i++) sc.slots.push_back(path_slots(seed_path, i));
    uint32_t left = best_key.first, right = best_key.second;
    const std::vector<uint32_t> seen = {left, right};  // never grows in the reference (:708-710)
    while (!pool.empty()) {
        bool grow_left = false, any = false;
        uint64_t top = 0;
        Key pick = best_key;
        for (uint32_t n : anchors) {
            if (std::find(seen.begin(), seen.end(), n) != seen.end()) continue;
            Key lk{n, left};  // the cursors move inside the scan (:722-738)
            auto it = pool.find(lk);
            if (it != pool.end() && top < it->second.second) { top = it->second.second; pick = lk; grow_left = true; any = true; left = n; }
            Key rk{right, n};
            it = pool.find(rk);
            if (it != pool.end() && top < it->second.second) { top = it->second.second; pick = rk; grow_left = false; any = true; right = n; }
        }
        if (!any) break;
        int64_t id = pool[pick].first;
        size_t ne = path_edges(id);
        if (!grow_left) {
            for (size_t i = 0; i < ne; i++) { sc.nodes.push_back(path_nodes(id, i + 1)); sc.slots.push_back(path_slots(id, i)); }
        } else {
            std::vector<uint32_t> pn, pslots;
            for (size_t i = 0; i < ne; i++) { pn.push_back(path_nodes(id, i)); pslots.push_back(path_slots(id, i)); }
            sc.nodes.insert(sc.nodes.begin(), pn.begin(), pn.end());
            sc.slots.insert(sc.slots.begin(), pslots.begin(), pslots.end());
        }
        pool.erase(pick);
    }
    sc.ok = true;
    return sc;
}

// ------------------------------------------------------------------------------------------
// Scaffolder, src/Scaffolder.cpp:12-147
// ------------------------------------------------------------------------------------------
bool load_fasta(const char *path, std::vector<std::string> &names, std::vector<std::string> &seqs) {
    const size_t n = std::strlen(path);
    const bool fastq = n > 0 && path[n - 1] == 'q';
    std::ifstream f(path);
    if (!f) return false;
    std::string line;
    bool want_name = true;
    while (std::getline(f, line)) {
        if (want_name) {
            std::istringstream iss(line);
            iss >> line;
            names.emplace_back(line.empty() ? line : line.substr(1));
        } else {
            seqs.push_back(line);
            if (fastq) { std::getline(f, line); std::getline(f, line); }
        }
        want_name = !want_name;
    }
    return true;
}

bool stitch_scaffold(const Graph &g, const Scaffold &sc, const std::vector<std::string> &names,
                     const std::vector<std::string> &seqs, std::string &fasta_text, std::string &log_line) {
    auto seq_of = [&](uint32_t node) -> const std::string * {
        for (size_t i = 0; i < names.size(); i++) if (names[i] == g.node_name[node]) return i < seqs.size() ? &seqs[i] : nullptr;
        return nullptr;
    };
    // substring(str, start, end) = str.substr(start, end - start + 1) with unsigned arithmetic (:110-113)
    auto piece = [](const std::string &s, uint64_t start, uint64_t end, std::string &out) -> bool {
        if (start > s.size()) return false;  // reference: std::out_of_range
        out = s.substr(start, end - start + 1);
        return true;
    };
    auto rev = [](std::string s) { std::reverse(s.begin(), s.end()); return s; };
    fasta_text = ">Resulting_HERA_scaffold\n";
    if (sc.nodes.empty()) return false;
    if (sc.nodes.size() == 1) {
        const std::string *s = seq_of(sc.nodes[0]);
        if (!s) return false;
        log_line = g.node_name[sc.nodes[0]] + "\n";
        fasta_text += rev(*s);
        return true;
    }
    std::string seq, tmp;
    log_line = g.node_name[sc.nodes[0]];
    const std::string *first = seq_of(sc.nodes[0]);
    if (!first || !piece(*first, static_cast<uint64_t>(g.slot_te(sc.slots[0])) + 1, first->size(), tmp)) return false;
    seq += rev(tmp);
    for (size_t i = 1; i < sc.slots.size(); i++) {
        const uint64_t start = g.slot_te(sc.slots[i]);     // next edge, own-side end
        const uint64_t end = g.slot_qe(sc.slots[i - 1]);   // previous edge, neighbour-side end
        if (start < end) {
            const std::string *s = seq_of(sc.nodes[i]);
            if (!s) return false;
            std::string body = *s;
            if (g.e_strand[g.adj_rec[sc.slots[i - 1]]]) {
                for (char &c : body) c = c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'G' ? 'C' : c == 'C' ? 'G' : c;
            }
            log_line += "-" + g.node_name[sc.nodes[i]];
            if (!piece(body, start + 1, end, tmp)) return false;
            seq += rev(tmp);
        } else if (start > end) {
            uint64_t keep = static_cast<uint64_t>(seq.size()) - (start - end);  // wraps like the reference's ulong arithmetic
            seq = seq.substr(0, keep);
        }
    }
    log_line += "-" + g.node_name[sc.nodes.back()] + "\n";
    const std::string *last = seq_of(sc.nodes.back());
    if (!last || !piece(*last, 0, g.slot_qe(sc.slots.back()), tmp)) return false;
    seq += rev(tmp);
    std::reverse(seq.begin(), seq.end());
    fasta_text += seq + "\n";
    return true;
}

// ------------------------------------------------------------------------------------------
// Reports (text must match the reference's stdout)
// ------------------------------------------------------------------------------------------
std::string graph_stats(const Graph &g) {
    std::stringstream str;
    uint64_t min_len = UINT64_MAX, max_len = 0, min_con = UINT64_MAX, max_con = 0, anchors = 0, reads = 0;
    for (size_t i = 0; i < g.nodes(); i++) {
        g.node_anchor[i] ? anchors++ : reads++;
        uint64_t deg = g.adj_off[i + 1] - g.adj_off[i];
        min_len = std::min<uint64_t>(min_len, g.node_len[i]);
        max_len = std::max<uint64_t>(max_len, g.node_len[i]);
        min_con = std::min(min_con, deg);
        max_con = std::max(max_con, deg);
    }
    uint64_t min_ol = UINT64_MAX, max_ol = 0;
    float min_os = FLT_MAX, max_os = 0, min_es = FLT_MAX, max_es = 0, min_si = FLT_MAX, max_si = 0;
    for (size_t e = 0; e < g.edges(); e++) {
        if (min_os > g.e_os[e]) min_os = g.e_os[e];
        if (max_os < g.e_os[e]) max_os = g.e_os[e];
        if (min_es > g.e_es[e]) min_es = g.e_es[e];
        if (max_es < g.e_es[e]) max_es = g.e_es[e];
        if (min_si > g.e_si[e]) min_si = g.e_si[e];
        if (max_si < g.e_si[e]) max_si = g.e_si[e];
        int ol = static_cast<int>(std::max(g.e_qe[e] - g.e_qs[e], g.e_te[e] - g.e_ts[e]));
        uint64_t olu = static_cast<uint64_t>(static_cast<int64_t>(ol));
        if (min_ol > olu) min_ol = olu;
        if (max_ol < olu) max_ol = olu;
    }
    str << "Nodes" << '\n'
        << "-  anchor: " << anchors << '\n'
        << "-    read: " << reads << '\n'
        << "-   total: " << g.nodes() << '\n'
        << "- min_len: " << min_len << '\n'
        << "- max_len: " << max_len << '\n'
        << "- min_con: " << min_con << '\n'
        << "- max_con: " << max_con << '\n'
        << "Edges" << '\n'
        << "-   total: " << g.edges() << '\n'
        << "-  min_OL: " << min_ol << '\n'
        << "-  max_OL: " << max_ol << '\n'
        << "-  min_OS: " << min_os << '\n'
        << "-  max_OS: " << max_os << '\n'
        << "-  min_ES: " << min_es << '\n'
        << "-  max_ES: " << max_es << '\n'
        << "-  min_SI: " << min_si << '\n'
        << "-  max_SI: " << max_si << '\n'
        << std::endl;
    return str.str();
}

std::string path_stats(const PathSet &ps) {
    std::stringstream str;
    uint64_t min_len = UINT64_MAX, max_len = 0, sum = 0, negatives = 0;
    for (int64_t l : ps.len) {
        uint64_t u = static_cast<uint64_t>(l);
        sum += u;
        min_len = std::min(min_len, u);
        max_len = std::max(max_len, u);
        if (l < 0) negatives++;
    }
    str << "Paths" << '\n'
        << "- total_num: " << ps.size() << '\n'
        << "-   min_len: " << min_len << '\n'
        << "-   max_len: " << max_len << '\n'
        << "-   avg_len: " << (ps.size() > 0 ? sum / ps.size() : 0) << std::endl;
    if (negatives > 0) str << "- negatives: " << negatives << std::endl;
    return str.str();
}

}  // namespace hera_oracle
