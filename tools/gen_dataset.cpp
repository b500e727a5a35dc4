// Synthetic HERA dataset generator (run.sh dataset contract: reads.fasta, contigs.fasta,
// rr_overlaps.paf, cr_overlaps.paf in one directory).  Deterministic for a given option set
// (own splitmix64 stream; no libc rand), so the GPU box regenerates byte-identical inputs.
//
// Geometry follows SURVEY.md Appendix A, which was derived from what the reference's filter
// and path-length code reward (src/OverlapGraph.cpp:161-201, src/Path.cpp:12-15):
//   * every read pair whose genome intervals intersect by >= min_ovl gives two PAF records
//     (a as query / b as target and the dual), grouped by query like minimap2 --dual=yes;
//   * reads covering different copies of one repeat family get repeat-induced records;
//   * contig-read records put the contig in the QUERY column when the read hangs off the
//     contig's right end, else in the TARGET column; contig names start with "ctg";
//   * exactly 16 whitespace-separated columns per record (src/OverlapGraph.cpp:109-128).
//
// Usage: gen_dataset --out DIR [--genome N] [--read-len N] [--coverage X] [--contigs N]
//          [--gap N] [--min-ovl N] [--families N] [--copies N] [--seed N] [--jitter N]
//          [--minus-frac F] [--no-fasta] [--max-rr N] [--dry-run]
#include <algorithm>
#include <cinttypes>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed) {}
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    uint64_t below(uint64_t n) { return n ? next() % n : 0; }
    double unit() { return (next() >> 11) * (1.0 / 9007199254740992.0); }
};

static uint64_t mix(uint64_t a, uint64_t b, uint64_t c) {
    Rng r(a * 0x9E3779B97F4A7C15ull ^ (b + 0x632BE59BD9B4E019ull) * 0xD6E8FEB86659FD93ull ^ c);
    r.next();
    return r.next();
}

struct Read { int64_t start, len; };
struct Copy { int family; int64_t start, len; };

struct Out {
    FILE *f;
    std::vector<char> buf;
    explicit Out(const std::string &p) : f(std::fopen(p.c_str(), "wb")) { buf.reserve(1 << 22); }
    void flush() { if (!buf.empty()) { std::fwrite(buf.data(), 1, buf.size(), f); buf.clear(); } }
    void write(const char *s, size_t n) { buf.insert(buf.end(), s, s + n); if (buf.size() > (1u << 22) - 4096) flush(); }
    ~Out() { flush(); if (f) std::fclose(f); }
};

static bool g_dry_run = false;   // --dry-run: count the records, write nothing (used to size a configuration)

static void paf_record(Out &o, const char *qn, int64_t ql, int64_t qs, int64_t qe, char strand, const char *tn,
                       int64_t tl, int64_t ts, int64_t te, int64_t matches, int64_t block) {
    if (g_dry_run) return;
    char line[512];
    int n = std::snprintf(line, sizeof line,
                          "%s\t%" PRId64 "\t%" PRId64 "\t%" PRId64 "\t%c\t%s\t%" PRId64 "\t%" PRId64 "\t%" PRId64
                          "\t%" PRId64 "\t%" PRId64 "\t0\ttp:A:S\tcm:i:%" PRId64 "\ts1:i:%" PRId64 "\tdv:f:0.1000\n",
                          qn, ql, qs, qe, strand, tn, tl, ts, te, matches, block, matches / 10, matches);
    o.write(line, static_cast<size_t>(n));
}

int main(int argc, char **argv) {
    std::string out_dir;
    int64_t G = 600000, L = 10000, gap = 50000, min_ovl = 1000, jitter = 0, max_rr = 0;
    double coverage = 30.0, minus_frac = 0.0;
    int n_contigs = 3, families = 0, copies = 3;
    uint64_t seed = 7;
    bool fasta = true;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto val = [&]() -> const char * { if (i + 1 >= argc) { std::fprintf(stderr, "missing value for %s\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--out") out_dir = val();
        else if (a == "--genome") G = std::atoll(val());
        else if (a == "--read-len") L = std::atoll(val());
        else if (a == "--coverage") coverage = std::atof(val());
        else if (a == "--contigs") n_contigs = std::atoi(val());
        else if (a == "--gap") gap = std::atoll(val());
        else if (a == "--min-ovl") min_ovl = std::atoll(val());
        else if (a == "--families") families = std::atoi(val());
        else if (a == "--copies") copies = std::atoi(val());
        else if (a == "--seed") seed = std::strtoull(val(), nullptr, 10);
        else if (a == "--jitter") jitter = std::atoll(val());
        else if (a == "--minus-frac") minus_frac = std::atof(val());
        else if (a == "--max-rr") max_rr = std::atoll(val());
        else if (a == "--no-fasta") fasta = false;
        else if (a == "--dry-run") { g_dry_run = true; fasta = false; }
        else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
    }
    if (out_dir.empty()) { std::fprintf(stderr, "--out DIR is required\n"); return 2; }

    Rng rng(seed * 0x2545F4914F6CDD1Dull + 1);

    // Reads: uniform starts, length L +-10 %, sorted by start (ties by length).
    int64_t n_reads = static_cast<int64_t>(static_cast<double>(G) * coverage / static_cast<double>(L));
    std::vector<Read> reads(static_cast<size_t>(n_reads));
    for (Read &r : reads) {
        r.len = L - L / 10 + static_cast<int64_t>(rng.below(static_cast<uint64_t>(L / 5 + 1)));
        r.start = static_cast<int64_t>(rng.below(static_cast<uint64_t>(G - r.len)));
    }
    std::sort(reads.begin(), reads.end(), [](const Read &a, const Read &b) { return a.start != b.start ? a.start < b.start : a.len < b.len; });

    // Contigs: n_contigs equal stretches separated by `gap` bases (half a gap at each genome end).
    struct Contig { int64_t start, len; };
    std::vector<Contig> contigs;
    int64_t stride = G / n_contigs;
    for (int c = 0; c < n_contigs; c++) contigs.push_back({c * stride + gap / 2, stride - gap});

    // Repeat families: `copies` copies of a 5-20 kb segment each, at random non-overlapping loci.
    std::vector<Copy> cps;
    for (int f = 0; f < families; f++) {
        int64_t rl = 5000 + static_cast<int64_t>(rng.below(15001));
        for (int c = 0; c < copies; c++) {
            for (int attempt = 0; attempt < 1000; attempt++) {
                int64_t s = static_cast<int64_t>(rng.below(static_cast<uint64_t>(G - rl)));
                bool clash = false;
                for (const Copy &o : cps) if (s < o.start + o.len && o.start < s + rl) { clash = true; break; }
                if (!clash) { cps.push_back({f, s, rl}); break; }
            }
        }
    }

    // ---- FASTA
    if (fasta) {
        std::string genome(static_cast<size_t>(G), 'A');
        Rng gr(seed ^ 0xA5A5A5A55A5A5A5Aull);
        for (int64_t i = 0; i < G; i += 32) {
            uint64_t x = gr.next();
            for (int64_t j = i; j < std::min(G, i + 32); j++, x >>= 2) genome[static_cast<size_t>(j)] = "ACGT"[x & 3];
        }
        for (int f = 0; f < families; f++) {
            const Copy *first = nullptr;
            for (const Copy &c : cps) {
                if (c.family != f) continue;
                if (!first) first = &c;
                else genome.replace(static_cast<size_t>(c.start), static_cast<size_t>(c.len), genome, static_cast<size_t>(first->start), static_cast<size_t>(c.len));
            }
        }
        Out fr(out_dir + "/reads.fasta"), fc(out_dir + "/contigs.fasta");
        char hdr[64];
        for (int64_t i = 0; i < n_reads; i++) {
            int n = std::snprintf(hdr, sizeof hdr, ">read%07" PRId64 "\n", i);
            fr.write(hdr, static_cast<size_t>(n));
            fr.write(genome.data() + reads[static_cast<size_t>(i)].start, static_cast<size_t>(reads[static_cast<size_t>(i)].len));
            fr.write("\n", 1);
        }
        for (size_t c = 0; c < contigs.size(); c++) {
            int n = std::snprintf(hdr, sizeof hdr, ">ctg%zu\n", c + 1);
            fc.write(hdr, static_cast<size_t>(n));
            fc.write(genome.data() + contigs[c].start, static_cast<size_t>(contigs[c].len));
            fc.write("\n", 1);
        }
    }

    auto pair_hash = [&](int64_t a, int64_t b, uint64_t salt) { return mix(static_cast<uint64_t>(std::min(a, b)), static_cast<uint64_t>(std::max(a, b)), seed ^ salt); };

    // ---- rr PAF, grouped by query.  Per query: genome-adjacent targets in start order, then
    // repeat-induced targets.
    struct RepHit { int copy; int64_t read; int64_t lo, hi; };  // read covers [lo,hi) of the repeat unit
    std::vector<std::vector<RepHit>> fam_hits(static_cast<size_t>(families));
    std::vector<std::vector<int>> read_hits(static_cast<size_t>(n_reads));  // indices into fam_hits[f] are stored per family below
    std::vector<std::vector<std::pair<int, int>>> read_fam(static_cast<size_t>(n_reads));
    for (size_t ci = 0; ci < cps.size(); ci++) {
        const Copy &c = cps[ci];
        // reads intersecting the copy: starts in (c.start - maxlen, c.start + c.len)
        int64_t maxlen = L + L / 10;
        auto lo_it = std::lower_bound(reads.begin(), reads.end(), c.start - maxlen, [](const Read &r, int64_t v) { return r.start < v; });
        for (auto it = lo_it; it != reads.end() && it->start < c.start + c.len; ++it) {
            int64_t lo = std::max(it->start, c.start), hi = std::min(it->start + it->len, c.start + c.len);
            if (hi - lo < min_ovl) continue;
            int64_t rid = it - reads.begin();
            fam_hits[static_cast<size_t>(c.family)].push_back({static_cast<int>(ci), rid, lo - c.start, hi - c.start});
            read_fam[static_cast<size_t>(rid)].push_back({c.family, static_cast<int>(fam_hits[static_cast<size_t>(c.family)].size() - 1)});
        }
    }

    int64_t rr_records = 0, cr_records = 0;
    {
        Out rr(out_dir + "/rr_overlaps.paf");
        char qn[32], tn[32];
        int64_t maxlen = L + L / 10;
        for (int64_t a = 0; a < n_reads && (max_rr == 0 || rr_records < max_rr); a++) {
            const Read &ra = reads[static_cast<size_t>(a)];
            std::snprintf(qn, sizeof qn, "read%07" PRId64, a);
            // candidates: reads with start in (ra.start - maxlen, ra.start + ra.len)
            auto lo_it = std::lower_bound(reads.begin(), reads.end(), ra.start - maxlen, [](const Read &r, int64_t v) { return r.start < v; });
            for (auto it = lo_it; it != reads.end() && it->start < ra.start + ra.len; ++it) {
                int64_t b = it - reads.begin();
                if (b == a) continue;
                const Read &rb = *it;
                int64_t lo = std::max(ra.start, rb.start), hi = std::min(ra.start + ra.len, rb.start + rb.len);
                uint64_t h = pair_hash(a, b, 0x11);
                if (jitter > 0) { lo += static_cast<int64_t>(h % static_cast<uint64_t>(jitter + 1)); hi -= static_cast<int64_t>((h >> 20) % static_cast<uint64_t>(jitter + 1)); }
                if (hi - lo < min_ovl) continue;
                double ident = 0.85 + 0.14 * static_cast<double>((h >> 40) & 0xFFFF) / 65535.0;
                int64_t ovl = hi - lo;
                int64_t m = static_cast<int64_t>(ident * static_cast<double>(ovl));
                char strand = (static_cast<double>((h >> 8) & 0xFFF) / 4096.0 < minus_frac) ? '-' : '+';
                std::snprintf(tn, sizeof tn, "read%07" PRId64, b);
                paf_record(rr, qn, ra.len, lo - ra.start, hi - ra.start, strand, tn, rb.len, lo - rb.start, hi - rb.start, m, ovl);
                rr_records++;
            }
            // repeat-induced: other copies of the same family
            for (auto &fh : read_fam[static_cast<size_t>(a)]) {
                const RepHit &ha = fam_hits[static_cast<size_t>(fh.first)][static_cast<size_t>(fh.second)];
                const Copy &ca = cps[static_cast<size_t>(ha.copy)];
                for (const RepHit &hb : fam_hits[static_cast<size_t>(fh.first)]) {
                    if (hb.copy == ha.copy || hb.read == a) continue;
                    int64_t lo = std::max(ha.lo, hb.lo), hi = std::min(ha.hi, hb.hi);
                    if (hi - lo < min_ovl) continue;
                    const Copy &cb = cps[static_cast<size_t>(hb.copy)];
                    const Read &rb = reads[static_cast<size_t>(hb.read)];
                    uint64_t h = pair_hash(a, hb.read, 0x22);
                    double ident = 0.85 + 0.14 * static_cast<double>((h >> 40) & 0xFFFF) / 65535.0;
                    int64_t ovl = hi - lo;
                    int64_t m = static_cast<int64_t>(ident * static_cast<double>(ovl));
                    std::snprintf(tn, sizeof tn, "read%07" PRId64, hb.read);
                    paf_record(rr, qn, ra.len, ca.start + lo - ra.start, ca.start + hi - ra.start, '+', tn, rb.len,
                               cb.start + lo - rb.start, cb.start + hi - rb.start, m, ovl);
                    rr_records++;
                }
            }
        }
    }
    {
        Out cr(out_dir + "/cr_overlaps.paf");
        char cn[32], rn[32];
        int64_t maxlen = L + L / 10;
        for (size_t c = 0; c < contigs.size(); c++) {
            const Contig &cg = contigs[c];
            std::snprintf(cn, sizeof cn, "ctg%zu", c + 1);
            auto lo_it = std::lower_bound(reads.begin(), reads.end(), cg.start - maxlen, [](const Read &r, int64_t v) { return r.start < v; });
            for (auto it = lo_it; it != reads.end() && it->start < cg.start + cg.len; ++it) {
                int64_t rid = it - reads.begin();
                const Read &rd = *it;
                int64_t lo = std::max(cg.start, rd.start), hi = std::min(cg.start + cg.len, rd.start + rd.len);
                if (hi - lo < min_ovl) continue;
                uint64_t h = pair_hash(static_cast<int64_t>(c), rid, 0x33);
                double ident = 0.85 + 0.14 * static_cast<double>((h >> 40) & 0xFFFF) / 65535.0;
                int64_t ovl = hi - lo;
                int64_t m = static_cast<int64_t>(ident * static_cast<double>(ovl));
                std::snprintf(rn, sizeof rn, "read%07" PRId64, rid);
                bool read_is_query = rd.start < cg.start || rd.start + rd.len <= cg.start + cg.len;
                if (read_is_query)
                    paf_record(cr, rn, rd.len, lo - rd.start, hi - rd.start, '+', cn, cg.len, lo - cg.start, hi - cg.start, m, ovl);
                else
                    paf_record(cr, cn, cg.len, lo - cg.start, hi - cg.start, '+', rn, rd.len, lo - rd.start, hi - rd.start, m, ovl);
                cr_records++;
            }
        }
    }
    std::printf("{\"reads\": %" PRId64 ", \"contigs\": %zu, \"rr_records\": %" PRId64 ", \"cr_records\": %" PRId64
                ", \"repeat_copies\": %zu}\n", n_reads, contigs.size(), rr_records, cr_records, cps.size());
    return 0;
}
