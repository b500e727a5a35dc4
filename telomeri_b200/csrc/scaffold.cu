// Scaffold: the chain of consensus paths main() builds and the sequence stitching of Scaffolder.
//
// Replaces, on device-resident results, PathManager::constructConsensusPath (reference
// src/PathManager.cpp:664-767, called from src/Main.cpp:344-353 with min_path_num = 10) and Scaffolder
// (src/Scaffolder.cpp:12-147): load of reads/contigs FASTA, name lookup, substring / complement / reverse
// stitching.  The chain itself is O(#anchor pairs) bookkeeping and runs on the host over the tables
// hera_groups produced; what is O(data) -- indexing the FASTA files (the reference keeps every read in
// a vector and finds names by linear search, src/Scaffolder.cpp:104-112) and moving sequence bytes --
// runs on the GPU: newline index by chunk counts + scan, one thread per record hashes the header word
// into an open-addressing table, the scaffold's nodes are looked up (bytes verified), and one kernel
// writes the scaffold with one thread per output byte.
#include <algorithm>
#include <cstring>
#include <fstream>
#include <map>

#include "ctx.cuh"
#include "primitives.cuh"

namespace hera {

namespace {

constexpr int FA_CHUNK = 256;

__device__ __forceinline__ bool fa_space(uint32_t c) { return c == ' ' || (c - 9u) <= 4u; }

__device__ __forceinline__ uint64_t fa_hash(const uint8_t *p, uint64_t len) {
    uint64_t h = 0xcbf29ce484222325ull;
    for (uint64_t i = 0; i < len; i++) h = (h ^ p[i]) * 0x100000001b3ull;
    h ^= h >> 32;
    return h ? h : 1;
}

__global__ void fa_count_kernel(const uint8_t *__restrict__ text, uint64_t n, uint32_t *__restrict__ counts) {
    const uint64_t ch = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t b = ch * FA_CHUNK;
    if (b >= n) return;
    const uint64_t e = min(n, b + FA_CHUNK);
    uint32_t c = 0;
    for (uint64_t i = b; i < e; i++) c += text[i] == '\n';
    counts[ch] = c;
}

// line_start[k] = first byte of line k (line 0 starts at 0; line k starts after the k-th newline)
__global__ void fa_lines_kernel(const uint8_t *__restrict__ text, uint64_t n, const uint32_t *__restrict__ count_scan, uint64_t *__restrict__ line_start) {
    const uint64_t ch = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t b = ch * FA_CHUNK;
    if (b >= n) return;
    if (ch == 0) line_start[0] = 0;
    const uint64_t e = min(n, b + FA_CHUNK);
    uint32_t k = count_scan[ch];
    for (uint64_t i = b; i < e; i++) if (text[i] == '\n') line_start[++k] = i + 1;
}

struct FastaDev {
    const uint8_t *text;
    uint64_t n;
    const uint64_t *line_start;   // lines + 1 entries, the last one = n + 1 sentinel handled by line_end
    uint64_t lines, step, records;
};

__device__ __forceinline__ uint64_t line_end(const FastaDev &f, uint64_t k) {   // exclusive, without the '\n'
    return k + 1 < f.lines ? f.line_start[k + 1] - 1 : (f.n > 0 && f.text[f.n - 1] == '\n' ? f.n - 1 : f.n);
}

// header word of record r: `iss >> word; name = word.substr(1)` (src/Scaffolder.cpp:24-27)
__device__ __forceinline__ void record_name(const FastaDev &f, uint64_t r, uint64_t &b, uint64_t &len) {
    const uint64_t l = r * f.step;
    uint64_t p = f.line_start[l];
    const uint64_t e = line_end(f, l);
    while (p < e && fa_space(f.text[p])) p++;
    uint64_t q = p;
    while (q < e && !fa_space(f.text[q])) q++;
    if (q > p) p++;   // drop '>' / '@'
    b = p;
    len = q - p;
}

__global__ void fa_insert_kernel(FastaDev f, uint32_t rec_base, unsigned long long *keys, uint32_t *first_rec, uint64_t mask) {
    const uint64_t r = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (r >= f.records) return;
    uint64_t b, len;
    record_name(f, r, b, len);
    const unsigned long long h = fa_hash(f.text + b, len);
    uint64_t slot = h & mask;
    for (;;) {
        unsigned long long cur = keys[slot];
        if (cur == 0) { cur = atomicCAS(&keys[slot], 0ull, h); if (cur == 0) cur = h; }
        if (cur == h) { atomicMin(&first_rec[slot], rec_base + static_cast<uint32_t>(r)); return; }   // std::find: the first record with that name
        slot = (slot + 1) & mask;
    }
}

struct SeqRef {
    uint64_t off, len;
    uint32_t file, found;
};

// One thread per scaffold node: name from the graph's name blob -> record -> its sequence line.
__global__ void fa_lookup_kernel(uint32_t count, const uint32_t *__restrict__ nodes, const uint64_t *__restrict__ name_off, const char *__restrict__ names,
                                 FastaDev f0, FastaDev f1, const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ first_rec, uint64_t mask,
                                 SeqRef *__restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint32_t x = nodes[i];
    const uint8_t *nm = reinterpret_cast<const uint8_t *>(names) + name_off[x];
    const uint64_t nl = name_off[x + 1] - name_off[x];
    const unsigned long long h = fa_hash(nm, nl);
    SeqRef res{0, 0, 0, 0};
    uint64_t slot = h & mask;
    for (uint64_t probes = 0; probes <= mask; probes++) {
        const unsigned long long cur = keys[slot];
        if (cur == 0) break;
        if (cur == h) {
            uint32_t r = first_rec[slot];
            const bool second = r >= f0.records;
            const FastaDev &f = second ? f1 : f0;
            const uint64_t rr = second ? r - f0.records : r;
            uint64_t b, len;
            record_name(f, rr, b, len);
            bool same = len == nl;
            for (uint64_t k = 0; same && k < nl; k++) same = f.text[b + k] == nm[k];
            const uint64_t sl = rr * f.step + 1;
            if (same && sl < f.lines) { res.off = f.line_start[sl]; res.len = line_end(f, sl) - res.off; res.file = second ? 1 : 0; res.found = 1; }
            break;
        }
        slot = (slot + 1) & mask;
    }
    out[i] = res;
}

struct Segment {   // out[dst .. dst + len) = (complement?) text[file][src .. src + len)
    uint64_t dst, src, len;
    uint32_t file, complement;
};

__global__ void stitch_kernel(uint64_t total, uint32_t nseg, const Segment *__restrict__ seg, const uint8_t *__restrict__ t0, const uint8_t *__restrict__ t1,
                              char *__restrict__ out) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= total) return;
    uint32_t lo = 0, hi = nseg;   // last segment with dst <= i
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (seg[mid].dst <= i) lo = mid; else hi = mid; }
    const Segment s = seg[lo];
    uint8_t c = (s.file ? t1 : t0)[s.src + (i - s.dst)];
    if (s.complement) c = c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'G' ? 'C' : c == 'C' ? 'G' : c;   // swapBases, src/Scaffolder.cpp:126-140
    out[i] = static_cast<char>(c);
}

struct SlotEnds {
    uint32_t q_end, t_end, strand;
};

// Coordinates of a path edge as its owner's list stores it (after the q/t swap of src/OverlapGraph.cpp:245-248).
__global__ void slot_ends_kernel(uint32_t count, const uint32_t *__restrict__ slots, const uint32_t *__restrict__ pay, const uint32_t *__restrict__ qe,
                                 const uint32_t *__restrict__ te, const uint8_t *__restrict__ flags, SlotEnds *__restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint32_t p = pay[slots[i]], e = p >> 1, side = p & 1;
    out[i] = SlotEnds{side ? te[e] : qe[e], side ? qe[e] : te[e], (flags[e] & KF_MINUS) ? 1u : 0u};
}

struct FastaFile {
    DBuf<uint8_t> text;
    DBuf<uint64_t> line_start;
    uint64_t n = 0, lines = 0, step = 2, records = 0;
    FastaDev dev() const { return FastaDev{text.p, n, line_start.p, lines, step, records}; }
};

void load_fasta(Ctx &c, const char *path, FastaFile &f) {
    std::ifstream in(path, std::ios::binary | std::ios::ate);
    if (!in) throw Error{1, std::string("Cannot open file: ") + path};
    f.n = static_cast<uint64_t>(in.tellg());
    in.seekg(0);
    const size_t plen = std::strlen(path);
    f.step = plen > 0 && path[plen - 1] == 'q' ? 4 : 2;   // fastq: '+' line and qualities are skipped (src/Scaffolder.cpp:33-36)
    HBuf<char> staging;
    staging.reserve(f.n + 1);
    if (f.n) in.read(staging.p, static_cast<std::streamsize>(f.n));
    HERA_REQUIRE(static_cast<uint64_t>(in.gcount()) == f.n || f.n == 0, 1, std::string("short read on ") + path);
    f.text.reserve(f.n + 1);
    if (f.n) HERA_CUDA(cudaMemcpyAsync(f.text.p, staging.p, f.n, cudaMemcpyHostToDevice, c.stream));
    f.lines = 0;
    f.records = 0;
    if (f.n == 0) { HERA_CUDA(cudaStreamSynchronize(c.stream)); return; }
    const uint64_t chunks = (f.n + FA_CHUNK - 1) / FA_CHUNK;
    TempPool tmp(c.stream);
    uint32_t *counts = tmp.get<uint32_t>(chunks), *d_total = tmp.get<uint32_t>(1);
    HERA_LAUNCH(c, fa_count_kernel, div_up(chunks, 128), 128, 0, f.text.p, f.n, counts);
    exclusive_scan_u32(c, counts, chunks, counts, d_total);
    uint32_t newlines = 0;
    HERA_CUDA(cudaMemcpyAsync(&newlines, d_total, sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
    HERA_CUDA(cudaStreamSynchronize(c.stream));   // also: staging may go
    f.line_start.reserve(static_cast<uint64_t>(newlines) + 2);
    HERA_LAUNCH(c, fa_lines_kernel, div_up(chunks, 128), 128, 0, f.text.p, f.n, counts, f.line_start.p);
    f.lines = newlines + (staging.p[f.n - 1] != '\n' ? 1 : 0);   // std::getline: a trailing newline does not open another line
    f.records = (f.lines + f.step - 1) / f.step;
    HERA_CUDA(cudaStreamSynchronize(c.stream));
}

}  // namespace

void build_scaffold(Ctx &c, const char *reads_path, const char *contigs_path, uint64_t min_path_num, hera_scaffold_view *out) {
    HERA_REQUIRE(c.groups_valid, 41, "hera_scaffold: call hera_groups first");
    const RowView rows = rows_view(c);   // coordinates of the path edges (collective on a sharded context)
    HERA_REQUIRE(reads_path && contigs_path, 101, "null path");
    const PathStore &ps = c.uniq_ready ? c.uniq : c.all;
    auto &m = c.grm;
    auto &sm = c.scm;
    sm.nodes.clear(); sm.slots.clear(); sm.logged.clear();
    *out = hera_scaffold_view{};
    Events<2> ev;   // events and temporaries go when the scope ends, exception or not
    const cudaEvent_t ev0 = ev[0], ev1 = ev[1];
    TempPool tmp(c.stream);

    // ---- the chain of consensus paths (src/PathManager.cpp:664-767), host bookkeeping over the group tables
    using Key = std::pair<uint32_t, uint32_t>;
    std::map<Key, std::pair<int64_t, uint64_t>> pool;   // anchor pair -> (consensus path, number of paths of the pair)
    std::vector<uint32_t> anchors;
    uint64_t best_num = 0;
    Key best_key{UINT32_MAX, UINT32_MAX};
    for (size_t p = 0; p < m.pair_a1.size(); p++) {
        if (m.pair_consensus[p] < 0) continue;
        const uint64_t num = m.pair_off[p + 1] - m.pair_off[p];
        if (num < min_path_num) continue;
        for (uint32_t a : {m.pair_a1[p], m.pair_a2[p]})
            if (std::find(anchors.begin(), anchors.end(), a) == anchors.end()) anchors.push_back(a);
        const Key k{m.pair_a1[p], m.pair_a2[p]};
        pool[k] = {m.pair_consensus[p], num};
        if (best_num < num) { best_num = num; best_key = k; }
    }
    if (!pool.count(best_key)) throw Error{40, "No consensus found!"};   // the reference's map::at throws here; main() prints this when no group has a consensus
    std::vector<uint64_t> h_off(ps.paths + 1);
    HERA_CUDA(cudaMemcpyAsync(h_off.data(), ps.off.p, (ps.paths + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, c.stream));
    HERA_CUDA(cudaStreamSynchronize(c.stream));
    auto fetch = [&](int64_t id, std::vector<uint32_t> &pn, std::vector<uint32_t> &pslots) {
        const uint64_t b = h_off[id], n = h_off[id + 1] - b;
        pn.resize(n);
        pslots.resize(n - 1);
        HERA_CUDA(cudaMemcpyAsync(pn.data(), ps.node.p + b, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaMemcpyAsync(pslots.data(), ps.slot.p + (b - static_cast<uint64_t>(id)), (n - 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
    };
    std::vector<uint32_t> pn, pslots;
    fetch(pool[best_key].first, pn, pslots);
    sm.nodes = pn;
    sm.slots = pslots;
    uint32_t left = best_key.first, right = best_key.second;
    const std::vector<uint32_t> seen = {left, right};   // never grows in the reference (:708-710)
    while (!pool.empty()) {
        bool grow_left = false, any = false;
        uint64_t top = 0;
        Key pick = best_key;
        for (uint32_t n : anchors) {
            if (std::find(seen.begin(), seen.end(), n) != seen.end()) continue;
            const Key lk{n, left};   // the cursors move inside the scan (:722-738)
            auto it = pool.find(lk);
            if (it != pool.end() && top < it->second.second) { top = it->second.second; pick = lk; grow_left = true; any = true; left = n; }
            const Key rk{right, n};
            it = pool.find(rk);
            if (it != pool.end() && top < it->second.second) { top = it->second.second; pick = rk; grow_left = false; any = true; right = n; }
        }
        if (!any) break;
        fetch(pool[pick].first, pn, pslots);
        if (!grow_left) {
            sm.nodes.insert(sm.nodes.end(), pn.begin() + 1, pn.end());
            sm.slots.insert(sm.slots.end(), pslots.begin(), pslots.end());
        } else {
            sm.nodes.insert(sm.nodes.begin(), pn.begin(), pn.end() - 1);
            sm.slots.insert(sm.slots.begin(), pslots.begin(), pslots.end());
        }
        pool.erase(pick);
    }
    const uint32_t nn = static_cast<uint32_t>(sm.nodes.size()), ne = static_cast<uint32_t>(sm.slots.size());
    HERA_REQUIRE(nn >= 2 && ne == nn - 1, 42, "scaffold path with fewer than two nodes");

    // ---- Scaffolder: FASTA index, name lookup, stitching (src/Scaffolder.cpp:12-147)
    HERA_CUDA(cudaEventRecord(ev0, c.stream));
    FastaFile reads, contigs;
    load_fasta(c, reads_path, reads);
    load_fasta(c, contigs_path, contigs);
    const uint64_t records = reads.records + contigs.records;
    HERA_REQUIRE(records < (1ull << 31), 43, "too many FASTA records");
    uint64_t cap = 1024;
    while (cap < 2 * records) cap <<= 1;
    unsigned long long *keys = tmp.get<unsigned long long>(cap);
    uint32_t *first_rec = tmp.get<uint32_t>(cap);
    HERA_CUDA(cudaMemsetAsync(keys, 0, cap * sizeof(unsigned long long), c.stream));
    HERA_CUDA(cudaMemsetAsync(first_rec, 0xff, cap * sizeof(uint32_t), c.stream));
    if (reads.records) HERA_LAUNCH(c, fa_insert_kernel, div_up(reads.records, 128), 128, 0, reads.dev(), 0u, keys, first_rec, cap - 1);
    if (contigs.records) HERA_LAUNCH(c, fa_insert_kernel, div_up(contigs.records, 128), 128, 0, contigs.dev(), static_cast<uint32_t>(reads.records), keys, first_rec, cap - 1);
    uint32_t *d_nodes = tmp.get<uint32_t>(nn), *d_slots = tmp.get<uint32_t>(ne);
    SeqRef *d_refs = tmp.get<SeqRef>(nn);
    SlotEnds *d_ends = tmp.get<SlotEnds>(ne);
    HERA_CUDA(cudaMemcpyAsync(d_nodes, sm.nodes.data(), nn * sizeof(uint32_t), cudaMemcpyHostToDevice, c.stream));
    HERA_CUDA(cudaMemcpyAsync(d_slots, sm.slots.data(), ne * sizeof(uint32_t), cudaMemcpyHostToDevice, c.stream));
    HERA_LAUNCH(c, fa_lookup_kernel, div_up(nn, 128), 128, 0, nn, d_nodes, c.g.node_name_off.p, c.g.node_names.p, reads.dev(), contigs.dev(), keys, first_rec, cap - 1, d_refs);
    HERA_LAUNCH(c, slot_ends_kernel, div_up(ne, 128), 128, 0, ne, d_slots, c.g.adj_pay.p, rows.qe, rows.te, rows.flags, d_ends);
    std::vector<SeqRef> refs(nn);
    std::vector<SlotEnds> ends(ne);
    HERA_CUDA(cudaMemcpyAsync(refs.data(), d_refs, nn * sizeof(SeqRef), cudaMemcpyDeviceToHost, c.stream));
    HERA_CUDA(cudaMemcpyAsync(ends.data(), d_ends, ne * sizeof(SlotEnds), cudaMemcpyDeviceToHost, c.stream));
    HERA_CUDA(cudaStreamSynchronize(c.stream));

    // pieces in the order the reference appends them; every piece stands for reverse(substring)
    struct Piece { uint64_t src, len; uint32_t file, comp; };
    std::vector<Piece> pieces;
    uint64_t seq_len = 0;
    auto append = [&](uint32_t i, uint64_t start, uint64_t end, bool comp) {
        HERA_REQUIRE(refs[i].found, 44, "a scaffold node has no sequence in the FASTA files");
        HERA_REQUIRE(start <= refs[i].len, 45, "scaffold substring starts behind the end of the sequence");   // std::out_of_range in the reference
        const uint64_t count = end - start + 1;   // unsigned, like substring() (:110-113)
        const uint64_t len = std::min<uint64_t>(count, refs[i].len - start);
        pieces.push_back(Piece{refs[i].off + start, len, refs[i].file, comp ? 1u : 0u});
        seq_len += len;
        sm.logged.push_back(i);
    };
    append(0, static_cast<uint64_t>(ends[0].t_end) + 1, refs[0].len, false);
    for (uint32_t i = 1; i < ne; i++) {
        const uint64_t start = ends[i].t_end;     // where the next overlap ends on this node
        const uint64_t end = ends[i - 1].q_end;   // where the previous overlap ends on this node
        if (start < end) {
            append(i, start + 1, end, ends[i - 1].strand != 0);
        } else if (start > end) {
            const uint64_t keep = seq_len - (start - end);   // wraps like the reference's ulong arithmetic: then nothing is cut
            if (keep < seq_len) {
                uint64_t cut = seq_len - keep;
                seq_len = keep;
                while (cut > 0 && !pieces.empty()) {   // the tail of reverse(piece) is the head of the piece
                    Piece &p = pieces.back();
                    const uint64_t x = std::min(cut, p.len);
                    p.src += x; p.len -= x; cut -= x;
                    if (p.len == 0) pieces.pop_back();
                }
            }
        }
    }
    append(nn - 1, 0, ends[ne - 1].q_end, false);
    // the final global reverse puts the pieces in reverse order, each one forward
    const char header[] = ">Resulting_HERA_scaffold\n";
    const uint64_t hl = sizeof(header) - 1;
    std::vector<Segment> segs;
    uint64_t dst = 0;
    for (size_t k = pieces.size(); k-- > 0;) {
        if (!pieces[k].len) continue;
        segs.push_back(Segment{dst, pieces[k].src, pieces[k].len, pieces[k].file, pieces[k].comp});
        dst += pieces[k].len;
    }
    sm.fasta.reserve(hl + seq_len + 2);
    std::memcpy(sm.fasta.p, header, hl);
    if (seq_len) {
        Segment *d_seg = tmp.get<Segment>(segs.size());
        char *d_out = tmp.get<char>(seq_len);
        HERA_CUDA(cudaMemcpyAsync(d_seg, segs.data(), segs.size() * sizeof(Segment), cudaMemcpyHostToDevice, c.stream));
        HERA_LAUNCH(c, stitch_kernel, div_up(seq_len, 256), 256, 0, seq_len, static_cast<uint32_t>(segs.size()), d_seg, reads.text.p, contigs.text.p, d_out);
        HERA_CUDA(cudaMemcpyAsync(sm.fasta.p + hl, d_out, seq_len, cudaMemcpyDeviceToHost, c.stream));
    }
    HERA_CUDA(cudaEventRecord(ev1, c.stream));
    HERA_CUDA(cudaStreamSynchronize(c.stream));
    sm.fasta.p[hl + seq_len] = '\n';
    float ms = 0;
    HERA_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
    out->nodes = nn; out->node = sm.nodes.data(); out->slot = sm.slots.data();
    out->logged = sm.logged.size(); out->logged_index = sm.logged.data();
    out->fasta_bytes = hl + seq_len + 1; out->fasta = sm.fasta.p; out->sequence_length = seq_len; out->ms = ms;
}

}  // namespace hera
