// K1 + K2: PAF text -> kept overlaps, one pass over the text.
//
// Replaces the read loop, filter and score block of OverlapGraph::load / filter / buildFrom
// (reference src/OverlapGraph.cpp:97-202, 221-231, 297-324).  The reference consumes the file
// as a whitespace-separated token stream, 16 tokens per record (12 typed + 4 skipped), so a
// record boundary is "token index % 16 == 0", not a newline.
//
// Persistent CTAs (320 threads, 3 per SM) work on 30 KiB tiles of text (plus a 1 KiB halo so the records that
// start in a tile can be finished).  Tiles are handed out by an atomic ticket so that a tile never waits for a
// tile that has not started (chained scans with decoupled look-back); a CTA keeps two tiles staged in shared
// memory and orders its work so that the chains are published early and looked up late:
//   1. asynchronous 16-byte copies (cp.async) stage tile+halo in shared memory; every thread classifies the
//      chunks it copied into a non-space bitmask with SWAR arithmetic (one bit per byte).
//   2. token starts = non-space after space; the tile's token count is published to the first chain; the
//      global index of the tile's first token is looked up an iteration later.  A sharded load adds the
//      tokens of the earlier ranks' byte ranges (ParseArgs::tok_base0).
//   3. every 16th global token opens a record; its offset is found with popcount selection.  One
//      thread takes one record: token boundaries come from the bitmask (funnel-shifted 32-bit
//      windows, ffs), integer fields are converted eight digits at a time from whole words, names
//      are hashed word-wise, then the fp32 filter and scores run in the reference's operation order
//      (explicit round-to-nearest intrinsics, no FMA contraction).
//   4. kept records are counted, the count is published to the second chain, the CTA stages its next tile,
//      then looks the chain up and writes its kept records at their final file-order position.
// A record is "regular" when its nine integer tokens are [+-]?digits within int range and the
// strand token is one character.  The first irregular record (stream extraction would fail or
// desynchronise there) ends the parallel part; a one-thread tail kernel continues from its byte
// offset with the exact character-stream rules of formatted extraction.
#include <algorithm>
#include <mutex>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <ctime>

#include "ctx.cuh"
#include "primitives.cuh"

namespace hera {

#ifndef PF_MIN_CTAS_D
#define PF_MIN_CTAS_D 3
#endif
#ifndef PF_TILE_D
#define PF_TILE_D 30720
#endif
#ifndef PF_THREADS_D
#define PF_THREADS_D 320
#endif
constexpr int PF_THREADS = PF_THREADS_D;
constexpr int PF_MIN_CTAS = PF_MIN_CTAS_D;
constexpr int PF_TILE = PF_TILE_D;
constexpr int PF_HALO = 1024;
constexpr int PF_WIN = PF_TILE + PF_HALO;
constexpr int PF_CHUNKS = PF_WIN / 16;
constexpr int PF_WPT = PF_TILE / 32 / PF_THREADS;   // 32-byte mask words owned by one thread
static_assert(PF_WPT * 32 * PF_THREADS == PF_TILE, "tile = threads x whole mask words");
constexpr int PF_MAXREC = PF_TILE / 32 + 2;   // a record is at least 16 tokens + 16 separators

struct FilterDev {
    int mode;
    float min_len_f, min_pct, max_ohl_f, max_pct;
};

struct ParseCounters {  // device block, zeroed per load
    unsigned int ticket;
    unsigned int name_too_long;         // a kept record has a sequence name of 65536 bytes or more (the name reference holds 16 bits of length)
    unsigned long long total_tokens;
    unsigned long long kept;
    unsigned long long first_bad_rec;   // min record index that is irregular (init UINT64_MAX)
    unsigned long long first_bad_off;   // its byte offset
    unsigned long long tail_records;    // records consumed by the tail kernel
    unsigned long long tail_kept;
};

struct StagedRecord {
    uint32_t qs, qe, ts, te, qlen, tlen, rec;
    float os, si, es;
    uint64_t qhash, thash, qname, tname;
    uint8_t flags;
};

__device__ __forceinline__ bool dev_is_space(uint32_t c) { return c == ' ' || (c - 9u) <= 4u; }

__device__ __forceinline__ uint64_t hash_finish(uint64_t h) {
    h ^= h >> 33; h *= 0xff51afd7ed558ccdull; h ^= h >> 33; h *= 0xc4ceb9fe1a85ec53ull; h ^= h >> 33;
    return h ? h : 1;  // 0 is the empty marker of the interning table
}
// Name hash: the name as little-endian 32-bit words (last one zero-padded) through two 32-bit
// multiplicative lanes, length folded in, mixed to 64 bits at the end.  add() feeds bytes (serial
// readers); the tile parser feeds whole words taken from shared memory (hash_name_window).
struct NameHash {
    uint32_t x = 0x811C9DC5u, y = 0x9747B28Cu, w = 0, nb = 0;
    __device__ __forceinline__ void mix(uint32_t word) { x = (x ^ word) * 0x01000193u; y = (y ^ word) * 0x85EBCA6Bu; }
    __device__ __forceinline__ void add(uint32_t c) {
        w |= c << (8 * (nb & 3));
        nb++;
        if ((nb & 3) == 0) { mix(w); w = 0; }
    }
    __device__ __forceinline__ uint64_t finish() {
        if (nb & 3) mix(w);
        return hash_finish((static_cast<uint64_t>(y ^ nb) << 32) | x);
    }
};

// x / 2.0f of the reference: the operand is an int converted to float (0 or |x| >= 1), so halving is exact
// and x * 0.5f gives the same bits as the IEEE division.
__device__ __forceinline__ float half_exact(float x) { return __fmul_rn(x, 0.5f); }

// fp32 filter, reference src/OverlapGraph.cpp:156-202 (name equality is checked by the caller).
__device__ __forceinline__ bool filter_fields(const FilterDev &f, int qlen, int qs, int qe, int tlen, int ts, int te) {
    const int q_ovl = static_cast<int>(static_cast<uint32_t>(qe) - static_cast<uint32_t>(qs));
    const int t_ovl = static_cast<int>(static_cast<uint32_t>(te) - static_cast<uint32_t>(ts));
    const int q_ohg = static_cast<int>(static_cast<uint32_t>(qlen) - static_cast<uint32_t>(qe));
    const int t_ohg = ts;
    float ovl, ohg, tot;
    if (f.mode == HERA_MODE_MIN) {
        ovl = __int2float_rn(min(q_ovl, t_ovl)); ohg = __int2float_rn(min(q_ohg, t_ohg)); tot = __int2float_rn(min(qlen, tlen));
    } else if (f.mode == HERA_MODE_MAX) {
        ovl = __int2float_rn(max(q_ovl, t_ovl)); ohg = __int2float_rn(max(q_ohg, t_ohg)); tot = __int2float_rn(max(qlen, tlen));
    } else {
        const float so = __int2float_rn(static_cast<int>(static_cast<uint32_t>(q_ovl) + static_cast<uint32_t>(t_ovl)));
        const float sh = __int2float_rn(static_cast<int>(static_cast<uint32_t>(q_ohg) + static_cast<uint32_t>(t_ohg)));
        const float st = __int2float_rn(static_cast<int>(static_cast<uint32_t>(qlen) + static_cast<uint32_t>(tlen)));
        if (f.mode == HERA_MODE_AVG) {
            ovl = half_exact(so); ohg = half_exact(sh); tot = half_exact(st);
        } else if (f.mode == HERA_MODE_SUM) {
            ovl = so; ohg = sh; tot = st;
        } else {
            return false;
        }
    }
    if (ovl < f.min_len_f || __fdiv_rn(ovl, tot) < f.min_pct) return false;
    return ohg <= f.max_ohl_f && __fdiv_rn(ohg, ovl) <= f.max_pct;
}

// SI / OS / ES, reference src/OverlapGraph.cpp:221-231,321-324, left to right in fp32.
__device__ __forceinline__ void score_fields(int qlen, int qs, int qe, int tlen, int ts, int te, int matches, float &si, float &os, float &es) {
    const int q_ovl = static_cast<int>(static_cast<uint32_t>(qe) - static_cast<uint32_t>(qs));
    const int t_ovl = static_cast<int>(static_cast<uint32_t>(te) - static_cast<uint32_t>(ts));
    const int q_ohg = static_cast<int>(static_cast<uint32_t>(qlen) - static_cast<uint32_t>(qe));
    si = __fdiv_rn(__int2float_rn(matches), __int2float_rn(min(qlen, tlen)));
    const float half_ovl = half_exact(__int2float_rn(static_cast<int>(static_cast<uint32_t>(q_ovl) + static_cast<uint32_t>(t_ovl))));
    os = __fmul_rn(half_ovl, si);
    const float ext = half_exact(__int2float_rn(qs));
    const float ohg = half_exact(__int2float_rn(static_cast<int>(static_cast<uint32_t>(q_ohg) + static_cast<uint32_t>(ts))));
    es = fabsf(__fsub_rn(__fadd_rn(os, ext), ohg));
}

// ---- exact character-stream reader (one thread) ---------------------------------------------
// Mirrors `fs >> string >> int ...` (src/OverlapGraph.cpp:109-127): string = run of non-space
// bytes, int = [+-]?digits stopping at the first other byte (fail: no digit / out of int range),
// char = one non-space byte.
struct SerialCursor {
    const uint8_t *t;
    uint64_t n, p;
    bool good;
    __device__ void skip() { while (p < n && dev_is_space(t[p])) p++; }
    __device__ bool word(uint64_t &off, uint32_t &len, uint64_t &hash) {
        if (!good) return false;
        skip();
        if (p >= n) return good = false;
        off = p;
        NameHash h;
        while (p < n && !dev_is_space(t[p])) { h.add(t[p]); p++; }
        len = static_cast<uint32_t>(p - off);
        hash = h.finish();
        return true;
    }
    __device__ bool integer(int &out) {
        if (!good) return false;
        skip();
        if (p >= n) return good = false;
        bool neg = false;
        if (t[p] == '+' || t[p] == '-') { neg = t[p] == '-'; p++; }
        unsigned long long v = 0;
        uint32_t digits = 0;
        bool over = false;
        while (p < n && t[p] >= '0' && t[p] <= '9') {
            if (!over) { v = v * 10 + (t[p] - '0'); if (v > (neg ? 2147483648ull : 2147483647ull)) over = true; }
            digits++; p++;
        }
        if (digits == 0 || over) return good = false;
        out = neg ? static_cast<int>(0u - static_cast<uint32_t>(v)) : static_cast<int>(v);
        return true;
    }
    __device__ bool single(uint32_t &ch) {
        if (!good) return false;
        skip();
        if (p >= n) return good = false;
        ch = t[p++];
        return true;
    }
};

struct SerialRecord {
    uint64_t qoff, toff, qhash, thash;
    uint32_t qnlen, tnlen, strand;
    int qlen, qs, qe, tlen, ts, te, matches, block, mapq;
};

__device__ bool serial_read_record(SerialCursor &c, SerialRecord &r) {
    bool ok = c.word(r.qoff, r.qnlen, r.qhash) && c.integer(r.qlen) && c.integer(r.qs) && c.integer(r.qe) && c.single(r.strand) &&
              c.word(r.toff, r.tnlen, r.thash) && c.integer(r.tlen) && c.integer(r.ts) && c.integer(r.te) && c.integer(r.matches) &&
              c.integer(r.block) && c.integer(r.mapq);
    if (!ok) return false;
    uint64_t o, h; uint32_t l;
    for (int k = 0; k < 4; k++) c.word(o, l, h);
    return true;
}

__device__ __forceinline__ bool names_equal(const uint8_t *text, uint64_t a, uint64_t b, uint32_t len) {
    for (uint32_t i = 0; i < len; i++) if (text[a + i] != text[b + i]) return false;
    return true;
}

__device__ __forceinline__ bool is_ctg(const uint8_t *text, uint64_t off, uint32_t len) {
    return len >= 3 && (text[off] | 0x20) == 'c' && (text[off + 1] | 0x20) == 't' && (text[off + 2] | 0x20) == 'g';
}

// Turns parsed fields into a staged record; returns keep flag.
__device__ __forceinline__ bool make_record(const FilterDev &f, const uint8_t *text, uint32_t file_id, bool anchors, const SerialRecord &r,
                                            uint32_t rec_idx, StagedRecord &out, ParseCounters *ctr, uint64_t name_off0 = 0) {
    if (r.qhash == r.thash && r.qnlen == r.tnlen && names_equal(text, r.qoff, r.toff, r.qnlen)) return false;  // :157
    if (!filter_fields(f, r.qlen, r.qs, r.qe, r.tlen, r.ts, r.te)) return false;
    if ((r.qnlen | r.tnlen) > 0xffffu) ctr->name_too_long = 1;   // (the reference takes names of any length; here the load fails with code 22)
    out.qs = static_cast<uint32_t>(r.qs); out.qe = static_cast<uint32_t>(r.qe);
    out.ts = static_cast<uint32_t>(r.ts); out.te = static_cast<uint32_t>(r.te);
    out.qlen = static_cast<uint32_t>(r.qlen); out.tlen = static_cast<uint32_t>(r.tlen);
    out.rec = rec_idx;
    score_fields(r.qlen, r.qs, r.qe, r.tlen, r.ts, r.te, r.matches, out.si, out.os, out.es);
    out.qhash = r.qhash; out.thash = r.thash;
    out.qname = pack_name(file_id, name_off0 + r.qoff, r.qnlen);   // (name_off0: where `text` starts in the loaded text, chunked loads)
    out.tname = pack_name(file_id, name_off0 + r.toff, r.tnlen);
    out.flags = (r.strand == '-' ? KF_MINUS : 0) | (anchors ? KF_CR : 0) | ((anchors && is_ctg(text, r.qoff, r.qnlen)) ? KF_QCTG : 0);
    return true;
}

__device__ __forceinline__ void store_record(const KeptPtrs &k, uint64_t i, const StagedRecord &s) {
    if (i >= k.cap) return;   // the table was sized optimistically and this text keeps more rows: the host sees kept > room and runs again
    k.qs[i] = s.qs; k.qe[i] = s.qe; k.ts[i] = s.ts; k.te[i] = s.te; k.qlen[i] = s.qlen; k.tlen[i] = s.tlen; k.rec[i] = s.rec;
    k.os[i] = s.os; k.si[i] = s.si; k.es[i] = s.es; k.flags[i] = s.flags;
    k.qhash[i] = s.qhash; k.thash[i] = s.thash; k.qname[i] = s.qname; k.tname[i] = s.tname;
}

// ---- look-back chain ------------------------------------------------------------------------
// state word: bits 63..62 = 0 empty / 1 tile aggregate / 2 inclusive prefix; low 62 bits = value.
constexpr unsigned long long ST_AGG = 1ull << 62, ST_INC = 2ull << 62, ST_MASK = 3ull << 62;

__device__ __forceinline__ unsigned long long ld_state(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_state(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// The chain is used in two steps so that a tile never has to wait: chain_publish makes the tile's aggregate
// visible as soon as it is known; chain_resolve, called (by warp 0) after the CTA has done other work, walks
// back over earlier tiles until it meets an inclusive prefix, publishes this tile's inclusive prefix and
// returns the exclusive one (same value in every lane).
__device__ __forceinline__ void chain_publish(unsigned long long *state, unsigned int tile, unsigned long long aggregate) {
    st_state(&state[tile], (tile == 0 ? ST_INC : ST_AGG) | aggregate);
}
// First look-back window of a tile, loaded early (lane l: state of tile - 1 - l) so that the round trip to L2
// overlaps with other work; chain_resolve re-reads whatever was still empty.
__device__ __forceinline__ unsigned long long chain_probe(const unsigned long long *state, unsigned int tile) {
    const int idx = static_cast<int>(tile) - 1 - static_cast<int>(lane_id());
    return idx >= 0 ? ld_state(&state[idx]) : ST_INC;   // tiles before 0 behave like an inclusive prefix of 0
}
__device__ unsigned long long chain_resolve(unsigned long long *state, unsigned int tile, unsigned long long aggregate, unsigned long long probe) {
    const uint32_t lane = lane_id();
    if (tile == 0) return 0;
    unsigned long long excl = 0;
    int look = static_cast<int>(tile) - 1;
    unsigned long long s = probe;
    for (;;) {
        const int idx = look - static_cast<int>(lane);
        if (idx >= 0) {
            while ((s & ST_MASK) == 0) s = ld_state(&state[idx]);
        }
        const uint32_t inc_mask = __ballot_sync(0xffffffffu, (s & ST_MASK) == ST_INC);
        const uint32_t upto = inc_mask ? static_cast<uint32_t>(__ffs(inc_mask)) : 32u;  // lanes [0, upto) contribute
        unsigned long long v = lane < upto ? (s & ~ST_MASK) : 0ull;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        excl += v;
        if (inc_mask) break;
        look -= 32;
        const int nidx = look - static_cast<int>(lane);
        s = nidx >= 0 ? ld_state(&state[nidx]) : ST_INC;
    }
    if (lane == 0) st_state(&state[tile], ST_INC | (excl + aggregate));
    return excl;
}

// Look-back for chains with MANY short tiles (the warp-autonomous kernel: one tile per 3 KiB): K windows of 32
// predecessors are requested together, so one round trip to L2 covers 32 K tiles.  With a single window the chain
// could resolve at most 32 tiles per round trip, less than the rate at which the warps of the GPU finish tiles.
template <int K>
__device__ unsigned long long chain_resolve_wide(unsigned long long *state, unsigned int tile, unsigned long long aggregate) {
    const uint32_t lane = lane_id();
    if (tile == 0) return 0;
    unsigned long long acc = 0;
    int look = static_cast<int>(tile) - 1;
    for (bool done = false; !done; look -= 32 * K) {
        unsigned long long s[K];
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int idx = look - static_cast<int>(lane) - 32 * k;
            s[k] = idx >= 0 ? ld_state(&state[idx]) : ST_INC;   // tiles before 0 behave like an inclusive prefix of 0
        }
#pragma unroll
        for (int k = 0; k < K; k++) {
            if (done) continue;   // warp-uniform
            const int idx = look - static_cast<int>(lane) - 32 * k;
            if (idx >= 0) {
                while ((s[k] & ST_MASK) == 0) s[k] = ld_state(&state[idx]);
            }
            const uint32_t inc_mask = __ballot_sync(0xffffffffu, (s[k] & ST_MASK) == ST_INC);
            const uint32_t upto = inc_mask ? static_cast<uint32_t>(__ffs(inc_mask)) : 32u;  // lanes [0, upto) contribute
            if (lane < upto) acc += s[k] & ~ST_MASK;
            done = inc_mask != 0;
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if (lane == 0) st_state(&state[tile], ST_INC | (acc + aggregate));
    return acc;
}

struct ParseArgs {
    const uint8_t *text;        // padded: readable up to tiles * PF_TILE + PF_HALO, bytes >= n are '\n' (unless `bare`)
    int bare;                   // the text is a caller's buffer: nothing behind its last 16-byte chunk may be read, bytes >= n count as white space
    uint64_t n;                 // bytes of text available (whole file, or slice + look-ahead)
    uint64_t owned;             // tokens that start at a byte >= owned belong to the next rank (== n for a whole file)
    uint64_t tok_base0;         // global index of the first token that starts in this text (0 for a whole file)
    const unsigned long long *tok_base0_dev;   // sharded load: the same number, computed on the device from all ranks' counts
    int prev_nonws;             // the byte before text[0] is part of a token
    int at_eof;                 // text[n] is the end of the file
    uint32_t tiles;
    uint32_t file_id;
    int anchors;
    uint64_t rec_limit;         // records with index >= rec_limit are ignored (test_load_num)
    uint64_t kept_base;         // kept records of earlier loads
    FilterDev filter;
    KeptPtrs kept;
    unsigned long long *tok_state, *kept_state;
    ParseCounters *ctr;
    // chunked load (a host text parsed chunk by chunk as it lands, one launch per chunk): rows kept by the earlier chunks
    // of this load, a device value (paf_parse_kernel only; ctr->kept and the chains stay relative to the chunk)
    const unsigned long long *kept_base_dev;
    uint64_t name_off0;         // where `text` starts in the loaded text: name references hold offsets into the whole text
};

// Exact reader for a record that does not fit the shared-memory window (very long names).
// Returns 1 = regular record in r, 0 = irregular (the tail kernel will deal with it).
__device__ __noinline__ int slow_path_record(const ParseArgs &a, uint64_t start, SerialRecord &r) {
    SerialCursor cur{a.text, a.n, start, true};
    if (!serial_read_record(cur, r)) return 0;
    // regular iff every typed token ended at white space
    SerialCursor chk{a.text, a.n, start, true};
    uint64_t o, hh; uint32_t l; int iv; uint32_t chh;
    bool whole = true;
    chk.word(o, l, hh);
    for (int f = 0; f < 3 && whole; f++) { chk.integer(iv); whole = chk.p >= a.n || dev_is_space(a.text[chk.p]); }
    if (whole) { chk.single(chh); whole = chk.p >= a.n || dev_is_space(a.text[chk.p]); }
    if (whole) chk.word(o, l, hh);
    for (int f = 0; f < 6 && whole; f++) { chk.integer(iv); whole = chk.p >= a.n || dev_is_space(a.text[chk.p]); }
    return whole ? 1 : 0;
}

// ---- field readers over the staged window -------------------------------------------------------
// `mask` has one bit per byte of the window (1 = not white space), so a token is a run of ones;
// text words are read whole and shifted into place -- no per-byte loops on the regular path.

// First position >= p whose mask bit equals `one`; >= PF_WIN when the window has none.
template <int WIN>
__device__ __noinline__ uint32_t next_bit(const uint32_t *mask, uint32_t p, bool one) {
    while (p < static_cast<uint32_t>(WIN)) {
        const uint32_t w = p >> 5;
        uint32_t win = __funnelshift_r(mask[w], mask[w + 1], p & 31u);
        if (!one) win = ~win;
        if (win) return p + __ffs(win) - 1;
        p += 32;
    }
    return WIN;
}

// Hash of the token [b, e): same function as NameHash fed byte by byte.
__device__ __forceinline__ uint64_t hash_name_window(const uint8_t *s, uint32_t b, uint32_t e) {
    const uint32_t *wp = reinterpret_cast<const uint32_t *>(s) + (b >> 2);   // s is 16-byte aligned
    const uint32_t sh = (b & 3u) * 8u, len = e - b;
    NameHash h;
    uint32_t lo = wp[0];
    uint32_t i = 0;
    for (; i + 4 <= len; i += 4) {
        const uint32_t hi = *++wp;
        h.mix(__funnelshift_r(lo, hi, sh));
        lo = hi;
    }
    if (len & 3u) h.mix(__funnelshift_r(lo, wp[1], sh) & ((1u << ((len & 3u) * 8u)) - 1u));
    h.nb = len;
    h.w = 0;
    return hash_finish((static_cast<uint64_t>(h.y ^ h.nb) << 32) | h.x);
}

// Four ASCII digits (first character in the low byte, already XORed with '0') -> value.
__device__ __forceinline__ uint32_t digits4(uint32_t d) {
    const uint32_t t = ((d * 2561u) >> 8) & 0x00ff00ffu;        // b0*10+b1 | b2*10+b3
    return (t * 6553601u) >> 16;                                // the product's upper half is (b0*10+b1)*100 + b2*10+b3
}

// Integer token [b, e), general form: [+-]?digits within int range.  Returns false when formatted extraction
// would fail or stop early on it (the caller declares the record irregular).
constexpr long long INT_TOKEN_BAD = 1ll << 40;
// (result by value: an out-parameter of a non-inlined call would force the caller's fields into local memory)
__device__ __noinline__ long long int_token_slow(const uint8_t *s, uint32_t b, uint32_t e) {
    const uint32_t c0 = s[b];
    const bool neg = c0 == '-';
    if (neg || c0 == '+') b++;
    if (e <= b) return INT_TOKEN_BAD;
    unsigned long long acc = 0;
    for (uint32_t i = b; i < e; i++) {
        const uint32_t d = static_cast<uint32_t>(s[i]) - '0';
        if (d > 9u) return INT_TOKEN_BAD;
        if (acc <= 4294967295ull) acc = acc * 10 + d;
    }
    if (acc > (neg ? 2147483648ull : 2147483647ull)) return INT_TOKEN_BAD;
    return neg ? -static_cast<long long>(acc) : static_cast<long long>(acc);
}

// Integer token [b, e).  Fast path: one to eight digits, read as the eight bytes that END at e (three aligned
// words, funnel-shifted), bytes before b cleared, all remaining bytes checked to be digits, then two
// four-digit conversions.  Signs, longer tokens and anything that is not a digit go to int_token_slow.
__device__ __forceinline__ bool int_token(const uint8_t *s, uint32_t b, uint32_t e, int &out) {
    const uint32_t len = e - b;
    const uint32_t q = e + 8u;                 // s - 16 is readable, so (e - 8) never underflows the buffer
    const uint32_t *wp = reinterpret_cast<const uint32_t *>(s - 16) + (q >> 2);
    const uint32_t sh = (q & 3u) * 8u;
    const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2];
    const uint32_t drop = (8u - len) * 8u;     // bits of the eight-byte window that belong to earlier text
    const uint32_t keep_lo = __funnelshift_lc(0u, 0xffffffffu, drop);                                      // drop >= 32: nothing kept
    const uint32_t keep_hi = __funnelshift_lc(0u, 0xffffffffu, static_cast<uint32_t>(max(static_cast<int>(drop) - 32, 0)));
    const uint32_t lo = (__funnelshift_r(w0, w1, sh) ^ 0x30303030u) & keep_lo;   // characters e-8 .. e-5
    const uint32_t hi = (__funnelshift_r(w1, w2, sh) ^ 0x30303030u) & keep_hi;   // characters e-4 .. e-1
    const uint32_t not_digit = (((lo + 0x76767676u) | lo) | ((hi + 0x76767676u) | hi)) & 0x80808080u;
    if (len - 1u > 7u || not_digit) {
        const long long v = int_token_slow(s, b, e);
        out = static_cast<int>(v);
        return v != INT_TOKEN_BAD;
    }
    out = static_cast<int>(digits4(lo) * 10000u + digits4(hi));
    return true;
}

// Span of the token that starts at b (a set mask bit, or PF_WIN when the window is used up): e = its end,
// nb = start of the next token.  One 32-bit window of the mask usually holds both.  Everything stays <= PF_WIN
// (the mask is zero behind the window), so a record that runs out of window just yields empty tokens.
template <int WIN>
__device__ __forceinline__ void token_span(const uint32_t *mask, uint32_t b, uint32_t &e, uint32_t &nb) {
    const uint32_t w = b >> 5;
    const uint32_t win = __funnelshift_r(mask[w], mask[w + 1], b & 31u);
    const uint32_t zeros = ~win;
    if (zeros) {
        const uint32_t eo = __ffs(zeros) - 1;
        e = b + eo;
        const uint32_t ones = win & (0xffffffffu << eo);
        nb = ones ? b + __ffs(ones) - 1 : next_bit<WIN>(mask, b + 32, true);
    } else {
        e = next_bit<WIN>(mask, b + 32, false);
        nb = next_bit<WIN>(mask, e, true);
    }
}

// One thread reads one record from the staged window: twelve tokens in straight-line code (fields stay in
// registers).  A token that is not what formatted extraction expects makes the record irregular, unless the
// window ran out first (then the exact reader decides).  Returns the keep flag.
template <int WIN>
__device__ __forceinline__ bool parse_one(const ParseArgs &a, const uint8_t *s, const uint32_t *mask, uint32_t p0, uint64_t tile_byte0,
                                          bool reaches_eof, uint64_t rec_idx, StagedRecord &out) {
    if (rec_idx >= a.rec_limit) return false;
    SerialRecord r;
    uint32_t b = p0, e, nb;
    bool bad = false;
    auto int_field = [&](int &v) {
        b = nb;
        token_span<WIN>(mask, b, e, nb);
        const bool ok = int_token(s, b, e, v);
        bad |= !ok && e < static_cast<uint32_t>(WIN);
    };
    token_span<WIN>(mask, b, e, nb);
    const uint32_t qb = b, qe = e;
    int_field(r.qlen); int_field(r.qs); int_field(r.qe);
    b = nb;
    token_span<WIN>(mask, b, e, nb);
    r.strand = s[b];
    bad |= (e - b != 1u) && e < static_cast<uint32_t>(WIN);
    b = nb;
    token_span<WIN>(mask, b, e, nb);
    const uint32_t tb = b, te = e;
    int_field(r.tlen); int_field(r.ts); int_field(r.te); int_field(r.matches); int_field(r.block); int_field(r.mapq);
    int status = bad ? 0 : (e >= static_cast<uint32_t>(WIN) ? 2 : 1);   // 1 regular, 0 irregular, 2 window exhausted
    if (status == 1) {
        r.qoff = tile_byte0 + qb; r.qnlen = qe - qb; r.qhash = hash_name_window(s, qb, qe);
        r.toff = tile_byte0 + tb; r.tnlen = te - tb; r.thash = hash_name_window(s, tb, te);
    } else if (status == 2) {
        if (reaches_eof) return false;  // trailing record with fewer than 12 tokens: not a record
        SerialRecord exact;   // (its address goes to a non-inlined call: keep it apart from r, which lives in registers)
        status = slow_path_record(a, tile_byte0 + p0, exact);
        r = exact;
    }
    if (status == 0) {
        atomicMin(&a.ctr->first_bad_rec, static_cast<unsigned long long>(rec_idx));
        atomicMin(&a.ctr->first_bad_off, static_cast<unsigned long long>(tile_byte0 + p0));
        return false;
    }
    return make_record(a.filter, a.text, a.file_id, a.anchors != 0, r, static_cast<uint32_t>(rec_idx), out, a.ctr, a.name_off0);
}

// Position of the k-th (0-based) set bit of a mask that has more than k bits set.
__device__ __forceinline__ uint32_t select_bit32(uint32_t m, uint32_t k) {
    uint32_t pos = 0, c;
    c = __popc(m & 0xffffu); if (k >= c) { k -= c; pos = 16; m >>= 16; }
    c = __popc(m & 0xffu);   if (k >= c) { k -= c; pos += 8; m >>= 8; }
    c = __popc(m & 0xfu);    if (k >= c) { k -= c; pos += 4; m >>= 4; }
    c = __popc(m & 0x3u);    if (k >= c) { k -= c; pos += 2; m >>= 2; }
    c = m & 1u;              if (k >= c) pos += 1;
    return pos;
}

// 4 bits: byte b of w is not white space (SWAR, no per-byte loop).
__device__ __forceinline__ uint32_t nonspace4(uint32_t w) {
    const uint32_t low7 = w & 0x7f7f7f7fu;
    const uint32_t x = w ^ 0x20202020u;
    const uint32_t is20 = ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u;          // byte == 0x20
    const uint32_t ge9 = low7 + 0x77777777u, ge14 = low7 + 0x72727272u;                      // bit 7: low7 >= 9 / >= 14
    const uint32_t ctl = ge9 & ~ge14 & ~w & 0x80808080u;                                     // 9 <= byte <= 13
    const uint32_t ns = ~(is20 | ctl) & 0x80808080u;
    return (((ns >> 7) * 0x00204081u) >> 21) & 0xfu;
}

// Shared memory of one staged tile: text window with 16 spare bytes on both sides (the field readers load whole
// words around a token), then the non-space bit per byte: chunk ch at nonws[ch + 2] (so that 32-byte units are
// aligned words), bit 15 of nonws[1] = the byte before the tile, two zero words of padding behind the window.
constexpr int PF_RAW_BYTES = PF_WIN + 32;
constexpr int PF_BUF_BYTES = (PF_RAW_BYTES + (PF_CHUNKS + 2 + 4) * 2 + 15) / 16 * 16;
constexpr int PF_SMEM_BYTES = 2 * PF_BUF_BYTES;
static_assert(PF_RAW_BYTES % 16 == 0, "mask words must stay aligned");

struct TileView {
    uint8_t *raw;        // PF_RAW_BYTES
    uint16_t *nonws;     // PF_CHUNKS + 6
    uint32_t mbar;       // shared-memory address of the buffer's mbarrier (bulk copies complete on it)
    __device__ __forceinline__ uint8_t *text() const { return raw + 16; }
    __device__ __forceinline__ const uint32_t *mask() const { return reinterpret_cast<const uint32_t *>(nonws) + 1; }  // word i = bytes [32 i, 32 i + 32)
};

// Token starts of the PF_WPT mask words (32 bytes each) thread tid owns: consecutive, so thread order = text order.
__device__ __forceinline__ uint32_t token_starts(const ParseArgs &a, const uint32_t *mask, uint64_t tile_byte0, uint32_t tid, uint32_t (&starts)[PF_WPT]) {
    uint32_t carry = mask[static_cast<int>(tid * PF_WPT) - 1] >> 31;   // is the byte before my stretch part of a token?
#pragma unroll
    for (int k = 0; k < PF_WPT; k++) {
        const uint32_t m = mask[tid * PF_WPT + k];
        starts[k] = m & ~((m << 1) | carry);
        carry = m >> 31;
    }
    if (tile_byte0 + PF_TILE > a.owned) {   // only the tile that holds the end of this rank's byte range
#pragma unroll
        for (int k = 0; k < PF_WPT; k++) {
            const uint64_t b0 = tile_byte0 + static_cast<uint64_t>(tid * PF_WPT + k) * 32;
            if (b0 + 32 > a.owned) starts[k] = b0 >= a.owned ? 0u : (starts[k] & ((1u << static_cast<uint32_t>(a.owned - b0)) - 1u));
        }
    }
    uint32_t n = 0;
#pragma unroll
    for (int k = 0; k < PF_WPT; k++) n += __popc(starts[k]);
    return n;
}

// A tile whose window lies inside the readable text is staged with ONE bulk copy (cp.async.bulk, the TMA unit: UBLKCP in
// SASS) issued by one thread and completing on the buffer's mbarrier; only the last tiles of a caller's buffer
// (hera_graph_load_device_view: nothing behind byte n may be read) fall back to per-thread 16-byte cp.async with zero fill.
__device__ __forceinline__ bool tile_is_bulk(const ParseArgs &a, uint32_t tile) {
    return !(a.bare && static_cast<uint64_t>(tile) * PF_TILE + PF_WIN > a.n);
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
    } while (!done);
}

// Stage A of a tile, first half: asynchronous copy of tile + halo into shared memory (the data is looked at later,
// after the CTA has done other work).  Called by all threads after a barrier: nobody reads the buffer any more.
__device__ __forceinline__ void stage_issue(const ParseArgs &a, uint32_t tile, const TileView &tv) {
    const uint64_t tile_byte0 = static_cast<uint64_t>(tile) * PF_TILE;
    const uint8_t *src = a.text + tile_byte0;
    const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(tv.text()));
    if (tile_is_bulk(a, tile)) {
        if (threadIdx.x == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the generic-proxy reads of the old tile come before the async-proxy writes
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tv.mbar), "r"(static_cast<uint32_t>(PF_WIN)) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst), "l"(src), "r"(static_cast<uint32_t>(PF_WIN)), "r"(tv.mbar) : "memory");
        }
        return;
    }
    if (a.bare && tile_byte0 + PF_WIN > a.n) {
        // the window reaches behind the caller's buffer: chunks that start behind byte n are not read at all (zero filled)
        for (int ch = threadIdx.x; ch < PF_CHUNKS; ch += PF_THREADS) {
            const uint64_t off = tile_byte0 + static_cast<uint64_t>(ch) * 16;
            const uint32_t take = off < a.n ? 16u : 0u;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + ch * 16), "l"(take ? src + static_cast<size_t>(ch) * 16 : a.text), "r"(take) : "memory");
        }
    } else {
        for (int ch = threadIdx.x; ch < PF_CHUNKS; ch += PF_THREADS)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + ch * 16), "l"(src + static_cast<size_t>(ch) * 16) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

// Second half, part one: byte classification (every thread reads back the chunks it copied).  The mask may be
// read by other threads after the next barrier.
__device__ __forceinline__ void stage_classify(const ParseArgs &a, uint32_t tile, const TileView &tv, uint32_t parity) {
    const uint32_t tid = threadIdx.x;
    const uint64_t tile_byte0 = static_cast<uint64_t>(tile) * PF_TILE;
    if (tile_is_bulk(a, tile)) mbar_wait(tv.mbar, parity);   // the whole window has landed (any thread may read any chunk)
    else asm volatile("cp.async.wait_group 0;" ::: "memory");
    const uint8_t *s_text = tv.text();
    if (a.bare && tile_byte0 + PF_WIN > a.n) {
        // bytes at and behind n (whatever the last chunk of the caller's buffer holds there, zeros behind it) are white space
        for (int ch = tid; ch < PF_CHUNKS; ch += PF_THREADS) {
            const uint4 v = reinterpret_cast<const uint4 *>(s_text)[ch];
            uint32_t m = nonspace4(v.x) | (nonspace4(v.y) << 4) | (nonspace4(v.z) << 8) | (nonspace4(v.w) << 12);
            const uint64_t off = tile_byte0 + static_cast<uint64_t>(ch) * 16;
            if (off + 16 > a.n) m = off >= a.n ? 0u : (m & ((1u << static_cast<uint32_t>(a.n - off)) - 1u));
            tv.nonws[ch + 2] = static_cast<uint16_t>(m);
        }
    } else {
#pragma unroll 2
        for (int ch = tid; ch < PF_CHUNKS; ch += PF_THREADS) {
            const uint4 v = reinterpret_cast<const uint4 *>(s_text)[ch];
            tv.nonws[ch + 2] = static_cast<uint16_t>(nonspace4(v.x) | (nonspace4(v.y) << 4) | (nonspace4(v.z) << 8) | (nonspace4(v.w) << 12));
        }
    }
    if (tid == 0) {
        uint32_t prev_nonws = a.prev_nonws ? 1u : 0u;  // does the byte before the tile belong to a token?
        if (tile_byte0 > 0) prev_nonws = dev_is_space(a.text[tile_byte0 - 1]) ? 0u : 1u;
        tv.nonws[1] = static_cast<uint16_t>(prev_nonws << 15);
        for (int k = 0; k < 4; k++) tv.nonws[PF_CHUNKS + 2 + k] = 0;
        for (int k = 0; k < 4; k++) { reinterpret_cast<uint32_t *>(tv.raw)[k] = 0x20202020u; reinterpret_cast<uint32_t *>(tv.raw + 16 + PF_WIN)[k] = 0x20202020u; }
    }
}

// Block-wide exclusive scan with ONE barrier: every warp leaves its total in s_tot, every thread adds up the
// totals of the warps before it.  The caller keeps a barrier between two scans that use the same s_tot.
constexpr int PF_WARPS = PF_THREADS / 32;
static_assert(PF_WARPS <= 32, "one lane per warp total");
__device__ __forceinline__ uint32_t scan_part1(uint32_t v, uint32_t *s_tot) {   // returns the inclusive prefix inside the warp
    const uint32_t lane = lane_id();
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= static_cast<uint32_t>(d)) incl += up;
    }
    if (lane == 31) s_tot[warp_id()] = incl;
    return incl;
}
__device__ __forceinline__ uint32_t scan_part2(const uint32_t *s_tot, uint32_t *total) {   // after the barrier: sum of the warps before mine
    const uint32_t lane = lane_id(), warp = warp_id();
    const uint32_t t = lane < PF_WARPS ? s_tot[lane] : 0u;
    *total = __reduce_add_sync(0xffffffffu, t);
    return __reduce_add_sync(0xffffffffu, lane < warp ? t : 0u);
}

struct TileFrame {   // what the record stage needs to know about a staged tile (same in every thread)
    unsigned long long tok_base;
    uint32_t first_rec_ord, nrec;
};

// Stage C of a tile whose chain 1 has been resolved (base = tokens before the tile in this text): note the
// offset of every record that starts in the tile (every 16th token).  s_rec_off is complete after the next barrier.
__device__ __forceinline__ TileFrame frame_tile(const ParseArgs &a, uint32_t tile, const TileView &tv, const uint16_t *s_excl, uint32_t tile_tokens,
                                                unsigned long long base, uint16_t *s_rec_off) {
    const uint32_t tid = threadIdx.x;
    const uint64_t tile_byte0 = static_cast<uint64_t>(tile) * PF_TILE;
    TileFrame f;
    f.tok_base = (a.tok_base0_dev ? *a.tok_base0_dev : a.tok_base0) + base;
    f.first_rec_ord = static_cast<uint32_t>((16 - (f.tok_base & 15)) & 15);
    f.nrec = tile_tokens > f.first_rec_ord ? (tile_tokens - f.first_rec_ord + 15) / 16 : 0;  // records starting in this tile
    uint32_t starts[PF_WPT];
    const uint32_t my_tokens = token_starts(a, tv.mask(), tile_byte0, tid, starts);
    const uint32_t excl = s_excl[tid];
    // a stretch of 32 PF_WPT bytes rarely holds more than one record start
    for (uint32_t ord = static_cast<uint32_t>((16 - ((f.tok_base + excl) & 15)) & 15); ord < my_tokens; ord += 16) {
        uint32_t r = ord, word = starts[0], k = 0;
#pragma unroll
        for (int j = 1; j < PF_WPT; j++) {
            const uint32_t cnt = __popc(starts[j - 1]);
            if (k == static_cast<uint32_t>(j - 1) && r >= cnt) { r -= cnt; word = starts[j]; k = j; }
        }
        s_rec_off[(excl + ord - f.first_rec_ord) >> 4] = static_cast<uint16_t>((tid * PF_WPT + k) * 32 + select_bit32(word, r));
    }
    return f;
}

// Persistent CTAs; tiles are handed out by an atomic ticket (so a tile only ever depends on tiles that are
// already running).  A CTA keeps two tiles staged and orders its work so that both chains are looked up long
// after every earlier tile has published, and the copy of a new tile is in flight while other work is done:
//     records of cur: parse, filter, score  ->  publish kept count (cur), take a ticket
//     ->  start the asynchronous copy of the new tile into cur's buffer  ->  classify it, count its tokens
//     ->  warp 0: publish the new tile's token count, resolve chain 1 of next and chain 2 of cur
//     ->  record offsets of next, write cur's kept records  ->  cur = next
// Six barriers per tile.
__global__ void __launch_bounds__(PF_THREADS, PF_MIN_CTAS) paf_parse_kernel(const ParseArgs a) {
    extern __shared__ __align__(16) uint8_t s_dyn[];
    __shared__ uint16_t s_rec_off[PF_MAXREC];
    __shared__ uint16_t s_excl[2][PF_THREADS];
    __shared__ uint32_t s_tot_tok[PF_WARPS], s_tot_kept[PF_WARPS];
    __shared__ uint32_t s_tokens[2];
    __shared__ unsigned int s_ticket;
    __shared__ unsigned long long s_tok_base, s_kept_base;
    __shared__ __align__(8) unsigned long long s_mbar[2];

    const uint32_t tid = threadIdx.x, lane = lane_id(), warp = warp_id();
    TileView tv[2];
#pragma unroll
    for (int p = 0; p < 2; p++) {
        tv[p].raw = s_dyn + p * PF_BUF_BYTES;
        tv[p].nonws = reinterpret_cast<uint16_t *>(s_dyn + p * PF_BUF_BYTES + PF_RAW_BYTES);
        tv[p].mbar = static_cast<uint32_t>(__cvta_generic_to_shared(&s_mbar[p]));
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tv[0].mbar) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tv[1].mbar) : "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    uint32_t phases = 0;   // bit q: parity of the next bulk copy into buffer q (same in every thread)
    // Classifies the copied tile `t` in buffer q, counts its tokens and lets warp 0 publish the count; warp 0 also
    // resolves chain 1 of tile `n` (-> s_tok_base) and chain 2 of tile `c` with aggregate kc (-> s_kept_base) while
    // the other warps wait at the closing barrier.  Tiles >= a.tiles are skipped.
    auto stage_and_service = [&](uint32_t t, int q, uint32_t n, uint32_t n_tokens, uint32_t c, uint32_t kc) {
        uint32_t my_tokens = 0, incl = 0;
        if (t < a.tiles) {
            stage_classify(a, t, tv[q], (phases >> q) & 1u);
            if (tile_is_bulk(a, t)) phases ^= 1u << q;
            __syncthreads();
            uint32_t starts[PF_WPT];
            my_tokens = token_starts(a, tv[q].mask(), static_cast<uint64_t>(t) * PF_TILE, tid, starts);
            incl = scan_part1(my_tokens, s_tot_tok);
        }
        __syncthreads();
        if (t < a.tiles) {
            uint32_t total;
            const uint32_t before = scan_part2(s_tot_tok, &total);
            s_excl[q][tid] = static_cast<uint16_t>(before + incl - my_tokens);
            if (tid == 0) { s_tokens[q] = total; chain_publish(a.tok_state, t, total); }
        }
        if (warp == 0) {
            // both first look-back windows are requested before either is used (one round trip to L2, not two)
            const unsigned long long probe_tok = n < a.tiles ? chain_probe(a.tok_state, n) : 0ull;
            const unsigned long long probe_kept = c < a.tiles ? chain_probe(a.kept_state, c) : 0ull;
            if (n < a.tiles) {
                const unsigned long long base = chain_resolve(a.tok_state, n, n_tokens, probe_tok);
                if (lane == 0) {
                    s_tok_base = base;
                    if (n == a.tiles - 1) a.ctr->total_tokens = base + n_tokens;
                }
            }
            if (c < a.tiles) {
                const unsigned long long base = chain_resolve(a.kept_state, c, kc, probe_kept);
                if (lane == 0) {
                    s_kept_base = base + (a.kept_base_dev ? *a.kept_base_dev : 0ull);
                    if (c == a.tiles - 1) a.ctr->kept = base + kc;
                }
            }
        }
        __syncthreads();
    };

    // Tickets are taken one at a time: tiles must reach the record stage in roughly ticket order everywhere, or a
    // CTA's second tile (handled an iteration later) would hold up every CTA whose first tile comes after it.
    if (tid == 0) s_ticket = atomicAdd(&a.ctr->ticket, 1u);
    __syncthreads();
    uint32_t tile = s_ticket;
    if (tile >= a.tiles) return;
    stage_issue(a, tile, tv[0]);
    __syncthreads();
    if (tid == 0) s_ticket = atomicAdd(&a.ctr->ticket, 1u);
    __syncthreads();
    uint32_t next = s_ticket;
    if (next < a.tiles) stage_issue(a, next, tv[1]);
    stage_and_service(tile, 0, a.tiles, 0, a.tiles, 0);
    stage_and_service(next, 1, tile, s_tokens[0], a.tiles, 0);
    TileFrame f = frame_tile(a, tile, tv[0], s_excl[0], s_tokens[0], s_tok_base, s_rec_off);
    __syncthreads();
    int p = 0;
    // The kept rows of a tile are written ONE ITERATION AFTER the tile was parsed: its count is published to chain 2 at once,
    // but the chain is resolved (and the rows stored at their file-order place) only while the next tile's successor is
    // staged.  Resolving in the same iteration left ~5 us between a tile's publication and its successor's look-up; the
    // CTAs of an SM slow each other down by more than that, so on a long text almost every look-up met a predecessor
    // that had not published yet and 57 % of all warp samples sat at the barrier behind warp 0's spin (ncu, configs[3]).
    // The parked row lives in registers / local memory (the compiler's choice) for the length of one iteration.
    uint32_t prev_tile = a.tiles, prev_kept = 0, e_prev = 0;
    bool keep_prev = false;
    StagedRecord mine_prev;

    while (tile < a.tiles) {
        const uint64_t tile_byte0 = static_cast<uint64_t>(tile) * PF_TILE;
        const uint8_t *s_text = tv[p].text();
        const uint32_t *mask = tv[p].mask();
        const bool reaches_eof = a.at_eof && tile_byte0 + PF_WIN >= a.n;
        const uint32_t nrec = f.nrec;
        const bool rounds = nrec > PF_THREADS;

        // one thread per record: parse, filter, score.  Tiles with more than PF_THREADS records (tiny records) take
        // further rounds; their records are parsed again after the chain instead of being kept in registers.
        const uint64_t first_rec = (f.tok_base + f.first_rec_ord) >> 4;
        StagedRecord mine;
        bool keep0 = false;
        uint32_t e0 = 0, tot0 = 0, tile_kept = 0;
        for (uint32_t b0 = 0; b0 < nrec; b0 += PF_THREADS) {
            const uint32_t j = b0 + tid;
            StagedRecord r;
            bool keep = false;
            if (j < nrec) keep = parse_one<PF_WIN>(a, s_text, mask, s_rec_off[j], tile_byte0, reaches_eof, first_rec + j, r);
            const uint32_t votes = __ballot_sync(0xffffffffu, keep);
            if (lane == 31) s_tot_kept[warp] = __popc(votes);
            __syncthreads();
            uint32_t tot;
            const uint32_t ex = scan_part2(s_tot_kept, &tot) + __popc(votes & ((1u << lane) - 1u));
            if (b0 == 0) { mine = r; keep0 = keep; e0 = ex; tot0 = tot; }
            tile_kept += tot;
            if (rounds) __syncthreads();   // s_tot_kept is written again right away
        }
        if (tid == 0) {
            chain_publish(a.kept_state, tile, tile_kept);
            s_ticket = atomicAdd(&a.ctr->ticket, 1u);
        }

        if (rounds) {   // rare: this tile still needs its text and record offsets, finish it before anything else
            if (warp == 0) {
                const unsigned long long base = chain_resolve(a.kept_state, tile, tile_kept, chain_probe(a.kept_state, tile));
                if (lane == 0) {
                    s_kept_base = base + (a.kept_base_dev ? *a.kept_base_dev : 0ull);
                    if (tile == a.tiles - 1) a.ctr->kept = base + tile_kept;
                }
            }
            __syncthreads();
            const uint64_t out0 = a.kept_base + s_kept_base;
            if (keep0) store_record(a.kept, out0 + e0, mine);
            uint32_t run = tot0;
            for (uint32_t b0 = PF_THREADS; b0 < nrec; b0 += PF_THREADS) {
                const uint32_t j = b0 + tid;
                StagedRecord r;
                bool keep = false;
                if (j < nrec) keep = parse_one<PF_WIN>(a, s_text, mask, s_rec_off[j], tile_byte0, reaches_eof, first_rec + j, r);
                const uint32_t votes = __ballot_sync(0xffffffffu, keep);
                if (lane == 31) s_tot_kept[warp] = __popc(votes);
                __syncthreads();
                uint32_t tot;
                const uint32_t ex = scan_part2(s_tot_kept, &tot) + __popc(votes & ((1u << lane) - 1u));
                if (keep) store_record(a.kept, out0 + run + ex, r);
                run += tot;
                __syncthreads();
            }
        }

        __syncthreads();   // everybody is done with cur's text and record offsets; the ticket is visible
        const uint32_t after = s_ticket;
        if (after < a.tiles) stage_issue(a, after, tv[p]);
        stage_and_service(after, p, next, s_tokens[p ^ 1], prev_tile, prev_kept);   // chain 2 of the tile parsed one iteration ago
        TileFrame fn{};
        if (next < a.tiles) fn = frame_tile(a, next, tv[p ^ 1], s_excl[p ^ 1], s_tokens[p ^ 1], s_tok_base, s_rec_off);
        if (keep_prev) store_record(a.kept, a.kept_base + s_kept_base + e_prev, mine_prev);
        __syncthreads();   // record offsets of next are complete; s_kept_base has been read
        prev_tile = rounds ? a.tiles : tile;
        prev_kept = tile_kept;
        keep_prev = !rounds && keep0;
        e_prev = e0;
        mine_prev = mine;
        tile = next;
        next = after;
        f = fn;
        p ^= 1;
    }
    // the rows of the last tile this CTA parsed
    if (warp == 0 && prev_tile < a.tiles) {
        const unsigned long long base = chain_resolve(a.kept_state, prev_tile, prev_kept, chain_probe(a.kept_state, prev_tile));
        if (lane == 0) {
            s_kept_base = base + (a.kept_base_dev ? *a.kept_base_dev : 0ull);
            if (prev_tile == a.tiles - 1) a.ctr->kept = base + prev_kept;
        }
    }
    __syncthreads();
    if (keep_prev) store_record(a.kept, a.kept_base + s_kept_base + e_prev, mine_prev);
}

// ---- warp-autonomous variant (hera_set_option "parse_variant" = 1) -----------------------------------------------------
// Same algorithm, no CTA-wide barrier: every WARP takes its own 3 KiB sub-tile (+ 256 bytes of halo) by ticket, stages and
// classifies it in its own slice of shared memory, publishes / resolves both chains itself and parses one record per
// lane.  The CTA kernel above spends a third of its stall samples at barriers (warps of a CTA finish the record stage at
// different times); here a warp that waits -- for its copy, for a predecessor in a chain -- only stalls itself, the other
// warps of the SM keep the issue slots busy.  The chains have ten times more entries (one per sub-tile).  A record that
// does not end inside the halo (names of several hundred bytes) goes to the exact one-thread reader.
constexpr int WW_TILE = 3072, WW_HALO = 256, WW_WIN = WW_TILE + WW_HALO, WW_CHUNKS = WW_WIN / 16;
constexpr int WW_WARPS = 8, WW_THREADS = WW_WARPS * 32, WW_MIN_CTAS = 4;
constexpr int WW_WPL = WW_TILE / 32 / 32;                 // mask words (32 bytes each) per lane
static_assert(WW_WPL * 32 * 32 == WW_TILE, "sub-tile = 32 lanes x whole mask words");
constexpr int WW_MAXREC = WW_TILE / 32 + 2;
constexpr int WW_RAW_BYTES = WW_WIN + 32;
constexpr int WW_MASK_BYTES = ((WW_CHUNKS + 2 + 4) * 2 + 15) / 16 * 16;
constexpr int WW_REC_BYTES = (WW_MAXREC * 2 + 15) / 16 * 16;
constexpr int WW_BUF_BYTES = WW_RAW_BYTES + WW_MASK_BYTES + WW_REC_BYTES;
static_assert(WW_RAW_BYTES % 16 == 0 && WW_BUF_BYTES % 16 == 0, "per-warp buffers stay 16-byte aligned");

__global__ void __launch_bounds__(WW_THREADS, WW_MIN_CTAS) paf_parse_warp_kernel(const ParseArgs a) {
    __shared__ __align__(16) uint8_t s_buf[WW_WARPS * WW_BUF_BYTES];
    const uint32_t lane = lane_id(), warp = warp_id();
    uint8_t *raw = s_buf + warp * WW_BUF_BYTES;
    uint8_t *s_text = raw + 16;
    uint16_t *nonws = reinterpret_cast<uint16_t *>(raw + WW_RAW_BYTES);
    const uint32_t *mask = reinterpret_cast<const uint32_t *>(nonws) + 1;   // word i = bytes [32 i, 32 i + 32) of the sub-tile
    uint16_t *s_rec_off = reinterpret_cast<uint16_t *>(raw + WW_RAW_BYTES + WW_MASK_BYTES);
    const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(s_text));
    const unsigned long long tok_base0 = a.tok_base0_dev ? *a.tok_base0_dev : a.tok_base0;

    for (;;) {
        uint32_t tile = 0;
        if (lane == 0) tile = atomicAdd(&a.ctr->ticket, 1u);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= a.tiles) break;
        const uint64_t tile_byte0 = static_cast<uint64_t>(tile) * WW_TILE;
        const bool tail = a.bare && tile_byte0 + WW_WIN > a.n;   // the window reaches behind the caller's buffer
        __syncwarp();   // everybody is done with the previous sub-tile's buffers

        // 1. stage text + halo
        const uint8_t *src = a.text + tile_byte0;
        for (int ch = lane; ch < WW_CHUNKS; ch += 32) {
            const uint64_t off = tile_byte0 + static_cast<uint64_t>(ch) * 16;
            const uint32_t take = (!tail || off < a.n) ? 16u : 0u;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + ch * 16), "l"(take ? src + static_cast<size_t>(ch) * 16 : a.text), "r"(take) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        if (lane == 0) {
            uint32_t prev_nonws = a.prev_nonws ? 1u : 0u;  // does the byte before the sub-tile belong to a token?
            if (tile_byte0 > 0) prev_nonws = dev_is_space(a.text[tile_byte0 - 1]) ? 0u : 1u;
            nonws[1] = static_cast<uint16_t>(prev_nonws << 15);
            for (int k = 0; k < 4; k++) nonws[WW_CHUNKS + 2 + k] = 0;
            for (int k = 0; k < 4; k++) { reinterpret_cast<uint32_t *>(raw)[k] = 0x20202020u; reinterpret_cast<uint32_t *>(raw + 16 + WW_WIN)[k] = 0x20202020u; }
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        // 2. classify the chunks this lane copied
        for (int ch = lane; ch < WW_CHUNKS; ch += 32) {
            const uint4 v = reinterpret_cast<const uint4 *>(s_text)[ch];
            uint32_t m = nonspace4(v.x) | (nonspace4(v.y) << 4) | (nonspace4(v.z) << 8) | (nonspace4(v.w) << 12);
            if (tail) {
                const uint64_t off = tile_byte0 + static_cast<uint64_t>(ch) * 16;
                if (off + 16 > a.n) m = off >= a.n ? 0u : (m & ((1u << static_cast<uint32_t>(a.n - off)) - 1u));
            }
            nonws[ch + 2] = static_cast<uint16_t>(m);
        }
        __syncwarp();

        // 3. token starts of this lane's mask words (consecutive: lane order = text order)
        uint32_t starts[WW_WPL];
        uint32_t carry = mask[static_cast<int>(lane * WW_WPL) - 1] >> 31;
#pragma unroll
        for (int k = 0; k < WW_WPL; k++) {
            const uint32_t m = mask[lane * WW_WPL + k];
            starts[k] = m & ~((m << 1) | carry);
            carry = m >> 31;
        }
        if (tile_byte0 + WW_TILE > a.owned) {   // only the sub-tile that holds the end of this rank's byte range
#pragma unroll
            for (int k = 0; k < WW_WPL; k++) {
                const uint64_t b0 = tile_byte0 + static_cast<uint64_t>(lane * WW_WPL + k) * 32;
                if (b0 + 32 > a.owned) starts[k] = b0 >= a.owned ? 0u : (starts[k] & ((1u << static_cast<uint32_t>(a.owned - b0)) - 1u));
            }
        }
        uint32_t my_tokens = 0;
#pragma unroll
        for (int k = 0; k < WW_WPL; k++) my_tokens += __popc(starts[k]);
        uint32_t incl = my_tokens;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= static_cast<uint32_t>(d)) incl += up;
        }
        const uint32_t tile_tokens = __shfl_sync(0xffffffffu, incl, 31), excl = incl - my_tokens;

        // 4. chain 1: tokens before this sub-tile
        if (lane == 0) chain_publish(a.tok_state, tile, tile_tokens);
        const unsigned long long tbase = chain_resolve_wide<16>(a.tok_state, tile, tile_tokens);
        if (lane == 0 && tile == a.tiles - 1) a.ctr->total_tokens = tbase + tile_tokens;
        const unsigned long long tok_base = tok_base0 + tbase;

        // 5. offsets of the records that start here (every 16th token)
        const uint32_t first_rec_ord = static_cast<uint32_t>((16 - (tok_base & 15)) & 15);
        const uint32_t nrec = tile_tokens > first_rec_ord ? (tile_tokens - first_rec_ord + 15) / 16 : 0;
        for (uint32_t ord = static_cast<uint32_t>((16 - ((tok_base + excl) & 15)) & 15); ord < my_tokens; ord += 16) {
            uint32_t r = ord, word = starts[0], k = 0;
#pragma unroll
            for (int j = 1; j < WW_WPL; j++) {
                const uint32_t cnt = __popc(starts[j - 1]);
                if (k == static_cast<uint32_t>(j - 1) && r >= cnt) { r -= cnt; word = starts[j]; k = j; }
            }
            s_rec_off[(excl + ord - first_rec_ord) >> 4] = static_cast<uint16_t>((lane * WW_WPL + k) * 32 + select_bit32(word, r));
        }
        __syncwarp();

        // 6. one lane per record: parse, filter, score (further rounds only for tiny records; those are parsed again below)
        const bool reaches_eof = a.at_eof && tile_byte0 + WW_WIN >= a.n;
        const uint64_t first_rec = (tok_base + first_rec_ord) >> 4;
        StagedRecord mine;
        bool keep0 = false;
        uint32_t e0 = 0, tot0 = 0, tile_kept = 0;
        for (uint32_t b0 = 0; b0 < nrec; b0 += 32) {
            const uint32_t j = b0 + lane;
            StagedRecord r;
            bool keep = false;
            if (j < nrec) keep = parse_one<WW_WIN>(a, s_text, mask, s_rec_off[j], tile_byte0, reaches_eof, first_rec + j, r);
            const uint32_t votes = __ballot_sync(0xffffffffu, keep);
            if (b0 == 0) { mine = r; keep0 = keep; e0 = __popc(votes & ((1u << lane) - 1u)); tot0 = __popc(votes); }
            tile_kept += __popc(votes);
        }

        // 7. chain 2: kept records before this sub-tile, then the rows go to their final place
        if (lane == 0) chain_publish(a.kept_state, tile, tile_kept);
        const unsigned long long kbase = chain_resolve_wide<16>(a.kept_state, tile, tile_kept);
        if (lane == 0 && tile == a.tiles - 1) a.ctr->kept = kbase + tile_kept;
        const uint64_t out0 = a.kept_base + kbase;
        if (keep0) store_record(a.kept, out0 + e0, mine);
        uint32_t run = tot0;
        for (uint32_t b0 = 32; b0 < nrec; b0 += 32) {
            const uint32_t j = b0 + lane;
            StagedRecord r;
            bool keep = false;
            if (j < nrec) keep = parse_one<WW_WIN>(a, s_text, mask, s_rec_off[j], tile_byte0, reaches_eof, first_rec + j, r);
            const uint32_t votes = __ballot_sync(0xffffffffu, keep);
            if (keep) store_record(a.kept, out0 + run + __popc(votes & ((1u << lane) - 1u)), r);
            run += __popc(votes);
        }
    }
}

// More rows were kept than the table has room for (optimistic reservation): marked in the top bit of the count.
constexpr unsigned long long KEPT_OVERFLOW = 1ull << 63;
__global__ void kept_overflow_kernel(ParseCounters *ctr, uint64_t room) {
    if (ctr->kept > room) ctr->kept |= KEPT_OVERFLOW;
}

// ---- chunked load: a host text is copied in chunks and every chunk is parsed by its own launch as soon as it (and the
// head of its successor) has landed, so the parse hides behind the host-to-device copy.  A chunk is what a rank's slice is
// in a sharded load: it starts at an arbitrary byte, counts the tokens that START inside it and completes its last record
// from the bytes behind it.  Where a slice learns the tokens in front of it from the other ranks, a chunk reads them -- and
// the rows kept so far -- from the `carry` words the previous chunk's launch left behind.
__global__ void chunk_init_kernel(ParseCounters *part, uint32_t chunks) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= chunks) return;
    part[k].first_bad_rec = ~0ull;
    part[k].first_bad_off = ~0ull;
}
// merged: the counters of the whole load (what an unchunked launch leaves); carry[0] tokens, carry[1] kept rows so far
__global__ void chunk_carry_kernel(ParseCounters *merged, const ParseCounters *part, unsigned long long *carry, uint64_t byte0) {
    carry[0] += part->total_tokens;
    carry[1] += part->kept;
    merged->total_tokens = carry[0];
    merged->kept = carry[1];
    if (part->first_bad_rec < merged->first_bad_rec) {   // record indices are global; chunks come in file order
        merged->first_bad_rec = part->first_bad_rec;
        merged->first_bad_off = part->first_bad_off + byte0;
    }
    if (part->name_too_long) merged->name_too_long = 1;
}

// Finds how many kept records of this load precede the first irregular record.
__global__ void kept_before_kernel(const uint32_t *__restrict__ rec, uint64_t base, uint64_t n, uint64_t bad_rec, unsigned long long *out) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bool below = rec[base + i] < bad_rec;
    const bool next_below = (i + 1 < n) && rec[base + i + 1] < bad_rec;
    if (below && !next_below) *out = i + 1;
}

// One thread continues with the exact stream rules from the first irregular record.
__global__ void paf_tail_kernel(const uint8_t *text, uint64_t n, uint64_t start_off, uint64_t first_rec, uint64_t rec_limit, uint32_t file_id,
                                int anchors, FilterDev filter, KeptPtrs kept, uint64_t kept_pos, uint64_t kept_cap, ParseCounters *ctr) {
    SerialCursor cur{text, n, start_off, true};
    uint64_t rec = first_rec, nk = 0;
    while (rec < rec_limit) {
        SerialRecord r;
        if (!serial_read_record(cur, r)) break;
        StagedRecord s;
        if (make_record(filter, text, file_id, anchors != 0, r, static_cast<uint32_t>(rec), s, ctr)) {
            if (kept_pos + nk < kept_cap) store_record(kept, kept_pos + nk, s);
            nk++;
        }
        rec++;
    }
    ctr->tail_records = rec - first_rec;
    ctr->tail_kept = nk;
}

// Tokens that start in text[0, owned): what a rank of a sharded load tells the others before anyone
// can know where its records begin (record boundary = global token index % 16 == 0).
__global__ void __launch_bounds__(256) paf_count_tokens_kernel(const uint8_t *__restrict__ text, uint64_t owned, int prev_nonws, unsigned long long *out) {
    const uint64_t chunks = (owned + 15) / 16;
    unsigned long long mine = 0;
    for (uint64_t ch = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; ch < chunks; ch += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(text) + ch);
        const uint32_t cur = nonspace4(v.x) | (nonspace4(v.y) << 4) | (nonspace4(v.z) << 8) | (nonspace4(v.w) << 12);
        const uint32_t prev = ch ? (dev_is_space(text[ch * 16 - 1]) ? 0u : 1u) : (prev_nonws ? 1u : 0u);
        uint32_t starts = cur & ~((cur << 1) | prev) & 0xffffu;
        const uint64_t b0 = ch * 16;
        if (b0 + 16 > owned) starts &= (1u << static_cast<uint32_t>(owned - b0)) - 1u;
        mine += __popc(starts);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, d);
    if (lane_id() == 0 && mine) atomicAdd(out, mine);
}

// tokens before rank `me`'s slice and tokens of all slices, from the gathered per-rank counts
__global__ void shard_prefix_kernel(const uint64_t *__restrict__ counts, int ranks, int me, unsigned long long *out) {
    if (threadIdx.x) return;
    unsigned long long before = 0, all = 0;
    for (int r = 0; r < ranks; r++) { if (r < me) before += counts[r]; all += counts[r]; }
    out[0] = before;
    out[1] = all;
}

__global__ void pad_text_kernel(uint8_t *text, uint64_t n, uint64_t padded) {
    const uint64_t i = n + static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < padded) text[i] = '\n';
}

static KeptPtrs kept_ptrs(KeptTable &k) {
    return KeptPtrs{k.qs.p, k.qe.p, k.ts.p, k.te.p, k.qlen.p, k.tlen.p, k.rec.p, k.os.p, k.si.p, k.es.p, k.flags.p, k.qhash.p, k.thash.p, k.qname.p, k.tname.p,
                    k.qs.cap};
}

uint64_t padded_text_size(uint64_t n) { return (n / PF_TILE + 2) * static_cast<uint64_t>(PF_TILE); }

static_assert(LOAD_CHUNK_GRAIN == PF_TILE, "chunks of a host text are whole tiles");
static bool host_is_space(uint8_t ch) { return ch == ' ' || (ch >= 9 && ch <= 13); }

static double host_now_ms() {   // wall clock shared by the ranks of one box (HERA_DEBUG_TIMING)
    timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    return static_cast<double>(ts.tv_sec % 1000) * 1e3 + static_cast<double>(ts.tv_nsec) * 1e-6;
}

void load_text_resident(Ctx &c, LoadedText &text, uint32_t file_id, bool anchors, uint32_t test_load_num, const ShardSpec *shard, hera_load_result *out,
                        const HostSource *host) {
    const bool debug_timing = getenv("HERA_DEBUG_TIMING") != nullptr;
    const double t_entry = debug_timing ? host_now_ms() : 0.0;
    double t_gather = 0.0, t_launched = 0.0;
    // HERA_DEBUG_TIMING=2: synchronise after every phase and note the host clock (which phase does the host wait for?)
    const bool debug_phases = debug_timing && std::atoi(getenv("HERA_DEBUG_TIMING")) >= 2;
    std::string phase_log;
    auto mark = [&](const char *name) {
        if (!debug_phases) return;
        cudaStreamSynchronize(c.stream);
        char buf[64];
        snprintf(buf, sizeof buf, " %s +%.3f", name, host_now_ms() - t_entry);
        phase_log += buf;
    };
    cudaEvent_t ev_in = nullptr, ev_out = nullptr;   // (debug) device time in front of and behind the timed region
    if (debug_timing) {
        cudaEventCreate(&ev_in);
        cudaEventCreate(&ev_out);
        cudaEventRecord(ev_in, c.stream);
    }
    const uint64_t n = text.n;
    const uint64_t padded = padded_text_size(n);
    const bool bare = text.view != nullptr;
    const uint8_t *text_p = text.data();
    if (!bare) {
        HERA_REQUIRE(text.bytes.cap >= padded, 3, "internal: text buffer not padded");
        HERA_LAUNCH(c, pad_text_kernel, div_up(padded - n, 256), 256, 0, text.bytes.p, n, padded);
    }
    const bool sharded = shard != nullptr;
    const int ranks = xch_size(c), me = xch_rank(c);
    // The text is still in host memory: copy it on the copy stream, chunk by chunk, an event behind every chunk.
    struct Chunk { uint64_t off, owned; uint32_t tiles, tile0; cudaEvent_t landed; };
    std::vector<Chunk> chunks;
    struct ChunkEvents {
        std::vector<Chunk> &v;
        cudaEvent_t t0 = nullptr, t1 = nullptr;
        ~ChunkEvents() {
            for (Chunk &k : v) if (k.landed) cudaEventDestroy(k.landed);
            if (t0) cudaEventDestroy(t0);
            if (t1) cudaEventDestroy(t1);
        }
    } chunk_events{chunks};
    if (host) {
        HERA_REQUIRE(!bare && !sharded && host->chunk_bytes >= LOAD_CHUNK_MIN_GRAINS * PF_TILE && host->chunk_bytes % PF_TILE == 0, 3, "internal: chunked load of a view / a slice");
        if (!c.copy_stream) HERA_CUDA(cudaStreamCreateWithFlags(&c.copy_stream, cudaStreamNonBlocking));
        HERA_CUDA(cudaEventCreate(&chunk_events.t0));
        HERA_CUDA(cudaEventCreate(&chunk_events.t1));
        {   // the buffer may be one an earlier graph of this context was parsed from: the copy starts behind the stream's work
            Events<1> entry(cudaEventDisableTiming);
            HERA_CUDA(cudaEventRecord(entry[0], c.stream));
            HERA_CUDA(cudaStreamWaitEvent(c.copy_stream, entry[0], 0));
        }
        HERA_CUDA(cudaEventRecord(chunk_events.t0, c.copy_stream));
        uint32_t tile0 = 0;
        for (uint64_t off = 0; off < n; off += host->chunk_bytes) {
            Chunk k{off, std::min<uint64_t>(host->chunk_bytes, n - off), 0, tile0, nullptr};
            k.tiles = static_cast<uint32_t>(k.owned / PF_TILE + 1);
            tile0 += k.tiles;
            chunks.push_back(k);
            HERA_CUDA(cudaEventCreateWithFlags(&chunks.back().landed, cudaEventDisableTiming));
            HERA_CUDA(cudaMemcpyAsync(text.bytes.p + off, host->text + off, k.owned, cudaMemcpyHostToDevice, c.copy_stream));
            HERA_CUDA(cudaEventRecord(chunks.back().landed, c.copy_stream));
        }
        HERA_CUDA(cudaEventRecord(chunk_events.t1, c.copy_stream));
    }
    HERA_REQUIRE(!sharded || c.xch, 21, "hera_graph_load_shard needs a communicator (hera_comm_init_nccl / hera_comm_attach_local)");
    const uint64_t owned = sharded ? shard->owned : n;
    HERA_REQUIRE(owned <= n, 101, "shard: owned bytes exceed the bytes provided");

    const bool warp_variant = c.opt_parse_variant == 1;
    bool chunked = chunks.size() > 1 && !warp_variant;
    uint32_t tiles = static_cast<uint32_t>(owned / (warp_variant ? WW_TILE : PF_TILE) + 1);
    // worst case every record is kept: records <= bytes / 24 (12 one-byte tokens + separators).  For a large text that
    // bound would reserve far more rows than the file can hold records (42 GB of PAF -> 128 GB of rows), so the tokens
    // are counted first (one more pass over the text) and the rows are reserved for tokens / 16 records.
    // A context that has loaded a text of this kind before skips the counting pass: the rows it holds already are taken as
    // the reservation (a PAF record has > 100 bytes; owned / 160 rows is the least that is worth trying), the kernel drops
    // rows that do not fit, the overflow is noticed right after the kernel (on every rank of a sharded load: the mark
    // travels with the kept counts) and the load is done again with the count.
    uint64_t max_new = owned / 24 + 2;
    const uint64_t room_now = c.kept.qs.cap > c.kept.n ? c.kept.qs.cap - c.kept.n : 0;
    const uint64_t count_rows = static_cast<uint64_t>(c.opt_parse_count_rows);
    const bool optimistic = max_new > count_rows && !c.parse_force_count && room_now >= owned / 160 + 2;
    if (!chunks.empty() && !(chunked && (optimistic || max_new <= count_rows))) {
        // one chunk, or the tokens are counted first (a pass over the whole text): wait for the text, then as for a resident one
        HERA_CUDA(cudaStreamWaitEvent(c.stream, chunks.back().landed, 0));
        chunked = false;
    }
    if (chunked) tiles = chunks.back().tile0 + chunks.back().tiles;
    if (optimistic) {
        max_new = room_now;
    } else if (max_new > count_rows) {
        c.counters.reserve(16);
        unsigned long long *d_tok = reinterpret_cast<unsigned long long *>(c.counters.p + 13);
        HERA_CUDA(cudaMemsetAsync(d_tok, 0, sizeof(unsigned long long), c.stream));
        HERA_LAUNCH(c, paf_count_tokens_kernel, 148 * 8, 256, 0, text_p, owned, sharded && shard->prev_nonws ? 1 : 0, d_tok);
        c.h_counters.reserve(16);
        HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 13, d_tok, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        max_new = std::min<uint64_t>(max_new, c.h_counters.p[13] / 16 + 2);
    }
    if (!optimistic) c.kept.reserve(c.kept.n + max_new, c.stream);
    const uint64_t room = c.kept.qs.cap - c.kept.n;

    // device block: [0, 8) ParseCounters, [12] cut, [13] my token count, [14] tokens before my slice, [15] tokens of all
    // slices, [16, 16 + 2 tiles) chain states, then what the ranks tell each other: token counts, {kept, first bad} pairs
    const uint64_t gather_at = 16 + 2ull * tiles;
    const uint64_t parts_at = gather_at + 3ull * ranks + 8, part_words = sizeof(ParseCounters) / sizeof(uint64_t);   // per-chunk counters, then carry[2]
    static_assert(sizeof(ParseCounters) % sizeof(uint64_t) == 0, "counter blocks are laid out in 64-bit words");
    const uint64_t block_words = parts_at + (chunked ? chunks.size() * part_words + 2 : 0);
    c.counters.reserve(block_words);
    ParseCounters *ctr = reinterpret_cast<ParseCounters *>(c.counters.p);
    unsigned long long *tok_state = reinterpret_cast<unsigned long long *>(c.counters.p + 16);
    unsigned long long *kept_state = tok_state + tiles;
    HERA_CUDA(cudaMemsetAsync(c.counters.p, 0, (chunked ? block_words : gather_at + 3ull * ranks) * sizeof(uint64_t), c.stream));
    ParseCounters init{};
    init.first_bad_rec = ~0ull;
    init.first_bad_off = ~0ull;
    HERA_CUDA(cudaMemcpyAsync(ctr, &init, sizeof(init), cudaMemcpyHostToDevice, c.stream));
    c.h_counters.reserve(16 + 3ull * ranks);

    FilterDev fd{c.fp.mode, static_cast<float>(c.fp.min_overlap_length), c.fp.min_overlap_percentage,
                 static_cast<float>(c.fp.max_overhang_length), c.fp.max_overhang_percentage};
    const uint64_t rec_limit = test_load_num > 0 ? static_cast<uint64_t>(test_load_num) + 1 : ~0ull;

    Events<3> ev;   // destroyed when the scope ends, exception or not
    const cudaEvent_t ev0 = ev[0], ev1 = ev[1], ev2 = ev[2];
    maybe_flush_l2(c);
    mark("setup");
    HERA_CUDA(cudaEventRecord(ev0, c.stream));

    // sharded: where do my records start?  tokens before my slice = sum over the earlier ranks, exchanged and added
    // up on the device (no host round trip before the parse kernel)
    uint64_t *d_tok_all = c.counters.p + gather_at, *d_kb_all = d_tok_all + ranks;
    if (sharded) {
        unsigned long long *d_tok = reinterpret_cast<unsigned long long *>(c.counters.p + 13);
        if (owned) HERA_LAUNCH(c, paf_count_tokens_kernel, std::min<unsigned>(div_up(div_up(owned, 16), 256), 148 * 8), 256, 0, text_p, owned, shard->prev_nonws ? 1 : 0, d_tok);
        mark("count");
        if (debug_timing) t_gather = host_now_ms();
        c.xch->all_gather_dev(c, c.counters.p + 13, d_tok_all, 1);
        mark("gather1");
        HERA_LAUNCH(c, shard_prefix_kernel, 1, 32, 0, d_tok_all, ranks, me, reinterpret_cast<unsigned long long *>(c.counters.p + 14));
        HERA_CUDA(cudaEventRecord(ev2, c.stream));   // (reused below: only read here when HERA_DEBUG_TIMING is set)
    }

    ParseArgs a{text_p, bare ? 1 : 0, n, owned, 0, sharded ? reinterpret_cast<const unsigned long long *>(c.counters.p + 14) : nullptr, sharded && shard->prev_nonws ? 1 : 0, (!sharded || shard->at_eof) ? 1 : 0,
                tiles, file_id, anchors ? 1 : 0, rec_limit, c.kept.n, fd, kept_ptrs(c.kept), tok_state, kept_state, ctr};
    int sms = 148;
    HERA_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c.device));
    HERA_CUDA(cudaFuncSetAttribute(paf_parse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PF_SMEM_BYTES));   // per device: cheap, idempotent
    if (chunked) {
        ParseCounters *part = reinterpret_cast<ParseCounters *>(c.counters.p + parts_at);
        unsigned long long *carry = reinterpret_cast<unsigned long long *>(c.counters.p + parts_at + chunks.size() * part_words);
        HERA_LAUNCH(c, chunk_init_kernel, div_up(chunks.size(), 128), 128, 0, part, static_cast<uint32_t>(chunks.size()));
        for (size_t k = 0; k < chunks.size(); k++) {
            const Chunk &ck = chunks[k];
            // a chunk's last record ends in the bytes behind it: wait for the next chunk too (PF_HALO << a chunk)
            HERA_CUDA(cudaStreamWaitEvent(c.stream, chunks[std::min(k + 1, chunks.size() - 1)].landed, 0));
            ParseArgs ak = a;
            ak.text = text_p + ck.off;
            ak.n = n - ck.off;
            ak.owned = ck.owned;
            ak.tok_base0_dev = carry;
            ak.prev_nonws = ck.off > 0 && !host_is_space(host->text[ck.off - 1]) ? 1 : 0;
            ak.at_eof = 1;   // text[n] is the end of the file for every chunk: ak.n reaches to the end of the text
            ak.tiles = ck.tiles;
            ak.tok_state = tok_state + ck.tile0;
            ak.kept_state = kept_state + ck.tile0;
            ak.ctr = part + k;
            ak.kept_base_dev = carry + 1;
            ak.name_off0 = ck.off;
            HERA_LAUNCH(c, paf_parse_kernel, std::min<uint32_t>(ck.tiles, static_cast<uint32_t>(sms) * PF_MIN_CTAS), PF_THREADS, PF_SMEM_BYTES, ak);
            HERA_LAUNCH(c, chunk_carry_kernel, 1, 1, 0, ctr, part + k, carry, ck.off);
        }
    } else if (warp_variant) {
        HERA_LAUNCH(c, paf_parse_warp_kernel, std::min<uint32_t>(div_up(tiles, WW_WARPS), static_cast<uint32_t>(sms) * WW_MIN_CTAS), WW_THREADS, 0, a);
    } else {
        HERA_LAUNCH(c, paf_parse_kernel, std::min<uint32_t>(tiles, static_cast<uint32_t>(sms) * PF_MIN_CTAS), PF_THREADS, PF_SMEM_BYTES, a);
    }
    HERA_CUDA(cudaEventRecord(ev1, c.stream));
    mark("parse");
    HERA_LAUNCH(c, kept_overflow_kernel, 1, 1, 0, ctr, room);

    if (sharded) {   // every rank learns every rank's kept count and whether anybody met an irregular record
        c.xch->all_gather_dev(c, c.counters.p + 2, d_kb_all, 2);   // ParseCounters::kept, first_bad_rec
        mark("gather2");
        HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 16, d_tok_all, 3ull * ranks * sizeof(uint64_t), cudaMemcpyDeviceToHost, c.stream));
    }
    HERA_CUDA(cudaMemcpyAsync(c.h_counters.p, ctr, sizeof(ParseCounters), cudaMemcpyDeviceToHost, c.stream));
    if (debug_timing) { t_launched = host_now_ms(); cudaEventRecord(ev_out, c.stream); }
    HERA_CUDA(cudaStreamSynchronize(c.stream));
    if (debug_timing) {
        float lead = 0, trail = 0;
        cudaEventElapsedTime(&lead, ev_in, ev0);
        cudaEventElapsedTime(&trail, ev1, ev_out);
        fprintf(stderr, "[hera rank %d] load: device time entry -> timed region %.3f ms, parse kernel -> last copy %.3f ms; host clock at phase ends:%s end +%.3f\n", me, lead, trail,
                phase_log.c_str(), host_now_ms() - t_entry);
        cudaEventDestroy(ev_in);
        cudaEventDestroy(ev_out);
    }
    if (sharded && debug_timing) {
        float pre = 0, krn = 0;
        cudaEventElapsedTime(&pre, ev0, ev2);
        cudaEventElapsedTime(&krn, ev2, ev1);
        fprintf(stderr, "[hera rank %d] load: token exchange %.3f ms, parse kernel %.3f ms; host clock: entry %.3f, all-gather queued +%.3f, all queued +%.3f, synchronised +%.3f\n",
                me, pre, krn, t_entry, t_gather - t_entry, t_launched - t_entry, host_now_ms() - t_entry);
    }
    ParseCounters hc = *reinterpret_cast<ParseCounters *>(c.h_counters.p);
    bool overflow = (hc.kept & KEPT_OVERFLOW) != 0;
    for (int r = 0; sharded && r < ranks; r++) overflow = overflow || (c.h_counters.p[16 + ranks + 2 * r] & KEPT_OVERFLOW) != 0;
    if (overflow) {   // (only after an optimistic reservation; the same decision on every rank)
        ev.destroy();
        HERA_REQUIRE(!c.parse_force_count, 3, "internal: kept rows overflowed a counted reservation");
        c.parse_force_count = true;
        try {
            load_text_resident(c, text, file_id, anchors, test_load_num, shard, out, nullptr);   // (the text has landed by now)
        } catch (...) {
            c.parse_force_count = false;
            throw;
        }
        c.parse_force_count = false;
        return;
    }
    uint64_t tokens_all = 0;
    for (int r = 0; sharded && r < ranks; r++) tokens_all += c.h_counters.p[16 + r];

    const uint64_t T = sharded ? tokens_all : hc.total_tokens;
    uint64_t records = T / 16 + ((T % 16) >= 12 ? 1 : 0);
    if (records > rec_limit) records = rec_limit;
    uint64_t kept_new = hc.kept, kept_global = 0;
    if (sharded) {
        const uint64_t *gathered = c.h_counters.p + 16 + ranks;
        std::vector<uint64_t> counts(ranks), displs(ranks);
        uint64_t total = 0, first_bad = ~0ull;
        for (int r = 0; r < ranks; r++) {
            counts[r] = gathered[2 * r];
            displs[r] = total;
            total += counts[r];
            first_bad = std::min(first_bad, gathered[2 * r + 1]);
        }
        if (first_bad < records) {
            throw Error{20, "irregular PAF record " + std::to_string(first_bad) + ": formatted extraction leaves the 16-token grid there; load this file unsharded"};
        }
        // The kept rows stay where they were parsed (node ids and adjacency are built from every rank's own
        // rows, see finalize_graph); only their number per rank is known everywhere.
        LoadShape shape;
        shape.global0 = c.rows_total;
        shape.total = total;
        shape.counts = counts;
        shape.displs = displs;
        c.load_shapes.push_back(shape);
        HERA_CUDA(cudaEventRecord(ev2, c.stream));
        c.own_rows.push_back({c.rows_total + displs[me], c.rows_total + displs[me] + counts[me]});
        kept_new = counts[me];
        kept_global = total;
    } else if (hc.first_bad_rec < records) {
        // truncate to the records before the irregular one, then continue serially from there
        unsigned long long *d_cut = reinterpret_cast<unsigned long long *>(c.counters.p + 12);
        HERA_CUDA(cudaMemsetAsync(d_cut, 0, sizeof(unsigned long long), c.stream));
        if (kept_new) HERA_LAUNCH(c, kept_before_kernel, div_up(kept_new, 256), 256, 0, c.kept.rec.p, c.kept.n, kept_new, hc.first_bad_rec, d_cut);
        HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 12, d_cut, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        kept_new = c.h_counters.p[12];
        HERA_LAUNCH(c, paf_tail_kernel, 1, 1, 0, text_p, n, hc.first_bad_off, hc.first_bad_rec, rec_limit, file_id, anchors ? 1 : 0, fd,
                    kept_ptrs(c.kept), c.kept.n + kept_new, c.kept.qs.cap, ctr);
        HERA_CUDA(cudaMemcpyAsync(c.h_counters.p, ctr, sizeof(ParseCounters), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        hc = *reinterpret_cast<ParseCounters *>(c.h_counters.p);
        records = hc.first_bad_rec + hc.tail_records;
        kept_new += hc.tail_kept;
    }
    if (!sharded) {
        // A whole text on a sharded context (the exact reader's fallback for an irregular file): every rank parsed the
        // same rows; rank 0 keeps them, the others contribute none.
        kept_global = kept_new;
        if (me != 0) kept_new = 0;
        c.own_rows.push_back({c.rows_total, c.rows_total + kept_new});
        LoadShape shape;
        shape.global0 = c.rows_total;
        shape.total = kept_global;
        shape.counts.assign(ranks, 0);
        shape.counts[0] = kept_global;
        shape.displs.assign(ranks, kept_global);
        shape.displs[0] = 0;
        c.load_shapes.push_back(shape);
    }
    if (hc.name_too_long || records >= (1ull << 32)) {
        if (!c.own_rows.empty()) { c.own_rows.pop_back(); c.load_shapes.pop_back(); }
        HERA_REQUIRE(!hc.name_too_long, 22, "a kept record has a sequence name of 65536 bytes or more: not supported (name references hold 16 bits of length)");
        HERA_REQUIRE(records < (1ull << 32), 23, "more than 2^32 records in one load: not supported (record indices are 32 bits)");
    }
    c.kept.n += kept_new;
    c.rows_total += kept_global;
    c.gathered.valid = false;
    c.g.ready = false;
    if (out) {
        out->records = records;
        out->kept = kept_global;
        out->kept_total = c.rows_total;
        out->bytes = n;
        HERA_CUDA(cudaEventElapsedTime(&out->parse_ms, ev0, ev1));   // (a chunked load: from the first chunk's launch to the last one's end, waits for the copy included)
        if (host) {
            HERA_CUDA(cudaEventSynchronize(chunk_events.t1));   // recorded right behind the last chunk's event, which the stream has waited for
            HERA_CUDA(cudaEventElapsedTime(&out->h2d_ms, chunk_events.t0, chunk_events.t1));
        }
        out->exchange_ms = 0;
        if (sharded) {
            HERA_CUDA(cudaEventSynchronize(ev2));
            HERA_CUDA(cudaEventElapsedTime(&out->exchange_ms, ev1, ev2));
        }
    }
}

// ---- hera_paf_parse_noalloc: the tokenizer / filter / score kernel for callers that cannot allocate ---------------------
// (No counterpart in the reference.)  The parse kernel needs scratch memory --
// the ParseCounters block and the two look-back chains, one 64-bit state word per tile and chain -- which the ordinary entry
// points keep in a growable cudaMalloc'ed buffer of the context.  This variant takes it from a file-scope device array of
// fixed size, so the call path contains no cudaMalloc / cudaFree / stream-ordered allocation at all: usable under CUDA
// graph capture, in serving loops and in harnesses that forbid allocation.  The array limits the text to
// HERA_NOALLOC_MAX_TEXT_BYTES per call; every GPU that loads the library holds the array (~256 KB) until the process exits.
// All calls on a device share it and it keeps the previous call's contents: every call clears what it uses first, and two
// calls whose kernels overlap on the GPU (different streams, no dependency) race.
constexpr uint32_t NA_MAX_TILES = static_cast<uint32_t>(HERA_NOALLOC_MAX_TEXT_BYTES / PF_TILE) + 1;
constexpr uint32_t NA_HEAD_WORDS = 16;   // ParseCounters (8 words) + spare
constexpr uint64_t NA_WORDS = NA_HEAD_WORDS + 2ull * NA_MAX_TILES;
static_assert(sizeof(ParseCounters) <= NA_HEAD_WORDS * sizeof(unsigned long long), "counter block fits the head of the scratch array");
static __device__ __align__(256) unsigned long long g_noalloc_scratch[NA_WORDS];

// Clears the words this call uses (the head and 2 * tiles chain states) and arms the "first irregular record" minima.
__global__ void __launch_bounds__(256) noalloc_init_kernel(unsigned long long *scratch, uint64_t words) {
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < words; i += stride) scratch[i] = 0;
}
__global__ void noalloc_arm_kernel(ParseCounters *ctr) {
    ctr->first_bad_rec = ~0ull;
    ctr->first_bad_off = ~0ull;
}
// result[0] records consumed ("ctr", src/OverlapGraph.cpp:128), [1] kept overlaps, [2] index of the first record formatted
// extraction would fail or desynchronise on (UINT64_MAX: none), [3] tokens in the text
__global__ void noalloc_result_kernel(const ParseCounters *ctr, uint64_t rec_limit, unsigned long long *result) {
    const unsigned long long T = ctr->total_tokens;
    unsigned long long records = T / 16 + ((T % 16) >= 12 ? 1 : 0);
    if (records > rec_limit) records = rec_limit;
    result[0] = records;
    result[1] = ctr->kept;
    result[2] = ctr->first_bad_rec < records ? ctr->first_bad_rec : ~0ull;
    result[3] = T;
}

namespace {
struct NoAllocDevice {
    bool ready = false;
    unsigned long long *scratch = nullptr;   // DEVICE address of g_noalloc_scratch on this device (cudaGetSymbolAddress)
    int sms = 0;
};
NoAllocDevice g_noalloc_dev[64];
std::mutex g_noalloc_mutex;
}  // namespace

// Once per device, with that device current: the device address of the scratch array (in host code the array's name is
// only a stand-in for the symbol, not an address a kernel could use) and the launch attributes of the kernel.
static const NoAllocDevice &noalloc_setup(int device) {
    std::lock_guard<std::mutex> lk(g_noalloc_mutex);
    HERA_REQUIRE(device >= 0 && device < 64, 101, "hera_paf_parse_noalloc: device index out of range");
    NoAllocDevice &d = g_noalloc_dev[device];
    if (!d.ready) {
        void *ptr = nullptr;
        HERA_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        HERA_REQUIRE(ptr != nullptr, 2, "cudaGetSymbolAddress returned a null address for the scratch array");
        d.scratch = static_cast<unsigned long long *>(ptr);
        HERA_CUDA(cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, device));
        HERA_CUDA(cudaFuncSetAttribute(paf_parse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PF_SMEM_BYTES));
        d.ready = true;
    }
    return d;
}

void paf_parse_noalloc(int device, const void *device_text, uint64_t n, int anchors, const hera_filter_params &fp, uint32_t test_load_num,
                       const hera_kept_columns &out, uint64_t capacity, uint64_t *d_result, cudaStream_t stream) {
    HERA_REQUIRE(n <= HERA_NOALLOC_MAX_TEXT_BYTES, 60, "hera_paf_parse_noalloc: the text is longer than HERA_NOALLOC_MAX_TEXT_BYTES (" +
                 std::to_string(static_cast<unsigned long long>(HERA_NOALLOC_MAX_TEXT_BYTES)) + " bytes): use hera_graph_load_device_view");
    HERA_REQUIRE(device_text != nullptr && n > 0, 101, "hera_paf_parse_noalloc: empty text");
    HERA_REQUIRE((reinterpret_cast<uintptr_t>(device_text) & 15u) == 0, 101, "hera_paf_parse_noalloc: the text must start at a 16-byte aligned address");
    HERA_REQUIRE(d_result != nullptr, 101, "hera_paf_parse_noalloc: null result pointer");
    HERA_REQUIRE(capacity >= n / 24 + 2, 61, "hera_paf_parse_noalloc: the output columns must hold n / 24 + 2 rows (every record could be kept)");
    HERA_REQUIRE(out.qs && out.qe && out.ts && out.te && out.qlen && out.tlen && out.rec && out.os && out.si && out.es && out.flags && out.qhash &&
                 out.thash && out.qname && out.tname, 101, "hera_paf_parse_noalloc: null output column");
    HERA_CUDA(cudaSetDevice(device));
    const NoAllocDevice &d = noalloc_setup(device);
    const uint32_t tiles = static_cast<uint32_t>(n / PF_TILE + 1);
    ParseCounters *ctr = reinterpret_cast<ParseCounters *>(d.scratch);
    unsigned long long *tok_state = d.scratch + NA_HEAD_WORDS, *kept_state = tok_state + tiles;
    const uint64_t words = NA_HEAD_WORDS + 2ull * tiles;
    noalloc_init_kernel<<<std::min<unsigned>(div_up(words, 256), 148u), 256, 0, stream>>>(d.scratch, words);
    noalloc_arm_kernel<<<1, 1, 0, stream>>>(ctr);
    FilterDev fd{fp.mode, static_cast<float>(fp.min_overlap_length), fp.min_overlap_percentage, static_cast<float>(fp.max_overhang_length), fp.max_overhang_percentage};
    const uint64_t rec_limit = test_load_num > 0 ? static_cast<uint64_t>(test_load_num) + 1 : ~0ull;
    const KeptPtrs kp{out.qs, out.qe, out.ts, out.te, out.qlen, out.tlen, out.rec, out.os, out.si, out.es, out.flags,
                      reinterpret_cast<uint64_t *>(out.qhash), reinterpret_cast<uint64_t *>(out.thash), reinterpret_cast<uint64_t *>(out.qname),
                      reinterpret_cast<uint64_t *>(out.tname)};
    ParseArgs a{static_cast<const uint8_t *>(device_text), 1, n, n, 0, nullptr, 0, 1, tiles, 0u, anchors ? 1 : 0, rec_limit, 0, fd, kp, tok_state, kept_state, ctr};
    paf_parse_kernel<<<std::min<uint32_t>(tiles, static_cast<uint32_t>(d.sms) * PF_MIN_CTAS), PF_THREADS, PF_SMEM_BYTES, stream>>>(a);
    noalloc_result_kernel<<<1, 1, 0, stream>>>(ctr, rec_limit, reinterpret_cast<unsigned long long *>(d_result));
    HERA_CUDA(cudaGetLastError());
}

}  // namespace hera
