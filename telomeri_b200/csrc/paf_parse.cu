// K1 + K2: PAF text -> kept overlaps, one pass over the text.
//
// Replaces the read loop, filter and score block of OverlapGraph::load / filter / buildFrom
// (reference src/OverlapGraph.cpp:97-202, 221-231, 297-324).  The reference consumes the file
// as a whitespace-separated token stream, 16 tokens per record (12 typed + 4 skipped), so a
// record boundary is "token index % 16 == 0", not a newline.
//
// One CTA owns one 8 KiB tile of text (plus a 512 B halo so the records that start in the tile
// can be finished).  Tiles are handed out by an atomic ticket so that a tile never waits for a
// tile that has not started (chained scan with decoupled look-back):
//   1. 128-bit coalesced loads stage tile+halo in shared memory; every thread classifies its
//      16-byte chunks into a non-space bitmask; token starts = non-space after space.
//   2. the tile's token count is published; a warp looks back over earlier tiles for the global
//      index of the tile's first token (first chain).
//   3. token starts are compacted to shared-memory offsets; every 16th global token opens a
//      record.  A half-warp takes one record: lane k parses token k (names -> 64-bit hash,
//      integers -> int32 with validity, strand char); warp shuffles hand the twelve fields to
//      the half-warp leader, which runs the fp32 filter and scores in the reference's
//      operation order (explicit round-to-nearest intrinsics, no FMA contraction).
//   4. kept records are counted, the count goes through a second look-back chain, and the CTA
//      writes its kept records at their final file-order position.
// A record is "regular" when its nine integer tokens are [+-]?digits within int range and the
// strand token is one character.  The first irregular record (stream extraction would fail or
// desynchronise there) ends the parallel part; a one-thread tail kernel continues from its byte
// offset with the exact character-stream rules of formatted extraction.
#include "ctx.cuh"
#include "primitives.cuh"

namespace hera {

constexpr int PF_THREADS = 256;
constexpr int PF_TILE = 8192;
constexpr int PF_HALO = 512;
constexpr int PF_WIN = PF_TILE + PF_HALO;
constexpr int PF_CHUNKS = PF_WIN / 16;
constexpr int PF_MAXTOK = PF_WIN / 2 + 1;
constexpr int PF_MAXREC = PF_TILE / 2 / 16 + 2;

struct FilterDev {
    int mode;
    float min_len_f, min_pct, max_ohl_f, max_pct;
};

struct ParseCounters {  // device block, zeroed per load
    unsigned int ticket;
    unsigned int pad;
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
constexpr uint64_t HASH_SEED = 0xcbf29ce484222325ull;
constexpr uint64_t HASH_MUL = 0x100000001b3ull;

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
            ovl = __fdiv_rn(so, 2.0f); ohg = __fdiv_rn(sh, 2.0f); tot = __fdiv_rn(st, 2.0f);
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
    const float half_ovl = __fdiv_rn(__int2float_rn(static_cast<int>(static_cast<uint32_t>(q_ovl) + static_cast<uint32_t>(t_ovl))), 2.0f);
    os = __fmul_rn(half_ovl, si);
    const float ext = __fdiv_rn(__int2float_rn(qs), 2.0f);
    const float ohg = __fdiv_rn(__int2float_rn(static_cast<int>(static_cast<uint32_t>(q_ohg) + static_cast<uint32_t>(ts))), 2.0f);
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
        uint64_t h = HASH_SEED;
        while (p < n && !dev_is_space(t[p])) { h = (h ^ t[p]) * HASH_MUL; p++; }
        len = static_cast<uint32_t>(p - off);
        hash = hash_finish(h);
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
                                            uint32_t rec_idx, StagedRecord &out) {
    if (r.qhash == r.thash && r.qnlen == r.tnlen && names_equal(text, r.qoff, r.toff, r.qnlen)) return false;  // :157
    if (!filter_fields(f, r.qlen, r.qs, r.qe, r.tlen, r.ts, r.te)) return false;
    out.qs = static_cast<uint32_t>(r.qs); out.qe = static_cast<uint32_t>(r.qe);
    out.ts = static_cast<uint32_t>(r.ts); out.te = static_cast<uint32_t>(r.te);
    out.qlen = static_cast<uint32_t>(r.qlen); out.tlen = static_cast<uint32_t>(r.tlen);
    out.rec = rec_idx;
    score_fields(r.qlen, r.qs, r.qe, r.tlen, r.ts, r.te, r.matches, out.si, out.os, out.es);
    out.qhash = r.qhash; out.thash = r.thash;
    out.qname = pack_name(file_id, r.qoff, r.qnlen);
    out.tname = pack_name(file_id, r.toff, r.tnlen);
    out.flags = (r.strand == '-' ? KF_MINUS : 0) | (anchors ? KF_CR : 0) | ((anchors && is_ctg(text, r.qoff, r.qnlen)) ? KF_QCTG : 0);
    return true;
}

__device__ __forceinline__ void store_record(const KeptPtrs &k, uint64_t i, const StagedRecord &s) {
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

// Called by warp 0.  Publishes this tile's aggregate, walks back over earlier tiles, publishes
// the inclusive prefix and returns the exclusive prefix (same value in every lane).
__device__ unsigned long long chain_exclusive(unsigned long long *state, unsigned int tile, unsigned long long aggregate) {
    const uint32_t lane = lane_id();
    if (tile == 0) {
        if (lane == 0) st_state(&state[0], ST_INC | aggregate);
        return 0;
    }
    if (lane == 0) st_state(&state[tile], ST_AGG | aggregate);
    unsigned long long excl = 0;
    int look = static_cast<int>(tile) - 1;
    for (;;) {
        const int idx = look - static_cast<int>(lane);
        unsigned long long s = ST_INC;  // tiles before 0 behave like an inclusive prefix of 0
        if (idx >= 0) {
            do { s = ld_state(&state[idx]); } while ((s & ST_MASK) == 0);
        }
        const uint32_t inc_mask = __ballot_sync(0xffffffffu, (s & ST_MASK) == ST_INC);
        const uint32_t upto = inc_mask ? static_cast<uint32_t>(__ffs(inc_mask)) : 32u;  // lanes [0, upto) contribute
        unsigned long long v = lane < upto ? (s & ~ST_MASK) : 0ull;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        excl += v;
        if (inc_mask) break;
        look -= 32;
    }
    if (lane == 0) st_state(&state[tile], ST_INC | (excl + aggregate));
    return excl;
}

struct ParseArgs {
    const uint8_t *text;        // padded: readable up to tiles * PF_TILE + PF_HALO, bytes >= n are '\n'
    uint64_t n;
    uint32_t tiles;
    uint32_t file_id;
    int anchors;
    uint64_t rec_limit;         // records with index >= rec_limit are ignored (test_load_num)
    uint64_t kept_base;         // kept records of earlier loads
    FilterDev filter;
    KeptPtrs kept;
    unsigned long long *tok_state, *kept_state;
    ParseCounters *ctr;
};

__global__ void __launch_bounds__(PF_THREADS) paf_parse_kernel(const ParseArgs a) {
    __shared__ __align__(16) uint8_t s_text[PF_WIN];
    __shared__ uint16_t s_nonws[PF_CHUNKS + 1];
    __shared__ uint16_t s_chunk_pre[PF_CHUNKS + 1];
    __shared__ uint16_t s_tok_off[PF_MAXTOK];
    __shared__ StagedRecord s_rec[PF_MAXREC];
    __shared__ uint8_t s_keep[PF_MAXREC];
    __shared__ uint32_t s_scan[33];
    __shared__ unsigned int s_tile;
    __shared__ unsigned long long s_tok_base, s_kept_base;

    const uint32_t tid = threadIdx.x, lane = lane_id(), warp = warp_id();
    if (tid == 0) s_tile = atomicAdd(&a.ctr->ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint64_t tile_byte0 = static_cast<uint64_t>(tile) * PF_TILE;

    // 1. stage tile + halo, classify bytes
    const uint4 *src = reinterpret_cast<const uint4 *>(a.text + tile_byte0);
    for (int ch = tid; ch < PF_CHUNKS; ch += PF_THREADS) {
        const uint4 v = __ldg(src + ch);
        reinterpret_cast<uint4 *>(s_text)[ch] = v;
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint32_t m = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const uint32_t c = (w[j] >> (8 * b)) & 0xffu;
                m |= (dev_is_space(c) ? 0u : 1u) << (4 * j + b);
            }
        }
        s_nonws[ch + 1] = static_cast<uint16_t>(m);
    }
    if (tid == 0) {
        // byte before the tile decides whether the tile's first byte starts a token
        uint32_t prev_nonws = 0;
        if (tile_byte0 > 0) prev_nonws = dev_is_space(a.text[tile_byte0 - 1]) ? 0u : 1u;
        s_nonws[0] = static_cast<uint16_t>(prev_nonws << 15);
    }
    __syncthreads();

    // token-start masks and per-chunk counts; chunks of the tile proper count towards the chain
    uint32_t my_cnt[(PF_CHUNKS + PF_THREADS - 1) / PF_THREADS];
    uint32_t my_mask[(PF_CHUNKS + PF_THREADS - 1) / PF_THREADS];
    uint32_t tile_tokens_local = 0;
    // thread t owns chunks [t*3, t*3+3) so that the chunk order equals the thread order
    constexpr int CPT = (PF_CHUNKS + PF_THREADS - 1) / PF_THREADS;
#pragma unroll
    for (int k = 0; k < CPT; k++) {
        const int ch = tid * CPT + k;
        uint32_t starts = 0;
        if (ch < PF_CHUNKS) {
            const uint32_t cur = s_nonws[ch + 1];
            const uint32_t prev_top = (s_nonws[ch] >> 15) & 1u;
            starts = cur & ~((cur << 1) | prev_top) & 0xffffu;
        }
        my_mask[k] = starts;
        my_cnt[k] = __popc(starts);
        if (ch < PF_TILE / 16) tile_tokens_local += my_cnt[k];
    }
    uint32_t thread_total = 0;
#pragma unroll
    for (int k = 0; k < CPT; k++) thread_total += my_cnt[k];
    uint32_t win_tokens;
    uint32_t excl = block_exclusive_scan<uint32_t>(thread_total, s_scan, &win_tokens);
    uint32_t tile_tokens;
    block_exclusive_scan<uint32_t>(tile_tokens_local, s_scan, &tile_tokens);

    // 2. first chain: global index of the tile's first token
    if (warp == 0) {
        unsigned long long base = chain_exclusive(a.tok_state, tile, tile_tokens);
        if (lane == 0) {
            s_tok_base = base;
            if (tile == a.tiles - 1) a.ctr->total_tokens = base + tile_tokens;
        }
    }
    // 3a. compact token starts to offsets (all tokens of the window)
    {
        uint32_t pos = excl;
#pragma unroll
        for (int k = 0; k < CPT; k++) {
            const int ch = tid * CPT + k;
            uint32_t m = my_mask[k];
            while (m) {
                const int b = __ffs(m) - 1;
                m &= m - 1;
                if (pos < PF_MAXTOK) s_tok_off[pos] = static_cast<uint16_t>(ch * 16 + b);
                pos++;
            }
        }
    }
    __syncthreads();
    const unsigned long long tok_base = s_tok_base;
    const uint32_t first_rec_ord = static_cast<uint32_t>((16 - (tok_base & 15)) & 15);
    const uint32_t nrec = tile_tokens > first_rec_ord ? (tile_tokens - first_rec_ord + 15) / 16 : 0;  // records starting in this tile
    const bool window_reaches_eof = tile_byte0 + PF_WIN >= a.n;

    // 3b. one half-warp per record
    const uint32_t half = tid >> 4, k = tid & 15;
    for (uint32_t j0 = 0; j0 < nrec; j0 += PF_THREADS / 16) {
        const uint32_t j = j0 + half;
        const bool have_rec = j < nrec;
        const uint32_t ord = first_rec_ord + j * 16 + k;
        const bool exists = have_rec && ord < win_tokens;
        uint32_t off = exists ? s_tok_off[ord] : 0;
        // parse my token
        uint64_t h = HASH_SEED;
        unsigned long long val = 0;
        uint32_t len = 0, digits = 0, bad = 0, neg = 0, first_ch = 0;
        if (exists) {
            uint32_t c = s_text[off];
            first_ch = c;
            while (!dev_is_space(c)) {
                h = (h ^ c) * HASH_MUL;
                const uint32_t dgt = c - '0';
                if (dgt <= 9u) {
                    if (val <= 4294967295ull) val = val * 10 + dgt;
                    digits++;
                } else if (len == 0 && (c == '-' || c == '+')) {
                    neg = c == '-';
                } else {
                    bad = 1;
                }
                len++;
                if (off + len >= PF_WIN) break;
                c = s_text[off + len];
            }
        }
        const bool unterminated = exists && (off + len >= PF_WIN) && !window_reaches_eof;
        const bool int_ok = exists && !bad && digits > 0 && val <= (neg ? 2147483648ull : 2147483647ull);
        const int ival = neg ? static_cast<int>(0u - static_cast<uint32_t>(val)) : static_cast<int>(static_cast<uint32_t>(val));
        const bool is_int_tok = (k >= 1 && k <= 3) || (k >= 6 && k <= 11);
        const bool tok_ok = k >= 12 ? true : (is_int_tok ? int_ok : (k == 4 ? (exists && len == 1) : exists));
        const uint32_t hw_mask = 0xffffu << (lane & 16);
        const uint32_t ok_ballot = __ballot_sync(0xffffffffu, tok_ok) & hw_mask;
        const uint32_t exist_ballot = __ballot_sync(0xffffffffu, exists) & hw_mask;
        const uint32_t unterm_ballot = __ballot_sync(0xffffffffu, unterminated && k < 12) & hw_mask;
        const uint32_t base_lane = lane & 16;
        // gather
        SerialRecord r;
        r.qlen = __shfl_sync(0xffffffffu, ival, base_lane + 1); r.qs = __shfl_sync(0xffffffffu, ival, base_lane + 2);
        r.qe = __shfl_sync(0xffffffffu, ival, base_lane + 3); r.strand = __shfl_sync(0xffffffffu, first_ch, base_lane + 4);
        r.tlen = __shfl_sync(0xffffffffu, ival, base_lane + 6); r.ts = __shfl_sync(0xffffffffu, ival, base_lane + 7);
        r.te = __shfl_sync(0xffffffffu, ival, base_lane + 8); r.matches = __shfl_sync(0xffffffffu, ival, base_lane + 9);
        const uint64_t hf = hash_finish(h);
        r.qhash = __shfl_sync(0xffffffffu, hf, base_lane + 0); r.thash = __shfl_sync(0xffffffffu, hf, base_lane + 5);
        r.qnlen = __shfl_sync(0xffffffffu, len, base_lane + 0); r.tnlen = __shfl_sync(0xffffffffu, len, base_lane + 5);
        const uint32_t qoff_l = __shfl_sync(0xffffffffu, off, base_lane + 0), toff_l = __shfl_sync(0xffffffffu, off, base_lane + 5);
        r.qoff = tile_byte0 + qoff_l; r.toff = tile_byte0 + toff_l;
        if (k == 0 && have_rec) {
            const uint64_t rec_idx = (tok_base + first_rec_ord) / 16 + j;
            const uint32_t twelve = 0x0fffu << base_lane;
            bool keep = false;
            if (rec_idx < a.rec_limit) {
                const bool all_exist = (exist_ballot & twelve) == twelve;
                bool regular = all_exist && ((ok_ballot & twelve) == twelve);
                bool complete = all_exist;
                if ((!all_exist && !window_reaches_eof) || (unterm_ballot & twelve)) {
                    // record does not fit the window (very long names): exact serial reader from its first byte
                    SerialCursor cur{a.text, a.n, r.qoff, true};
                    SerialRecord sr;
                    const bool got = serial_read_record(cur, sr);
                    // regular iff the reader consumed whole tokens only; otherwise flag it for the tail kernel
                    bool whole = got;
                    if (got) {
                        // re-check token regularity: each integer token must end at white space
                        SerialCursor chk{a.text, a.n, r.qoff, true};
                        uint64_t o, hh; uint32_t l; int iv; uint32_t chh;
                        chk.word(o, l, hh);
                        for (int f = 0; f < 3 && whole; f++) { chk.integer(iv); whole = chk.p >= a.n || dev_is_space(a.text[chk.p]); }
                        if (whole) { chk.single(chh); whole = chk.p >= a.n || dev_is_space(a.text[chk.p]); }
                        if (whole) chk.word(o, l, hh);
                        for (int f = 0; f < 6 && whole; f++) { chk.integer(iv); whole = chk.p >= a.n || dev_is_space(a.text[chk.p]); }
                    }
                    complete = true;
                    regular = whole;
                    r = sr;
                }
                if (complete) {
                    if (!regular) {
                        atomicMin(&a.ctr->first_bad_rec, static_cast<unsigned long long>(rec_idx));
                        atomicMin(&a.ctr->first_bad_off, static_cast<unsigned long long>(tile_byte0 + s_tok_off[first_rec_ord + j * 16]));
                    } else {
                        keep = make_record(a.filter, a.text, a.file_id, a.anchors != 0, r, static_cast<uint32_t>(rec_idx), s_rec[j]);
                    }
                }
            }
            s_keep[j] = keep ? 1 : 0;
        }
    }
    __syncthreads();

    // 4. compact kept records of the tile, second chain, store
    uint32_t mine = 0;
    for (uint32_t j = tid; j < nrec; j += PF_THREADS) mine += s_keep[j];  // nrec <= PF_MAXREC <= 2 * PF_THREADS
    uint32_t tile_kept;
    // thread t owns records t and t + PF_THREADS; to keep file order, scan the two rounds separately
    uint32_t k0 = tid < nrec ? s_keep[tid] : 0;
    uint32_t k1 = tid + PF_THREADS < nrec ? s_keep[tid + PF_THREADS] : 0;
    uint32_t tot0, tot1;
    const uint32_t e0 = block_exclusive_scan<uint32_t>(k0, s_scan, &tot0);
    const uint32_t e1 = block_exclusive_scan<uint32_t>(k1, s_scan, &tot1);
    tile_kept = tot0 + tot1;
    (void) mine;
    if (warp == 0) {
        unsigned long long base = chain_exclusive(a.kept_state, tile, tile_kept);
        if (lane == 0) {
            s_kept_base = base;
            if (tile == a.tiles - 1) a.ctr->kept = base + tile_kept;
        }
    }
    __syncthreads();
    const uint64_t out0 = a.kept_base + s_kept_base;
    if (k0) store_record(a.kept, out0 + e0, s_rec[tid]);
    if (k1) store_record(a.kept, out0 + tot0 + e1, s_rec[tid + PF_THREADS]);
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
        if (make_record(filter, text, file_id, anchors != 0, r, static_cast<uint32_t>(rec), s)) {
            if (kept_pos + nk < kept_cap) store_record(kept, kept_pos + nk, s);
            nk++;
        }
        rec++;
    }
    ctr->tail_records = rec - first_rec;
    ctr->tail_kept = nk;
}

__global__ void pad_text_kernel(uint8_t *text, uint64_t n, uint64_t padded) {
    const uint64_t i = n + static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < padded) text[i] = '\n';
}

static KeptPtrs kept_ptrs(KeptTable &k) {
    return KeptPtrs{k.qs.p, k.qe.p, k.ts.p, k.te.p, k.qlen.p, k.tlen.p, k.rec.p, k.os.p, k.si.p, k.es.p, k.flags.p, k.qhash.p, k.thash.p, k.qname.p, k.tname.p};
}

uint64_t padded_text_size(uint64_t n) { return (n / PF_TILE + 2) * static_cast<uint64_t>(PF_TILE); }

void load_text_resident(Ctx &c, LoadedText &text, uint32_t file_id, bool anchors, uint32_t test_load_num, hera_load_result *out) {
    const uint64_t n = text.n;
    const uint64_t padded = padded_text_size(n);
    HERA_REQUIRE(text.bytes.cap >= padded, 3, "internal: text buffer not padded");
    HERA_LAUNCH(c, pad_text_kernel, div_up(padded - n, 256), 256, 0, text.bytes.p, n, padded);

    const uint32_t tiles = static_cast<uint32_t>(n / PF_TILE + 1);
    // worst case every record is kept: records <= bytes / 24 (12 one-byte tokens + separators)
    const uint64_t max_new = n / 24 + 2;
    c.kept.reserve(c.kept.n + max_new, c.stream);

    c.counters.reserve(64 + 2ull * tiles);
    ParseCounters *ctr = reinterpret_cast<ParseCounters *>(c.counters.p);
    unsigned long long *tok_state = reinterpret_cast<unsigned long long *>(c.counters.p + 16);
    unsigned long long *kept_state = tok_state + tiles;
    HERA_CUDA(cudaMemsetAsync(c.counters.p, 0, (16 + 2ull * tiles) * sizeof(uint64_t), c.stream));
    ParseCounters init{};
    init.first_bad_rec = ~0ull;
    init.first_bad_off = ~0ull;
    HERA_CUDA(cudaMemcpyAsync(ctr, &init, sizeof(init), cudaMemcpyHostToDevice, c.stream));

    FilterDev fd{c.fp.mode, static_cast<float>(c.fp.min_overlap_length), c.fp.min_overlap_percentage,
                 static_cast<float>(c.fp.max_overhang_length), c.fp.max_overhang_percentage};
    const uint64_t rec_limit = test_load_num > 0 ? static_cast<uint64_t>(test_load_num) + 1 : ~0ull;

    cudaEvent_t ev0, ev1;
    HERA_CUDA(cudaEventCreate(&ev0));
    HERA_CUDA(cudaEventCreate(&ev1));
    maybe_flush_l2(c);
    HERA_CUDA(cudaEventRecord(ev0, c.stream));
    ParseArgs a{text.bytes.p, n, tiles, file_id, anchors ? 1 : 0, rec_limit, c.kept.n, fd, kept_ptrs(c.kept), tok_state, kept_state, ctr};
    HERA_LAUNCH(c, paf_parse_kernel, tiles, PF_THREADS, 0, a);
    HERA_CUDA(cudaEventRecord(ev1, c.stream));

    c.h_counters.reserve(16);
    HERA_CUDA(cudaMemcpyAsync(c.h_counters.p, ctr, sizeof(ParseCounters), cudaMemcpyDeviceToHost, c.stream));
    HERA_CUDA(cudaStreamSynchronize(c.stream));
    ParseCounters hc = *reinterpret_cast<ParseCounters *>(c.h_counters.p);

    const uint64_t T = hc.total_tokens;
    uint64_t records = T / 16 + ((T % 16) >= 12 ? 1 : 0);
    if (records > rec_limit) records = rec_limit;
    uint64_t kept_new = hc.kept;
    if (hc.first_bad_rec < records) {
        // truncate to the records before the irregular one, then continue serially from there
        unsigned long long *d_cut = reinterpret_cast<unsigned long long *>(c.counters.p + 12);
        HERA_CUDA(cudaMemsetAsync(d_cut, 0, sizeof(unsigned long long), c.stream));
        if (kept_new) HERA_LAUNCH(c, kept_before_kernel, div_up(kept_new, 256), 256, 0, c.kept.rec.p, c.kept.n, kept_new, hc.first_bad_rec, d_cut);
        HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 12, d_cut, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        kept_new = c.h_counters.p[12];
        HERA_LAUNCH(c, paf_tail_kernel, 1, 1, 0, text.bytes.p, n, hc.first_bad_off, hc.first_bad_rec, rec_limit, file_id, anchors ? 1 : 0, fd,
                    kept_ptrs(c.kept), c.kept.n + kept_new, c.kept.qs.cap, ctr);
        HERA_CUDA(cudaMemcpyAsync(c.h_counters.p, ctr, sizeof(ParseCounters), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        hc = *reinterpret_cast<ParseCounters *>(c.h_counters.p);
        records = hc.first_bad_rec + hc.tail_records;
        kept_new += hc.tail_kept;
    }
    c.kept.n += kept_new;
    c.g.ready = false;
    if (out) {
        out->records = records;
        out->kept = kept_new;
        out->kept_total = c.kept.n;
        out->bytes = n;
        HERA_CUDA(cudaEventElapsedTime(&out->parse_ms, ev0, ev1));
    }
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
}

}  // namespace hera
