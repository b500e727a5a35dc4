// K3 + K4: node interning, CSR adjacency, per-metric rank permutations, canonical slots.
//
// K3 replaces OverlapGraph::nodeIndex and the node creation in buildFrom (reference
// src/OverlapGraph.cpp:208-218, 251-258): node id = rank of the name's first appearance among
// kept records (query before target, rr load before cr load).  Every rank interns the names of ITS kept
// rows in an open-addressing table keyed by the 64-bit name hash (atomicMin of the global occurrence
// position 2 * row + side); the distinct names of all ranks (one entry per name and rank: hash, first
// position, length, anchor rule) are all-gathered and merged into every rank's table, so the table then
// knows the global first appearance of every name.  A bitmap over occurrence positions with a
// popcount scan turns first appearances into node ids.  A verification pass compares every
// occurrence's bytes with its node's name, so a 64-bit hash collision is detected, never silently merged.
//
// K4 replaces the adjacency pushes (src/OverlapGraph.cpp:241-248): each kept overlap e yields
// entry 2e (owner = target, stored as read) and 2e+1 (owner = query, q/t swapped).  Every rank sorts
// the entries of its own rows by (owner node, load) with a stable LSD radix sort -- record order inside
// each node's list is the reference's insertion order -- and the per-node counts of all ranks are
// all-gathered: that fixes the global CSR offsets AND where each rank's run of a node's list starts, so
// each rank writes its finished 16-byte walk entries and slot->row words straight to their final place in
// EVERY rank's table (peer stores over NVLink; no all-to-all of rows, no pack / pull pass).  Rank
// permutations restate the per-step stable_sort of the greedy walk (src/PathManager.cpp:289-296);
// canon[] its first-parallel-edge lookup (:340-347); they are computed per node range and all-gathered.
#include <cstdio>
#include <cstring>
#include <vector>

#include "ctx.cuh"
#include "primitives.cuh"

namespace hera {

struct TextPtrs {
    const uint8_t *p[16];
};

RowPieces row_pieces(const Ctx &c) {
    RowPieces pc{};
    pc.n = static_cast<int>(c.own_rows.size());
    uint64_t local = 0;
    for (int i = 0; i < pc.n; i++) {
        pc.local0[i] = local;
        pc.global0[i] = c.own_rows[i].begin;
        pc.count[i] = c.own_rows[i].end - c.own_rows[i].begin;
        local += pc.count[i];
    }
    return pc;
}

// ---- K3 ------------------------------------------------------------------------------------------
// Probe for hash h: claims an empty slot or finds the name's slot, lowers its first position to pos.
__device__ __forceinline__ bool intern_probe(unsigned long long *keys, uint32_t *minpos, uint64_t mask, unsigned long long h, uint32_t pos,
                                             unsigned int *distinct, uint32_t *slot_out) {
    uint64_t slot = h & mask;
    for (uint64_t probes = 0; probes <= mask; probes++) {
        unsigned long long cur = keys[slot];
        if (cur == 0) {
            cur = atomicCAS(&keys[slot], 0ull, h);
            if (cur == 0) { atomicAdd(distinct, 1u); cur = h; }
        }
        if (cur == h) {
            // positions only ever go down: an occurrence behind the value already there cannot be the first one
            if (minpos[slot] > pos) atomicMin(&minpos[slot], pos);
            *slot_out = static_cast<uint32_t>(slot);
            return true;
        }
        slot = (slot + 1) & mask;
    }
    return false;
}

// occurrence o = 2 * local row + side (0 = query, 1 = target: the reference looks the query up first)
__global__ void intern_insert_kernel(uint64_t occ, RowPieces pc, const uint64_t *__restrict__ qhash, const uint64_t *__restrict__ thash,
                                     unsigned long long *keys, uint32_t *minpos, uint64_t mask, uint32_t *__restrict__ occ_slot,
                                     unsigned int *distinct, unsigned int *overflow) {
    const uint64_t o = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (o >= occ) return;
    const uint64_t l = o >> 1;
    const unsigned long long h = (o & 1) ? thash[l] : qhash[l];
    const uint32_t pos = static_cast<uint32_t>(2 * pc.global_of(l) + (o & 1));
    uint32_t slot = 0;
    if (intern_probe(keys, minpos, mask, h, pos, distinct, &slot)) occ_slot[o] = slot;
    else *overflow = 1;
}

// One entry per name of the local table: what another rank needs to know should this be the name's first appearance --
// length and anchor flag of THAT record (src/OverlapGraph.cpp:209-218), length of the name.
__global__ void intern_entries_kernel(uint64_t cap, const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ minpos, RowPieces pc,
                                      const uint32_t *__restrict__ qlen, const uint32_t *__restrict__ tlen, const uint8_t *__restrict__ flags,
                                      const uint64_t *__restrict__ qname, const uint64_t *__restrict__ tname, unsigned int *counter,
                                      uint64_t *__restrict__ ent_hash, uint32_t *__restrict__ ent_pos, uint32_t *__restrict__ ent_len, uint32_t *__restrict__ ent_meta) {
    const uint64_t slot = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (slot >= cap || keys[slot] == 0) return;
    const uint32_t pos = minpos[slot], side = pos & 1u;
    uint64_t l = 0;
    pc.local_of(pos >> 1, &l);   // the table only holds this rank's occurrences
    const uint8_t f = flags[l];
    const bool cr = f & KF_CR, qctg = f & KF_QCTG;
    const bool anchor = side ? (cr && !qctg) : (cr && qctg);   // ContigPosition::TARGET / QUERY
    const uint32_t i = atomicAdd(counter, 1u);
    ent_hash[i] = keys[slot];
    ent_pos[i] = pos;
    ent_len[i] = side ? tlen[l] : qlen[l];
    ent_meta[i] = name_len(side ? tname[l] : qname[l]) | (anchor ? 0x10000u : 0u);
}

__global__ void intern_merge_kernel(uint64_t T, const uint64_t *__restrict__ ent_hash, const uint32_t *__restrict__ ent_pos, unsigned long long *keys,
                                    uint32_t *minpos, uint64_t mask, unsigned int *distinct, unsigned int *overflow) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= T) return;
    uint32_t slot = 0;
    if (!intern_probe(keys, minpos, mask, ent_hash[i], ent_pos[i], distinct, &slot)) *overflow = 1;
}

// The entry whose position survived as the minimum describes the node; its bit is set in the first-appearance bitmap.
__global__ void intern_winner_kernel(uint64_t T, const uint64_t *__restrict__ ent_hash, const uint32_t *__restrict__ ent_pos,
                                     const uint32_t *__restrict__ ent_len, const uint32_t *__restrict__ ent_meta, const unsigned long long *__restrict__ keys,
                                     const uint32_t *__restrict__ minpos, uint64_t mask, uint32_t *__restrict__ slot_len, uint32_t *__restrict__ slot_meta,
                                     uint32_t *first_bits) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= T) return;
    const unsigned long long h = ent_hash[i];
    uint64_t slot = h & mask;
    while (keys[slot] != h) slot = (slot + 1) & mask;
    const uint32_t pos = ent_pos[i];
    if (minpos[slot] != pos) return;
    slot_len[slot] = ent_len[i];
    slot_meta[slot] = ent_meta[i];
    atomicOr(&first_bits[pos >> 5], 1u << (pos & 31u));
}

__global__ void popc_kernel(uint64_t words, const uint32_t *__restrict__ bits, uint32_t *__restrict__ cnt) {
    const uint64_t w = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (w < words) cnt[w] = __popc(bits[w]);
}

// node id = number of first appearances before this one (file order of the kept records, query before target)
__global__ void intern_nodes_kernel(uint64_t cap, const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ minpos,
                                    const uint32_t *__restrict__ first_bits, const uint32_t *__restrict__ word_prefix, const uint32_t *__restrict__ slot_len,
                                    const uint32_t *__restrict__ slot_meta, uint32_t *__restrict__ slot_id, uint32_t *__restrict__ node_len,
                                    uint8_t *__restrict__ node_anchor, uint32_t *__restrict__ node_pos, uint32_t *__restrict__ name_lens) {
    const uint64_t slot = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (slot >= cap || keys[slot] == 0) return;
    const uint32_t pos = minpos[slot], w = pos >> 5;
    const uint32_t id = word_prefix[w] + __popc(first_bits[w] & ((1u << (pos & 31u)) - 1u));
    slot_id[slot] = id;
    node_len[id] = slot_len[slot];
    node_anchor[id] = (slot_meta[slot] >> 16) & 1u;
    node_pos[id] = pos;
    name_lens[id] = slot_meta[slot] & 0xffffu;
}

__global__ void intern_assign_kernel(uint64_t L, const uint32_t *__restrict__ occ_slot, const uint32_t *__restrict__ slot_id, uint32_t *__restrict__ e_q,
                                     uint32_t *__restrict__ e_t) {
    const uint64_t l = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (l >= L) return;
    e_q[l] = slot_id[occ_slot[2 * l]];
    e_t[l] = slot_id[occ_slot[2 * l + 1]];
}

// Name bytes of the nodes whose creating row this rank parsed (the other ranks fill the rest).
__global__ void name_copy_kernel(uint64_t N, const uint32_t *__restrict__ node_pos, RowPieces pc, const uint64_t *__restrict__ qname,
                                 const uint64_t *__restrict__ tname, const uint64_t *__restrict__ offs, TextPtrs texts, char *__restrict__ blob) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const uint32_t pos = node_pos[i];
    uint64_t l = 0;
    if (!pc.local_of(pos >> 1, &l)) return;
    const uint64_t r = (pos & 1u) ? tname[l] : qname[l];
    const uint8_t *src = texts.p[name_file(r)] + name_off(r);
    char *dst = blob + offs[i];
    for (uint32_t k = 0; k < name_len(r); k++) dst[k] = static_cast<char>(src[k]);
}

// blob |= part[1..parts): every byte is non-zero in exactly one part.
__global__ void name_merge_kernel(uint64_t bytes, int parts, const char *__restrict__ gathered, char *__restrict__ blob) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= bytes) return;
    char v = 0;
    for (int p = 0; p < parts; p++) v |= gathered[static_cast<uint64_t>(p) * bytes + i];
    blob[i] = v;
}

// Verification: every occurrence this rank parsed has the same bytes as its node's name, so a
// 64-bit hash collision is an error, never a silent merge.
__global__ void intern_verify_kernel(uint64_t L, const uint32_t *__restrict__ e_q, const uint32_t *__restrict__ e_t,
                                     const uint64_t *__restrict__ qname, const uint64_t *__restrict__ tname, const uint64_t *__restrict__ name_off_tab,
                                     const char *__restrict__ blob, TextPtrs texts, unsigned int *collision) {
    // a small persistent grid (grid-stride loop): the kernel runs on a side stream beside the CSR build; a grid of L / 256 CTAs
    // would all be issued before the main stream's next kernels, which then wait for it
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    for (uint64_t e = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < L; e += stride) {
        const uint64_t mine[2] = {qname[e], tname[e]};
        const uint32_t ids[2] = {e_q[e], e_t[e]};
#pragma unroll
        for (int s = 0; s < 2; s++) {
            const uint64_t o = name_off_tab[ids[s]];
            const uint32_t len = static_cast<uint32_t>(name_off_tab[ids[s] + 1] - o);
            bool same = len == name_len(mine[s]);
            if (same) {
                const char *a = blob + o;
                const uint8_t *b = texts.p[name_file(mine[s])] + name_off(mine[s]);
                for (uint32_t i = 0; i < len; i++) if (static_cast<uint8_t>(a[i]) != b[i]) { same = false; break; }
            }
            if (!same) *collision = 1;
        }
    }
}

// ---- K4 ------------------------------------------------------------------------------------------
// Local entry i = 2 * local row + side; key = owner node * P + load (P loads): inside a node's list the entries of an
// earlier load come first whatever rank parsed them.
__global__ void csr_keys_kernel(uint64_t L, RowPieces pc, uint32_t P, const uint32_t *__restrict__ e_q, const uint32_t *__restrict__ e_t,
                                uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= 2 * L) return;
    const uint64_t l = i >> 1;
    uint32_t load = 0;
    for (int k = 1; k < pc.n; k++) if (l >= pc.local0[k]) load = k;
    keys[i] = ((i & 1) ? e_q[l] : e_t[l]) * P + load;
    vals[i] = static_cast<uint32_t>(i);
}

// run_start / run_end of every key in the sorted local entries (both zero for a key that does not occur)
__global__ void csr_runs_kernel(uint64_t S, const uint32_t *__restrict__ keys, uint32_t *__restrict__ run_start, uint32_t *__restrict__ run_end) {
    const uint64_t s = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const uint32_t k = keys[s];
    if (s == 0 || keys[s - 1] != k) run_start[k] = static_cast<uint32_t>(s);
    if (s + 1 == S || keys[s + 1] != k) run_end[k] = static_cast<uint32_t>(s + 1);
}

__global__ void csr_counts_kernel(uint64_t K, const uint32_t *__restrict__ run_start, const uint32_t *__restrict__ run_end, uint32_t *__restrict__ cnt) {
    const uint64_t k = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (k < K) cnt[k] = run_end[k] - run_start[k];
}

// cnt_all[r][x * P + p] = entries of node x rank r holds from load p.  Global list order of node x: loads in order, ranks in
// order inside a load, local (= file) order inside a rank's run.  deg[x] = all of them; before[x * P + p] = entries in
// front of this rank's run (x, p).
__global__ void csr_degrees_kernel(uint64_t N, uint32_t P, int ranks, int me, const uint32_t *__restrict__ cnt_all, uint32_t *__restrict__ deg,
                                   uint32_t *__restrict__ before) {
    const uint64_t x = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (x >= N) return;
    const uint64_t K = N * P;
    uint32_t run = 0;
    for (uint32_t p = 0; p < P; p++) {
        for (int r = 0; r < ranks; r++) {
            if (r == me) before[x * P + p] = run;
            run += cnt_all[static_cast<uint64_t>(r) * K + x * P + p];
        }
    }
    deg[x] = run;
}

constexpr int MAX_PEERS = 16;
struct PeerTables {
    uint4 *adj4[MAX_PEERS];
    uint32_t *pay[MAX_PEERS];
    int n;
};

// Sorted local entry s -> its final slot; the finished walk entry {neighbour | anchor bit, es, os, q_start - t_start
// (src/Path.cpp:13)} and the slot -> row word go to every rank's tables.
__global__ void csr_scatter_kernel(uint64_t S, uint32_t P, const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals,
                                   const uint32_t *__restrict__ run_start, const uint32_t *__restrict__ adj_off, const uint32_t *__restrict__ before,
                                   RowPieces pc, const uint32_t *__restrict__ e_q, const uint32_t *__restrict__ e_t, const uint32_t *__restrict__ qs,
                                   const uint32_t *__restrict__ ts, const float *__restrict__ es, const float *__restrict__ os,
                                   const uint8_t *__restrict__ node_anchor, PeerTables t) {
    const uint64_t s = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const uint32_t k = keys[s], v = vals[s], x = k / P;
    const uint64_t l = v >> 1;
    const uint32_t side = v & 1u;
    const uint32_t dest = adj_off[x] + before[k] + (static_cast<uint32_t>(s) - run_start[k]);
    const uint32_t nbr = side ? e_t[l] : e_q[l];
    const int64_t d = side ? static_cast<int64_t>(ts[l]) - static_cast<int64_t>(qs[l]) : static_cast<int64_t>(qs[l]) - static_cast<int64_t>(ts[l]);
    const uint4 entry = make_uint4(nbr | (node_anchor[nbr] ? 0x80000000u : 0u), __float_as_uint(es[l]), __float_as_uint(os[l]),
                                   static_cast<uint32_t>(static_cast<int32_t>(d)));
    const uint32_t pay = static_cast<uint32_t>(2 * pc.global_of(l) + side);
    for (int r = 0; r < t.n; r++) {
        t.adj4[r][dest] = entry;
        t.pay[r][dest] = pay;
    }
}

__device__ __forceinline__ float metric_value(int metric, float es, float os);

// One warp per node.  The roulette of the Monte-Carlo walk needs the fp64 sum of the candidates'
// metrics in list order (src/PathManager.cpp:63-66,107-113).  If every value of the node's list is a
// non-negative finite fp32 number, hence an integer multiple of q = 2^(smallest exponent - 23), and the
// sum of ALL of them stays below 2^53 * q, then every partial sum over every subset is an integer
// multiple of q below 2^53 * q: exactly representable, so any summation order gives the sequential
// result bit for bit.  Bit m of node_exact[x] records that for metric m.
// The same pass writes the 8-byte views {neighbour | anchor bit, metric} the Monte-Carlo walks stream (one per stored score;
// the walk stage is bound by the bytes it reads per examined entry, and a roulette scan needs neither the other score nor
// the coordinate delta of an entry it does not take).
__global__ void __launch_bounds__(128) node_exact_kernel(uint32_t N, const uint32_t *__restrict__ off, const uint4 *__restrict__ adj4,
                                                         uint8_t *__restrict__ node_exact, uint32_t *__restrict__ max_deg, uint2 *__restrict__ a8_es,
                                                         uint2 *__restrict__ a8_os) {
    const uint32_t x = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (x >= N) return;
    const uint32_t base = off[x], deg = off[x + 1] - base;
    if (lane == 0) atomicMax(max_deg, deg);
    uint32_t flags = 0;
    bool anchor_nbr = false;   // bit 4: some neighbour is an anchor (only then can a walk step be forced, src/PathManager.cpp:57-60)
    bool weird[4] = {false, false, false, false};
    int min_exp[4] = {1 << 20, 1 << 20, 1 << 20, 1 << 20};
    double part[4] = {0.0, 0.0, 0.0, 0.0};
    for (uint32_t k = lane; k < deg; k += 32) {
        const uint4 e = adj4[base + k];
        a8_es[base + k] = make_uint2(e.x, e.y);
        a8_os[base + k] = make_uint2(e.x, e.z);
        anchor_nbr |= (e.x >> 31) != 0;
#pragma unroll
        for (int m = 0; m < 4; m++) {
            const float v = metric_value(m, __uint_as_float(e.y), __uint_as_float(e.z));
            const uint32_t bits = __float_as_uint(v), ex = (bits >> 23) & 0xffu;
            if ((bits >> 31) || ex == 0xffu) weird[m] = true;
            else if (bits != 0) min_exp[m] = min(min_exp[m], static_cast<int>(ex ? ex : 1u) - 150);
            part[m] = __dadd_rn(part[m], static_cast<double>(v));
        }
    }
    if (__any_sync(0xffffffffu, anchor_nbr)) flags |= 16u;
#pragma unroll
    for (int m = 0; m < 4; m++) {
        int me = min_exp[m];
        double pt = part[m];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            me = min(me, __shfl_xor_sync(0xffffffffu, me, d));
            pt = __dadd_rn(pt, __shfl_xor_sync(0xffffffffu, pt, d));
        }
        bool exact = !__any_sync(0xffffffffu, weird[m]);
        if (exact && pt > 0.0) {
            const int total_exp = static_cast<int>((__double_as_longlong(pt) >> 52) & 0x7ff) - 1023;
            exact = total_exp < me + 53;
        }
        if (exact) flags |= 1u << m;
    }
    if (lane == 0) node_exact[x] = static_cast<uint8_t>(flags);
}

// adj_off at the node-range boundaries of the ranks (N r / ranks), widened for the host
__global__ void node_cut_kernel(const uint32_t *__restrict__ off, uint64_t N, int ranks, uint64_t *__restrict__ out) {
    for (int r = threadIdx.x; r <= ranks; r += blockDim.x) out[r] = off[N * r / ranks];
}

__global__ void anchor_list_kernel(uint64_t N, const uint8_t *__restrict__ node_anchor, const uint32_t *__restrict__ anchor_scan, uint32_t *__restrict__ anchor_ids) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < N && node_anchor[i]) anchor_ids[anchor_scan[i]] = static_cast<uint32_t>(i);
}

constexpr int RK_THREADS = 128;

__device__ __forceinline__ float metric_value(int metric, float es, float os) {
    switch (metric) {
        case HERA_EXTENSION_SCORE: return es;
        case HERA_OVERLAP_SCORE: return os;
        case HERA_EXTENSION_SCORE_SQRT: return __fsqrt_rn(es);
        default: return __fsqrt_rn(os);
    }
}
// The rank kernels read the 8-byte views {neighbour | anchor bit, score} (node_exact_kernel wrote them just before): a list
// is read once per metric and once for the canonical slots, 8 bytes per entry instead of the 16-byte walk entry.
struct SlotView { const uint2 *es, *os; };
__device__ __forceinline__ float slot_metric(const SlotView &v, int metric, uint32_t s) {
    const float x = __uint_as_float(((metric & 1) ? v.os : v.es)[s].y);   // metrics: 0 ES, 1 OS, 2 sqrt(ES), 3 sqrt(OS)
    return metric >= 2 ? __fsqrt_rn(x) : x;
}

// K4b: canonical slots and rank permutations of every node's list, staged in shared memory.
//   canon[i] = first slot of the list with the same neighbour (shared-memory hash table: neighbour -> min index).
//   rank[m]  = stable descending order by metric m, the order the reference recomputes at every greedy step
//              (src/PathManager.cpp:289-296): a bitonic sort of 64-bit keys {orderable metric bits, ~index}; equal
//              metrics keep list order because the index is part of the key (-0.0 is folded into +0.0 first, as
//              `>` sees them).  A list with a NaN metric, or longer than CAP, is ranked by counting instead:
//              position of slot i = #{j : m[j] > m[i]} + #{j < i : m[j] == m[i]}  (what `>` on NaN gives).
// This kernel takes the lists longer than LOW entries (the short ones go to rank_small_kernel below).
// Where the permutations and canonical slots go.  (The kernels can store into several ranks' tables at once; measured on
// 8 GPUs the 4-byte peer stores of the sorting warps reached 220 GB/s, so a rank now writes its own tables and
// slice_push_kernel copies the finished slices to the peers with 16-byte stores.)
struct RankPtrs {
    uint32_t *p[4][MAX_PEERS];
    uint32_t *canon[MAX_PEERS];
    int n;
    __device__ __forceinline__ void put_rank(int m, uint32_t at, uint32_t v) const {
        for (int r = 0; r < n; r++) p[m][r][at] = v;
    }
    __device__ __forceinline__ void put_canon(uint32_t at, uint32_t v) const {
        for (int r = 0; r < n; r++) canon[r][at] = v;
    }
};

__device__ __forceinline__ uint64_t rank_key(float v, uint32_t i, bool &nan) {
    uint32_t bits = __float_as_uint(v);
    if (v != v) nan = true;
    if (bits == 0x80000000u) bits = 0;
    const uint32_t u = (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
    return (static_cast<uint64_t>(u) << 32) | (0xffffffffu - i);
}

__device__ void rank_by_counting(const SlotView adj4, int metric, uint32_t base, uint32_t deg, const RankPtrs &out) {
    for (uint32_t i = threadIdx.x; i < deg; i += RK_THREADS) {
        const float mi = slot_metric(adj4, metric, base + i);
        uint32_t r = 0;
        for (uint32_t j = 0; j < deg; j++) { const float mj = slot_metric(adj4, metric, base + j); r += (mj > mi) || (mj == mi && j < i); }
        out.put_rank(metric, base + r, base + i);
    }
}

// Lists of up to RS_SMALL entries (almost all of them): one WARP per node, no CTA barrier.  Canonical slots through the
// same kind of shared-memory hash table.  Ranks: position of slot i in the stable descending order =
// #{j : K[j] > K[i]} with the 64-bit key K = {orderable metric bits, ~index} (all keys of a list are distinct, equal
// metrics keep list order).  Counting every pair costs deg^2 compares (the reference's configs have lists of 300+
// entries); instead 32 sampled keys, sorted in registers, split the list into 33 buckets: b(i) = #{splitters > K[i]}.
// Every key of a lower bucket is greater than every key of a higher one, so
//     position(i) = #{keys in buckets below b(i)} + #{K[j] in bucket b(i) : K[j] > K[i]}
// and only the ~deg/32 keys of the own bucket are compared.  Skewed samples only cost time, never correctness.
// A list with a NaN is ranked with float compares (what `>` does with a NaN is not an order).
constexpr int RB_WARPS = 4, RB_THREADS = RB_WARPS * 32;

__device__ void rank_by_counting_warp(const SlotView adj4, int metric, uint32_t base, uint32_t deg, const RankPtrs &out) {
    for (uint32_t i = threadIdx.x & 31u; i < deg; i += 32) {
        const float mi = slot_metric(adj4, metric, base + i);
        uint32_t r = 0;
        for (uint32_t j = 0; j < deg; j++) { const float mj = slot_metric(adj4, metric, base + j); r += (mj > mi) || (mj == mi && j < i); }
        out.put_rank(metric, base + r, base + i);
    }
}

// GROUP threads work on one node: a warp (GROUP = 32, RB_WARPS nodes per CTA, lists of 1..CAP = 512 entries) or the whole
// CTA (GROUP = 128, lists of LOW + 1..CAP = 513..2048 entries).
template <int GROUP, int CAP>
struct RankSmem {
    static constexpr int GROUPS = RB_THREADS / GROUP;
    unsigned long long key[GROUPS][CAP];   // K[i]; first the canon hash table (keys)
    unsigned long long bkt[GROUPS][CAP];   // keys in bucket order; first the canon hash table (values)
    unsigned long long split[GROUPS][32];
    uint32_t cnt[GROUPS][36];
    int nan[GROUPS];
};

template <int GROUP, int CAP>
__device__ __forceinline__ void rank_bucket_node(RankSmem<GROUP, CAP> &sm, const SlotView adj4, uint32_t metrics_mask, const RankPtrs &rank,
                                                 uint32_t base, uint32_t deg) {
    constexpr int PER = CAP / GROUP;
    static_assert(GROUP == 32 || GROUP == RB_THREADS, "a warp or the CTA per node");
    const uint32_t t = threadIdx.x % GROUP, w = threadIdx.x / GROUP, lane = threadIdx.x & 31u;
    auto group_sync = [] { if (GROUP == 32) __syncwarp(); else __syncthreads(); };
    auto &s_key = sm.key; auto &s_bkt = sm.bkt; auto &s_split = sm.split; auto &s_cnt = sm.cnt; auto &s_nan = sm.nan;
    uint32_t npad = 32;
    while (npad < deg) npad <<= 1;

    {   // canonical slots
        const uint32_t T = 2 * npad, shift = 32 - (31 - __clz(T));
        uint32_t *hkey = reinterpret_cast<uint32_t *>(s_key[w]), *hval = reinterpret_cast<uint32_t *>(s_bkt[w]);
        for (uint32_t i = t; i < T; i += GROUP) { hkey[i] = 0xffffffffu; hval[i] = 0xffffffffu; }
        group_sync();
        for (uint32_t i = t; i < deg; i += GROUP) {
            const uint32_t nbr = adj4.es[base + i].x;
            uint32_t h = (nbr * 2654435761u) >> shift;
            for (;;) {
                const uint32_t prev = atomicCAS(&hkey[h], 0xffffffffu, nbr);
                if (prev == 0xffffffffu || prev == nbr) { atomicMin(&hval[h], i); break; }
                h = (h + 1) & (T - 1);
            }
        }
        group_sync();
        for (uint32_t i = t; i < deg; i += GROUP) {
            const uint32_t nbr = adj4.es[base + i].x;
            uint32_t h = (nbr * 2654435761u) >> shift;
            while (hkey[h] != nbr) h = (h + 1) & (T - 1);
            rank.put_canon(base + i, base + hval[h]);
        }
    }

    unsigned long long *key = s_key[w], *bkt = s_bkt[w], *split = s_split[w];
    uint32_t *cnt = s_cnt[w];
    for (int m = 0; m < 4; m++) {
        if (!(metrics_mask & (1u << m))) continue;
        group_sync();   // the table / the previous metric's arrays are no longer read
        bool nan = false;
        for (uint32_t i = t; i < deg; i += GROUP) key[i] = rank_key(slot_metric(adj4, m, base + i), i, nan);
        if (t < 32) cnt[t] = 0;
        if (t < 2) cnt[32 + t] = 0;
        if (t == 0) s_nan[w] = 0;
        group_sync();
        if (nan) s_nan[w] = 1;
        group_sync();
        if (s_nan[w]) {
            if (GROUP == 32) rank_by_counting_warp(adj4, m, base, deg, rank);
            else rank_by_counting(adj4, m, base, deg, rank);
            continue;
        }
        uint32_t *out = reinterpret_cast<uint32_t *>(bkt);   // positions are staged here once the buckets have been read
        if (deg <= 64 && GROUP == 32) {   // short list: compare with everything
            uint32_t r[2];
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const uint32_t i = t + 32 * e;
                r[e] = 0;
                if (i < deg) {
                    const unsigned long long ki = key[i];
                    for (uint32_t j = 0; j < deg; j++) r[e] += key[j] > ki;
                }
            }
            group_sync();
#pragma unroll
            for (int e = 0; e < 2; e++) if (t + 32 * e < deg) out[r[e]] = t + 32 * e;
        } else {
            // 32 samples at equal distances, sorted descending in registers (bitonic network over the lanes of the first warp)
            if (t < 32) {
                unsigned long long sp = key[static_cast<uint32_t>((static_cast<uint64_t>(lane) * deg) >> 5)];
#pragma unroll
                for (uint32_t k = 2; k <= 32; k <<= 1) {
#pragma unroll
                    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                        const unsigned long long other = __shfl_xor_sync(0xffffffffu, sp, j);
                        const bool lower_lane = (lane & j) == 0, desc = (lane & k) == 0;
                        const bool take_max = lower_lane == desc;
                        sp = take_max ? (sp > other ? sp : other) : (sp < other ? sp : other);
                    }
                }
                split[lane] = sp;
            }
            group_sync();
            // bucket of every key, place inside the bucket in arrival order
            uint32_t bk[PER], at[PER];
#pragma unroll
            for (int e = 0; e < PER; e++) {
                const uint32_t i = t + GROUP * e;
                bk[e] = 0; at[e] = 0;
                if (i < deg) {
                    const unsigned long long ki = key[i];
                    uint32_t lo = 0, hi = 32;
#pragma unroll
                    for (int it = 0; it < 5; it++) { const uint32_t mid = (lo + hi) >> 1; if (split[mid] > ki) lo = mid + 1; else hi = mid; }
                    if (lo < 32 && split[lo] > ki) lo++;
                    bk[e] = lo;
                    at[e] = atomicAdd(&cnt[lo], 1u);
                }
            }
            group_sync();
            // exclusive scan of the 33 bucket sizes (first warp)
            if (t < 32) {
                const uint32_t c0 = cnt[lane], c32 = cnt[32];
                uint32_t incl = c0;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= static_cast<uint32_t>(d)) incl += up; }
                __syncwarp();
                cnt[lane] = incl - c0;
                if (lane == 31) { cnt[32] = incl; cnt[33] = incl + c32; }
            }
            group_sync();
#pragma unroll
            for (int e = 0; e < PER; e++) {
                const uint32_t i = t + GROUP * e;
                if (i < deg) bkt[cnt[bk[e]] + at[e]] = key[i];
            }
            group_sync();
            uint32_t r[PER];
#pragma unroll
            for (int e = 0; e < PER; e++) {
                const uint32_t i = t + GROUP * e;
                r[e] = 0;
                if (i < deg) {
                    const unsigned long long ki = key[i];
                    const uint32_t b0 = cnt[bk[e]], b1 = cnt[bk[e] + 1];
                    uint32_t g = b0;
                    for (uint32_t j = b0; j < b1; j++) g += bkt[j] > ki;
                    r[e] = g;
                }
            }
            group_sync();   // the buckets have been read: their memory takes the positions
#pragma unroll
            for (int e = 0; e < PER; e++) if (t + GROUP * e < deg) out[r[e]] = t + GROUP * e;
        }
        group_sync();
        for (uint32_t p = t; p < deg; p += GROUP) rank.put_rank(m, base + p, base + out[p]);
    }
}

// Lists of LOW + 1 .. CAP entries, one warp per node (node = warp index; a warp whose node belongs to another instance
// leaves at once).  Two instances (CAP 256 and 512): the shorter lists are the majority and need half the shared memory
// and half the unrolled iterations.
template <int CAP, int LOW>
__global__ void __launch_bounds__(RB_THREADS) rank_warp_kernel(const uint32_t *__restrict__ off, const SlotView adj4, uint32_t metrics_mask,
                                                               RankPtrs rank, uint32_t node0, uint32_t nodes) {
    __shared__ __align__(16) RankSmem<32, CAP> sm;
    const uint32_t xi = blockIdx.x * RB_WARPS + (threadIdx.x >> 5);
    if (xi >= nodes) return;
    const uint32_t x = node0 + xi;
    const uint32_t base = off[x], deg = off[x + 1] - base;
    if (deg <= static_cast<uint32_t>(LOW) || deg > static_cast<uint32_t>(CAP)) return;
    rank_bucket_node<32, CAP>(sm, adj4, metrics_mask, rank, base, deg);
}

// The nodes of a chunk of RANK_CHUNK consecutive nodes whose lists have LOW + 1 .. CAP entries, handled one after the other
// by the whole CTA.  The chunk is small: nodes with long lists come in runs (reads of one repeat copy are numbered
// consecutively), and a CTA that had to work through 128 of them was the critical path of the whole stage (4 ms whatever
// the number of nodes).
constexpr int RANK_CHUNK = 8;
template <int LOW, int CAP>
__device__ __forceinline__ uint32_t select_long_nodes(const uint32_t *__restrict__ off, uint32_t node0, uint32_t nodes, uint32_t *s_nodes, uint32_t *s_n) {
    if (threadIdx.x == 0) *s_n = 0;
    __syncthreads();
    const uint32_t xi = blockIdx.x * RANK_CHUNK + threadIdx.x;
    if (threadIdx.x < RANK_CHUNK && xi < nodes) {
        const uint32_t d = off[node0 + xi + 1] - off[node0 + xi];
        if (d > static_cast<uint32_t>(LOW) && (CAP == 0 || d <= static_cast<uint32_t>(CAP))) s_nodes[atomicAdd(s_n, 1u)] = node0 + xi;
    }
    __syncthreads();
    return *s_n;
}

// Lists of LOW + 1 .. CAP entries, the whole CTA per node (instances: CAP 1024 and 4096; dynamic shared memory).
constexpr int RS_SMALL = 512, RS_MID = 4096;
template <int CAP, int LOW>
__global__ void __launch_bounds__(RB_THREADS) rank_cta_kernel(const uint32_t *__restrict__ off, const SlotView adj4, uint32_t metrics_mask,
                                                              RankPtrs rank, uint32_t node0, uint32_t nodes) {
    extern __shared__ __align__(16) unsigned char rank_cta_smem[];
    RankSmem<RB_THREADS, CAP> &sm = *reinterpret_cast<RankSmem<RB_THREADS, CAP> *>(rank_cta_smem);
    __shared__ uint32_t s_nodes[RANK_CHUNK], s_n;
    const uint32_t n = select_long_nodes<LOW, CAP>(off, node0, nodes, s_nodes, &s_n);
    for (uint32_t q = 0; q < n; q++) {
        const uint32_t x = s_nodes[q];
        const uint32_t base = off[x], deg = off[x + 1] - base;
        rank_bucket_node<RB_THREADS, CAP>(sm, adj4, metrics_mask, rank, base, deg);
        __syncthreads();
    }
}

// Lists longer than RS_MID entries (hubs): counting with float compares from global memory.
__global__ void __launch_bounds__(RK_THREADS) rank_long_kernel(const uint32_t *__restrict__ off, const SlotView adj4, uint32_t metrics_mask,
                                                               RankPtrs rank, uint32_t node0, uint32_t nodes) {
    __shared__ uint32_t s_nodes[RANK_CHUNK], s_n;
    const uint32_t n = select_long_nodes<RS_MID, 0>(off, node0, nodes, s_nodes, &s_n);
    for (uint32_t q = 0; q < n; q++) {
        const uint32_t x = s_nodes[q];
        const uint32_t base = off[x], deg = off[x + 1] - base;
        for (uint32_t i = threadIdx.x; i < deg; i += RK_THREADS) {
            const uint32_t ni = adj4.es[base + i].x;
            uint32_t j = 0;
            while (adj4.es[base + j].x != ni) j++;
            rank.put_canon(base + i, base + j);
        }
        for (int m = 0; m < 4; m++)
            if (metrics_mask & (1u << m)) rank_by_counting(adj4, m, base, deg, rank);
    }
}


// The finished slices of this rank's node range (canon, rank permutations: contiguous slots [off[n_lo], off[n_hi])) go to
// the same place in every other rank's tables: 16-byte loads from local memory, 16-byte stores into peer memory over
// NVLink.  blockIdx.y = table, blockIdx.z = peer (visited in ring order so that no GPU is everybody's target at once).
struct SlicePush {
    const uint32_t *src[5];
    uint32_t *dst[5][MAX_PEERS];
    int tables, peers;
};
__global__ void __launch_bounds__(256) slice_push_kernel(SlicePush a, const uint32_t *__restrict__ off, uint32_t n_lo, uint32_t n_hi) {
    const uint64_t lo = off[n_lo], hi = off[n_hi];
    const uint32_t *src = a.src[blockIdx.y];
    uint32_t *dst = a.dst[blockIdx.y][blockIdx.z];
    const uint64_t up = (lo + 3) & ~3ull, down = hi & ~3ull;
    const uint64_t head = up < hi ? up : hi, tail = down > head ? down : head;   // [head, tail) is 16-byte aligned on both sides
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x, t = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    for (uint64_t i = lo + t; i < head; i += stride) dst[i] = src[i];
    const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
    uint4 *d4 = reinterpret_cast<uint4 *>(dst);
    uint64_t q = head / 4 + t;
    const uint64_t q_end = tail / 4;
    for (; q + 3 * stride < q_end; q += 4 * stride) {   // four loads in flight per thread
        const uint4 v0 = s4[q], v1 = s4[q + stride], v2 = s4[q + 2 * stride], v3 = s4[q + 3 * stride];
        d4[q] = v0; d4[q + stride] = v1; d4[q + 2 * stride] = v2; d4[q + 3 * stride] = v3;
    }
    for (; q < q_end; q += stride) d4[q] = s4[q];
    for (uint64_t i = tail + t; i < hi; i += stride) dst[i] = src[i];
}

static int bits_for(uint64_t n) {
    int b = 1;
    while ((1ull << b) < n) b++;
    return b;
}

namespace {
// Device time of the phases of a build, printed with HERA_DEBUG_TIMING (events are only created then).
struct PhaseLog {
    bool on;
    cudaStream_t s;
    std::vector<std::pair<const char *, cudaEvent_t>> ev;
    explicit PhaseLog(cudaStream_t st) : on(getenv("HERA_DEBUG_TIMING") != nullptr), s(st) { mark("start"); }
    void mark(const char *name) {
        if (!on) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, s);
        ev.push_back({name, e});
    }
    void print(int rank) {
        if (!on) return;
        cudaEventSynchronize(ev.back().second);
        std::string line;
        for (size_t i = 1; i < ev.size(); i++) {
            float ms = 0;
            cudaEventElapsedTime(&ms, ev[i - 1].second, ev[i].second);
            char buf[64];
            snprintf(buf, sizeof buf, " %s %.3f", ev[i].first, ms);
            line += buf;
        }
        fprintf(stderr, "[hera rank %d] finalize phases (ms):%s\n", rank, line.c_str());
        for (auto &e : ev) cudaEventDestroy(e.second);
        ev.clear();
    }
};

// An event a later point of a scope waits for.  Left early (exception), the scope still makes the stream wait for it
// before the temporaries declared in front of the guard go back to the pool (a kernel on a side stream may read them).
struct PendingEvent {
    cudaStream_t s;
    cudaEvent_t e = nullptr;
    ~PendingEvent() {
        if (!e) return;
        cudaStreamWaitEvent(s, e, 0);
        cudaEventDestroy(e);
    }
};
}  // namespace

void finalize_graph(Ctx &c, uint32_t metrics_mask, hera_graph_info *out) {
    GraphTables &g = c.g;
    const int ranks = xch_size(c), me = xch_rank(c);
    const uint64_t E = c.rows_total, S = 2 * E;        // all ranks
    const uint64_t L = c.kept.n, SL = 2 * L;            // this rank's rows / entries
    HERA_REQUIRE(S < (1ull << 32), 4, "too many kept overlaps for 32-bit slots");
    HERA_REQUIRE(c.own_rows.size() <= MAX_LOADS && c.texts.size() <= 16, 5, "more than 16 loads per graph are not supported");
    HERA_REQUIRE(ranks <= MAX_PEERS, 37, "more than 16 ranks are not supported");
    const RowPieces pc = row_pieces(c);
    const uint32_t P = static_cast<uint32_t>(std::max<size_t>(1, c.own_rows.size()));
    cudaEvent_t ev[4];
    for (auto &e : ev) HERA_CUDA(cudaEventCreate(&e));
    struct EvGuard { cudaEvent_t *e; ~EvGuard() { for (int i = 0; i < 4; i++) cudaEventDestroy(e[i]); } } ev_guard{ev};
    HERA_CUDA(cudaEventRecord(ev[0], c.stream));
    PhaseLog log(c.stream);
    g.ready = false;
    c.gathered.valid = false;
    g.E = E; g.S = S; g.anchors = 0; g.max_deg = 0;
    if (!c.cache_loaded) g.N = 0;
    for (bool &r : g.rank_ready) r = false;
    float intern_ms = 0, csr_ms = 0, rank_ms = 0;
    if (E > 0) {
        TempPool tmp(c.stream);
        uint64_t N = 0;
        uint32_t *d_late = tmp.zeroed<uint32_t>(4);   // [0] collision flag, [1] anchors, [2] max degree: looked at when the build is over
        PendingEvent verify_guard{c.stream};
        cudaEvent_t &verify_done = verify_guard.e;
        TextPtrs tp{};
        for (size_t i = 0; i < c.texts.size(); i++) tp.p[i] = c.texts[i]->data();
        if (!c.cache_loaded) {
            // ---- K3
            uint32_t *occ_slot = tmp.get<uint32_t>(SL);
            unsigned int *d_small = tmp.get<unsigned int>(8);   // [0] distinct names in the table, [1] table overflow, [2] local entries
            // the table has to hold the names of ALL ranks: sized from global quantities so that every rank takes the same decisions
            uint64_t cap = 1024;
            while (cap < 4 * E && cap < (1ull << 22)) cap <<= 1;
            unsigned long long *keys = nullptr;
            uint32_t *minpos = nullptr;
            uint64_t T = 0;                    // entries of all ranks (>= N)
            std::vector<uint64_t> ecount(ranks), edispl(ranks);
            c.h_counters.reserve(64 + ranks);
            for (;;) {
                keys = tmp.get<unsigned long long>(cap);
                minpos = tmp.get<uint32_t>(cap);
                HERA_CUDA(cudaMemsetAsync(keys, 0, cap * sizeof(unsigned long long), c.stream));
                HERA_CUDA(cudaMemsetAsync(minpos, 0xff, cap * sizeof(uint32_t), c.stream));
                HERA_CUDA(cudaMemsetAsync(d_small, 0, 8 * sizeof(unsigned int), c.stream));
                if (SL) HERA_LAUNCH(c, intern_insert_kernel, div_up(SL, 256), 256, 0, SL, pc, c.kept.qhash.p, c.kept.thash.p, keys, minpos, cap - 1, occ_slot, d_small, d_small + 1);
                HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 32, d_small, 2 * sizeof(unsigned int), cudaMemcpyDeviceToHost, c.stream));
                HERA_CUDA(cudaStreamSynchronize(c.stream));
                const unsigned int *hs = reinterpret_cast<const unsigned int *>(c.h_counters.p + 32);
                uint64_t mine[2] = {hs[0], hs[1]};
                std::vector<uint64_t> all(2 * static_cast<size_t>(ranks), 0);
                if (ranks > 1) c.xch->all_gather_host(c, mine, 2, all.data());
                else { all[0] = mine[0]; all[1] = mine[1]; }
                bool overflow = false;
                T = 0;
                for (int r = 0; r < ranks; r++) { ecount[r] = all[2 * r]; edispl[r] = T; T += ecount[r]; overflow = overflow || all[2 * r + 1] != 0; }
                if (!overflow && T * 2 <= cap) break;
                tmp.release(keys);
                tmp.release(minpos);
                while (cap < 4 * T) cap <<= 1;   // table too full: probe chains get long, rebuild larger
                if (overflow) cap <<= 1;
            }
            log.mark("insert");
            // this rank's distinct names -> entries, all ranks' entries -> every table
            const uint64_t mine_n = ecount[me];
            uint64_t *ent_hash = tmp.get<uint64_t>(T);
            uint32_t *ent_pos = tmp.get<uint32_t>(T), *ent_len = tmp.get<uint32_t>(T), *ent_meta = tmp.get<uint32_t>(T);
            uint64_t *loc_hash = ranks > 1 ? tmp.get<uint64_t>(mine_n) : ent_hash;
            uint32_t *loc_pos = ranks > 1 ? tmp.get<uint32_t>(mine_n) : ent_pos, *loc_len = ranks > 1 ? tmp.get<uint32_t>(mine_n) : ent_len,
                     *loc_meta = ranks > 1 ? tmp.get<uint32_t>(mine_n) : ent_meta;
            HERA_LAUNCH(c, intern_entries_kernel, div_up(cap, 256), 256, 0, cap, keys, minpos, pc, c.kept.qlen.p, c.kept.tlen.p, c.kept.flags.p, c.kept.qname.p,
                        c.kept.tname.p, d_small + 2, loc_hash, loc_pos, loc_len, loc_meta);
            if (ranks > 1) {
                const Exchange::Column cols[4] = {{ent_hash, 8, loc_hash}, {ent_pos, 4, loc_pos}, {ent_len, 4, loc_len}, {ent_meta, 4, loc_meta}};
                c.xch->gather_rows(c, cols, 4, 0, 0, ecount.data(), edispl.data());
                HERA_LAUNCH(c, intern_merge_kernel, div_up(T, 256), 256, 0, T, ent_hash, ent_pos, keys, minpos, cap - 1, d_small, d_small + 1);
            }
            log.mark("merge");
            const uint64_t words = (S + 31) / 32;
            uint32_t *first_bits = tmp.zeroed<uint32_t>(words), *word_prefix = tmp.get<uint32_t>(words + 1);
            uint32_t *slot_len = tmp.get<uint32_t>(cap), *slot_meta = tmp.get<uint32_t>(cap), *slot_id = tmp.get<uint32_t>(cap);
            uint32_t *name_lens = tmp.zeroed<uint32_t>(T + 1);
            HERA_LAUNCH(c, intern_winner_kernel, div_up(T, 256), 256, 0, T, ent_hash, ent_pos, ent_len, ent_meta, keys, minpos, cap - 1, slot_len, slot_meta, first_bits);
            HERA_LAUNCH(c, popc_kernel, div_up(words, 256), 256, 0, words, first_bits, word_prefix);
            exclusive_scan_u32(c, word_prefix, words, word_prefix, nullptr);
            // node tables: at most T nodes (N is read back together with the size of the name blob)
            g.node_len.reserve(T); g.node_anchor.reserve(T); g.node_pos.reserve(T); g.node_name_off.reserve(T + 2);
            g.e_q.reserve(L); g.e_t.reserve(L);
            HERA_LAUNCH(c, intern_nodes_kernel, div_up(cap, 256), 256, 0, cap, keys, minpos, first_bits, word_prefix, slot_len, slot_meta, slot_id, g.node_len.p,
                        g.node_anchor.p, g.node_pos.p, name_lens);
            if (L) HERA_LAUNCH(c, intern_assign_kernel, div_up(L, 256), 256, 0, L, occ_slot, slot_id, g.e_q.p, g.e_t.p);
            // node names: one blob in node order (lengths behind N are zero, so the offsets behind N all equal the blob size)
            exclusive_scan_u32_to_u64(c, name_lens, T + 1, g.node_name_off.p, g.node_name_off.p + T + 1);
            HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 32, d_small, 2 * sizeof(unsigned int), cudaMemcpyDeviceToHost, c.stream));
            HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 34, g.node_name_off.p + T + 1, sizeof(uint64_t), cudaMemcpyDeviceToHost, c.stream));
            HERA_CUDA(cudaStreamSynchronize(c.stream));
            {
                const unsigned int *hs = reinterpret_cast<const unsigned int *>(c.h_counters.p + 32);
                HERA_REQUIRE(hs[1] == 0, 6, "internal: the interning table overflowed while merging the ranks' names");
                N = hs[0];
            }
            g.N = N;
            const uint64_t blob_bytes = c.h_counters.p[34];
            g.name_bytes = blob_bytes;
            g.node_names.reserve(blob_bytes + 1);
            if (ranks > 1) HERA_CUDA(cudaMemsetAsync(g.node_names.p, 0, blob_bytes + 1, c.stream));
            HERA_LAUNCH(c, name_copy_kernel, div_up(N, 256), 256, 0, N, g.node_pos.p, pc, c.kept.qname.p, c.kept.tname.p, g.node_name_off.p, tp, g.node_names.p);
            if (ranks > 1 && blob_bytes) {
                char *parts = tmp.get<char>(blob_bytes * ranks);
                std::vector<uint64_t> counts(ranks, blob_bytes), displs(ranks);
                for (int r = 0; r < ranks; r++) displs[r] = static_cast<uint64_t>(r) * blob_bytes;
                c.xch->all_gatherv(c, g.node_names.p, parts, counts.data(), displs.data(), 1);
                HERA_LAUNCH(c, name_merge_kernel, div_up(blob_bytes, 256), 256, 0, blob_bytes, ranks, parts, g.node_names.p);
            }
            // Every occurrence's bytes against its node's name: only reads, so it runs beside the CSR build on a side
            // stream; the verdict is looked at together with the last results of this build.
            if (c.aux_streams.empty()) {
                cudaStream_t st;
                HERA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
                c.aux_streams.push_back(st);
            }
            if (L) {
                {
                    Events<1> names_ready(cudaEventDisableTiming);
                    HERA_CUDA(cudaEventRecord(names_ready[0], c.stream));
                    HERA_CUDA(cudaStreamWaitEvent(c.aux_streams[0], names_ready[0], 0));
                }
                intern_verify_kernel<<<std::min<unsigned>(div_up(L, 256), 148u * static_cast<unsigned>(c.opt_verify_ctas_per_sm)), 256, 0, c.aux_streams[0]>>>(L, g.e_q.p, g.e_t.p, c.kept.qname.p, c.kept.tname.p, g.node_name_off.p,
                                                                                  g.node_names.p, tp, d_late);
                c.launches.count++;
                HERA_CUDA(cudaGetLastError());
                HERA_CUDA(cudaEventCreateWithFlags(&verify_done, cudaEventDisableTiming));
                HERA_CUDA(cudaEventRecord(verify_done, c.aux_streams[0]));
            }
            log.mark("nodes");
        } else {
            N = g.N;   // node table, e_q / e_t restored from a cache file (hera_graph_load_cache)
        }
        // anchors in node order
        {
            uint32_t *anchor_scan = tmp.get<uint32_t>(N);
            exclusive_scan_u8_to_u32(c, g.node_anchor.p, N, anchor_scan, d_late + 1);
            g.anchor_ids.reserve(N + 1);
            HERA_LAUNCH(c, anchor_list_kernel, div_up(N, 256), 256, 0, N, g.node_anchor.p, anchor_scan, g.anchor_ids.p);
        }
        HERA_CUDA(cudaEventRecord(ev[1], c.stream));

        // ---- K4
        const uint64_t K = N * P;
        HERA_REQUIRE(K < (1ull << 32), 4, "too many nodes x loads for 32-bit sort keys");
        uint32_t *keys32 = tmp.get<uint32_t>(SL), *vals32 = tmp.get<uint32_t>(SL), *keys_alt = tmp.get<uint32_t>(SL), *vals_alt = tmp.get<uint32_t>(SL);
        uint32_t *run_start = tmp.zeroed<uint32_t>(K), *run_end = tmp.zeroed<uint32_t>(K);
        uint32_t *before = tmp.get<uint32_t>(K), *deg = tmp.get<uint32_t>(N + 1);
        if (SL) {
            HERA_LAUNCH(c, csr_keys_kernel, div_up(SL, 256), 256, 0, L, pc, P, g.e_q.p, g.e_t.p, keys32, vals32);
            radix_sort_pairs(c, keys32, vals32, keys_alt, vals_alt, SL, bits_for(K));
            HERA_LAUNCH(c, csr_runs_kernel, div_up(SL, 256), 256, 0, SL, keys32, run_start, run_end);
        }
        log.mark("sort");
        uint32_t *cnt_all = tmp.get<uint32_t>(K * ranks);
        HERA_LAUNCH(c, csr_counts_kernel, div_up(K, 256), 256, 0, K, run_start, run_end, cnt_all + static_cast<uint64_t>(me) * K);
        if (ranks > 1) {
            std::vector<uint64_t> counts(ranks, K), displs(ranks);
            for (int r = 0; r < ranks; r++) displs[r] = static_cast<uint64_t>(r) * K;
            const Exchange::Column col[1] = {{cnt_all, 4}};
            c.xch->gather_rows(c, col, 1, 0, displs[me], counts.data(), displs.data());
        }
        g.adj_off.reserve(N + 1);
        HERA_LAUNCH(c, csr_degrees_kernel, div_up(N, 256), 256, 0, N, P, ranks, me, cnt_all, deg, before);
        exclusive_scan_u32(c, deg, N, g.adj_off.p, g.adj_off.p + N);
        log.mark("degrees");
        // final tables: every rank's entries go straight into every rank's arrays
        PeerTables pt{};
        pt.n = 1;
        bool merged_by_reduce = false;
        if (ranks > 1) {
            void *p4[MAX_PEERS], *pp[MAX_PEERS];
            const bool mapped = c.xch->shared_buffers(c, 0, (S + 1) * sizeof(uint4), p4) & c.xch->shared_buffers(c, 1, (S + 1) * sizeof(uint32_t), pp);
            g.adj4.borrow(static_cast<uint4 *>(p4[me]), S + 1);
            g.adj_pay.borrow(static_cast<uint32_t *>(pp[me]), S + 1);
            if (mapped) {
                pt.n = ranks;
                for (int r = 0; r < ranks; r++) { pt.adj4[r] = static_cast<uint4 *>(p4[r]); pt.pay[r] = static_cast<uint32_t *>(pp[r]); }
                // nobody may still be reading the previous graph out of these tables: the ranks meet before the first store
                c.xch->all_gather_dev(c, c.counters.p, c.counters.p + 1, 1);
            } else {
                merged_by_reduce = true;   // the ranks fill disjoint slots of zeroed tables, the sum over the ranks is the table
                pt.adj4[0] = g.adj4.p; pt.pay[0] = g.adj_pay.p;
                HERA_CUDA(cudaMemsetAsync(g.adj4.p, 0, S * sizeof(uint4), c.stream));
                HERA_CUDA(cudaMemsetAsync(g.adj_pay.p, 0, S * sizeof(uint32_t), c.stream));
            }
        } else {
            g.adj4.reserve(S + 1);
            g.adj_pay.reserve(S + 1);
            pt.adj4[0] = g.adj4.p; pt.pay[0] = g.adj_pay.p;
        }
        g.node_exact.reserve(N);
        g.a8[0].reserve(S + 1); g.a8[1].reserve(S + 1);
        if (SL) HERA_LAUNCH(c, csr_scatter_kernel, div_up(SL, 256), 256, 0, SL, P, keys32, vals32, run_start, g.adj_off.p, before, pc, g.e_q.p, g.e_t.p, c.kept.qs.p,
                            c.kept.ts.p, c.kept.es.p, c.kept.os.p, g.node_anchor.p, pt);
        if (merged_by_reduce) {
            c.xch->all_reduce_sum_u32(c, reinterpret_cast<uint32_t *>(g.adj4.p), 4 * S);
            c.xch->all_reduce_sum_u32(c, g.adj_pay.p, S);
        } else if (ranks > 1) {
            c.xch->all_gather_dev(c, c.counters.p, c.counters.p + 1, 1);   // stream-ordered rendezvous: every rank's stores have been issued and fenced
        }
        log.mark("scatter");
        HERA_LAUNCH(c, node_exact_kernel, div_up(N * 32, 128), 128, 0, static_cast<uint32_t>(N), g.adj_off.p, g.adj4.p, g.node_exact.p, d_late + 2, g.a8[0].p, g.a8[1].p);
        log.mark("exact");
        HERA_CUDA(cudaEventRecord(ev[2], c.stream));

        // ---- rank permutations + canonical slots: each rank does its share of the nodes and stores the results into every
        // rank's tables (peer memory); when the ranks cannot address each other's memory the slices are all-gathered instead
        c.h_counters.reserve(64 + ranks);
        RankPtrs rp{};
        rp.n = 1;
        bool rank_pushed = false;
        SlicePush push{};
        if (ranks > 1) {
            void *ptrs[5][MAX_PEERS];
            bool mapped = c.xch->shared_buffers(c, 2, (S + 1) * sizeof(uint32_t), ptrs[4]);
            g.canon.borrow(static_cast<uint32_t *>(ptrs[4][me]), S + 1);
            for (int m = 0; m < 4; m++) {
                if (!(metrics_mask & (1u << m))) continue;
                mapped = c.xch->shared_buffers(c, 3 + m, (S + 1) * sizeof(uint32_t), ptrs[m]) && mapped;
                g.rank[m].borrow(static_cast<uint32_t *>(ptrs[m][me]), S + 1);
            }
            if (mapped) {
                rank_pushed = true;
                push.src[0] = static_cast<uint32_t *>(ptrs[4][me]);
                push.tables = 1;
                for (int m = 0; m < 4; m++) if (metrics_mask & (1u << m)) push.src[push.tables++] = static_cast<uint32_t *>(ptrs[m][me]);
                push.peers = ranks - 1;
                for (int k = 1; k < ranks; k++) {   // ring order: rank + 1 first
                    const int r = (me + k) % ranks;
                    int t = 0;
                    push.dst[t++][k - 1] = static_cast<uint32_t *>(ptrs[4][r]);
                    for (int m = 0; m < 4; m++) if (metrics_mask & (1u << m)) push.dst[t++][k - 1] = static_cast<uint32_t *>(ptrs[m][r]);
                }
                c.xch->all_gather_dev(c, c.counters.p, c.counters.p + 1, 1);   // no rank still reads the previous graph's tables
            }
        } else {
            g.canon.reserve(S);
        }
        PendingEvent cuts_guard{c.stream};
        cudaEvent_t &cuts_ready = cuts_guard.e;
        if (ranks > 1 && !rank_pushed) {   // slot ranges of the ranks' node ranges: read now, needed (on the host) once the rank kernels are queued
            uint64_t *d_cut = tmp.get<uint64_t>(ranks + 1);
            HERA_LAUNCH(c, node_cut_kernel, 1, 64, 0, g.adj_off.p, N, ranks, d_cut);
            HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 40, d_cut, (ranks + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, c.stream));
            HERA_CUDA(cudaEventCreateWithFlags(&cuts_ready, cudaEventDisableTiming));
            HERA_CUDA(cudaEventRecord(cuts_ready, c.stream));
        }
        const uint64_t n_lo = N * me / ranks, n_hi = N * (me + 1) / ranks;
        const unsigned my_nodes = static_cast<unsigned>(n_hi - n_lo);
        for (int m = 0; m < 4; m++) {
            if (!(metrics_mask & (1u << m))) continue;
            if (ranks == 1) g.rank[m].reserve(S);
            rp.p[m][0] = g.rank[m].p;
            g.rank_ready[m] = true;
        }
        rp.canon[0] = g.canon.p;
        const SlotView views{g.a8[0].p, g.a8[1].p};
        if (my_nodes) {
            HERA_LAUNCH(c, (rank_warp_kernel<256, 0>), div_up(my_nodes, RB_WARPS), RB_THREADS, 0, g.adj_off.p, views, metrics_mask, rp,
                        static_cast<uint32_t>(n_lo), my_nodes);
            HERA_LAUNCH(c, (rank_warp_kernel<RS_SMALL, 256>), div_up(my_nodes, RB_WARPS), RB_THREADS, 0, g.adj_off.p, views, metrics_mask, rp,
                        static_cast<uint32_t>(n_lo), my_nodes);
            log.mark("rank small");
            constexpr size_t smem1 = sizeof(RankSmem<RB_THREADS, 1024>), smem2 = sizeof(RankSmem<RB_THREADS, RS_MID>);
            HERA_CUDA(cudaFuncSetAttribute(rank_cta_kernel<RS_MID, 1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem2)));
            HERA_LAUNCH(c, (rank_cta_kernel<1024, RS_SMALL>), div_up(my_nodes, RANK_CHUNK), RB_THREADS, smem1, g.adj_off.p, views, metrics_mask, rp,
                        static_cast<uint32_t>(n_lo), my_nodes);
            HERA_LAUNCH(c, (rank_cta_kernel<RS_MID, 1024>), div_up(my_nodes, RANK_CHUNK), RB_THREADS, smem2, g.adj_off.p, views, metrics_mask, rp,
                        static_cast<uint32_t>(n_lo), my_nodes);
            log.mark("rank mid");
            HERA_LAUNCH(c, rank_long_kernel, div_up(my_nodes, RANK_CHUNK), RK_THREADS, 0, g.adj_off.p, views, metrics_mask, rp,
                        static_cast<uint32_t>(n_lo), my_nodes);
        }
        log.mark("rank long");
        if (rank_pushed) {
            if (my_nodes) HERA_LAUNCH(c, slice_push_kernel, dim3(148, push.tables, push.peers), 256, 0, push, g.adj_off.p, static_cast<uint32_t>(n_lo), static_cast<uint32_t>(n_hi));
            log.mark("rank push");
            c.xch->all_gather_dev(c, c.counters.p, c.counters.p + 1, 1);   // stream-ordered rendezvous: every rank's stores have been issued
            log.mark("rank rendezvous");
        } else if (ranks > 1) {
            HERA_CUDA(cudaEventSynchronize(cuts_ready));   // the slice boundaries left the device before the rank kernels started
            cudaEventDestroy(cuts_ready);
            cuts_ready = nullptr;
            const uint64_t *cut = c.h_counters.p + 40;
            std::vector<uint64_t> counts(ranks), displs(ranks);
            for (int r = 0; r < ranks; r++) { displs[r] = cut[r]; counts[r] = cut[r + 1] - cut[r]; }
            Exchange::Column cols[5] = {{g.canon.p, 4}};
            int ncols = 1;
            for (int m = 0; m < 4; m++)
                if (g.rank_ready[m]) cols[ncols++] = {g.rank[m].p, 4};
            c.xch->gather_rows(c, cols, ncols, 0, displs[me], counts.data(), displs.data());
            log.mark("rank gather");
        }
        HERA_CUDA(cudaEventRecord(ev[3], c.stream));
        if (verify_done) {
            HERA_CUDA(cudaStreamWaitEvent(c.stream, verify_done, 0));
            cudaEventDestroy(verify_done);
            verify_done = nullptr;
        }
        HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 36, d_late, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        const uint32_t *late = reinterpret_cast<const uint32_t *>(c.h_counters.p + 36);
        g.max_deg = late[2];
        g.anchors = late[1];
        uint64_t collided = late[0];
        if (ranks > 1) {
            std::vector<uint64_t> all(ranks);
            c.xch->all_gather_host(c, &collided, 1, all.data());
            for (uint64_t v : all) collided |= v;
        }
        HERA_REQUIRE(!collided, 6, "64-bit name hash collision detected while interning node names");
        HERA_CUDA(cudaEventElapsedTime(&intern_ms, ev[0], ev[1]));
        HERA_CUDA(cudaEventElapsedTime(&csr_ms, ev[1], ev[2]));
        HERA_CUDA(cudaEventElapsedTime(&rank_ms, ev[2], ev[3]));
        log.print(me);
        if (getenv("HERA_DEBUG_DEGREES") && me == 0) {   // list-length histogram (powers of two) of the finished graph
            std::vector<uint32_t> off(N + 1);
            HERA_CUDA(cudaMemcpy(off.data(), g.adj_off.p, (N + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost));
            uint64_t hist[33] = {0};
            for (uint64_t x = 0; x < N; x++) { const uint32_t d = off[x + 1] - off[x]; hist[d ? 32 - __builtin_clz(d) : 0]++; }
            std::string line;
            for (int b = 0; b < 33; b++) if (hist[b]) line += " <2^" + std::to_string(b) + ":" + std::to_string(hist[b]);
            fprintf(stderr, "[hera] list lengths of %llu nodes:%s\n", static_cast<unsigned long long>(N), line.c_str());
        }
    }
    if (E == 0) for (int m = 0; m < 4; m++) g.rank_ready[m] = (metrics_mask >> m) & 1u;  // empty permutations
    g.ready = true;
    if (out) {
        out->nodes = g.N; out->anchors = g.anchors; out->edges = g.E; out->slots = g.S;
        out->intern_ms = intern_ms; out->csr_ms = csr_ms; out->rank_ms = rank_ms;
    }
}

// ---- the file-order table of kept overlaps (edges_) -----------------------------------------------------------
// A single rank holds it; on a sharded context every rank holds the rows it parsed, and the first caller that needs the
// whole table (export, cache, scaffold, digest) makes every rank gather it: COLLECTIVE.
RowView rows_view(Ctx &c) {
    KeptTable &k = c.kept;
    GraphTables &g = c.g;
    RowView v{};
    v.n = c.rows_total;
    const int ranks = xch_size(c), me = xch_rank(c);
    if (ranks == 1) {
        v.qs = k.qs.p; v.qe = k.qe.p; v.ts = k.ts.p; v.te = k.te.p; v.qlen = k.qlen.p; v.tlen = k.tlen.p; v.rec = k.rec.p;
        v.os = k.os.p; v.si = k.si.p; v.es = k.es.p; v.flags = k.flags.p; v.e_q = g.e_q.p; v.e_t = g.e_t.p;
        return v;
    }
    auto &G = c.gathered;
    if (!G.valid) {
        HERA_REQUIRE(g.ready, 7, "the kept overlaps of a sharded graph are gathered after hera_graph_finalize");
        const uint64_t E = c.rows_total;
        G.qs.reserve(E + 1); G.qe.reserve(E + 1); G.ts.reserve(E + 1); G.te.reserve(E + 1); G.qlen.reserve(E + 1); G.tlen.reserve(E + 1); G.rec.reserve(E + 1);
        G.e_q.reserve(E + 1); G.e_t.reserve(E + 1); G.os.reserve(E + 1); G.si.reserve(E + 1); G.es.reserve(E + 1); G.flags.reserve(E + 1);
        const RowPieces pc = row_pieces(c);
        for (size_t i = 0; i < c.load_shapes.size(); i++) {
            const LoadShape &sh = c.load_shapes[i];
            if (!sh.total) continue;
            const uint64_t l0 = pc.local0[i];
            const Exchange::Column cols[13] = {{G.qs.p, 4, k.qs.p + l0}, {G.qe.p, 4, k.qe.p + l0}, {G.ts.p, 4, k.ts.p + l0}, {G.te.p, 4, k.te.p + l0},
                                               {G.qlen.p, 4, k.qlen.p + l0}, {G.tlen.p, 4, k.tlen.p + l0}, {G.rec.p, 4, k.rec.p + l0},
                                               {G.e_q.p, 4, g.e_q.p + l0}, {G.e_t.p, 4, g.e_t.p + l0}, {G.os.p, 4, k.os.p + l0}, {G.si.p, 4, k.si.p + l0},
                                               {G.es.p, 4, k.es.p + l0}, {G.flags.p, 1, k.flags.p + l0}};
            c.xch->gather_rows(c, cols, 13, sh.global0, 0, sh.counts.data(), sh.displs.data());
        }
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        c.xch->barrier(c);
        G.valid = true;
        (void)me;
    }
    v.qs = G.qs.p; v.qe = G.qe.p; v.ts = G.ts.p; v.te = G.te.p; v.qlen = G.qlen.p; v.tlen = G.tlen.p; v.rec = G.rec.p;
    v.os = G.os.p; v.si = G.si.p; v.es = G.es.p; v.flags = G.flags.p; v.e_q = G.e_q.p; v.e_t = G.e_t.p;
    return v;
}

// ---- export ---------------------------------------------------------------------------------
__global__ void export_adj_kernel(uint64_t S, const uint32_t *__restrict__ pay, uint32_t *__restrict__ rec, uint8_t *__restrict__ side) {
    const uint64_t s = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (s < S) { rec[s] = pay[s] >> 1; side[s] = pay[s] & 1; }
}

__global__ void export_strand_kernel(uint64_t E, const uint8_t *__restrict__ flags, uint8_t *__restrict__ strand) {
    const uint64_t e = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (e < E) strand[e] = (flags[e] & KF_MINUS) ? 1 : 0;
}

__global__ void widen_kernel(uint64_t n, const uint32_t *__restrict__ in, uint64_t *__restrict__ out) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[i];
}

template <typename T>
static void d2h(Ctx &c, HBuf<T> &dst, const T *src, uint64_t n) {
    dst.reserve(n + 1);
    if (n) HERA_CUDA(cudaMemcpyAsync(dst.p, src, n * sizeof(T), cudaMemcpyDeviceToHost, c.stream));
}

void export_graph(Ctx &c, hera_graph_view *out) {
    GraphTables &g = c.g;
    HERA_REQUIRE(g.ready, 7, "hera_graph_export: call hera_graph_finalize first");
    const uint64_t N = g.N, E = g.E, S = g.S;
    const RowView rv = rows_view(c);   // (collective on a sharded context: every rank gathers the rows of the others)
    auto &m = c.gm;
    *out = hera_graph_view{};
    out->nodes = N; out->edges = E; out->slots = S;
    m.node_name_off.reserve(N + 1); m.adj_off.reserve(N + 1); m.node_names.reserve(1);
    m.node_name_off.p[0] = 0; m.adj_off.p[0] = 0;
    m.canon.reserve(1);
    for (int k = 0; k < 4; k++) if (g.rank_ready[k]) m.rank[k].reserve(1);
    if (E > 0) {
        d2h(c, m.node_len, g.node_len.p, N);
        d2h(c, m.node_anchor, g.node_anchor.p, N);
        // names
        TempPool tmp(c.stream);
        uint64_t *offs = tmp.get<uint64_t>(N + 1);
        d2h(c, m.node_name_off, g.node_name_off.p, N + 1);
        m.node_names.reserve(g.name_bytes + 1);
        if (g.name_bytes) HERA_CUDA(cudaMemcpyAsync(m.node_names.p, g.node_names.p, g.name_bytes, cudaMemcpyDeviceToHost, c.stream));
        // edges
        d2h(c, m.e_q, rv.e_q, E); d2h(c, m.e_t, rv.e_t, E);
        d2h(c, m.e_qs, rv.qs, E); d2h(c, m.e_qe, rv.qe, E); d2h(c, m.e_ts, rv.ts, E); d2h(c, m.e_te, rv.te, E);
        d2h(c, m.e_os, rv.os, E); d2h(c, m.e_si, rv.si, E); d2h(c, m.e_es, rv.es, E);
        uint8_t *tmp8 = tmp.get<uint8_t>(S + 1);
        uint32_t *tmp32 = tmp.get<uint32_t>(S + 1);
        HERA_LAUNCH(c, export_strand_kernel, div_up(E, 256), 256, 0, E, rv.flags, tmp8);
        d2h(c, m.e_strand, tmp8, E);
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        HERA_LAUNCH(c, export_adj_kernel, div_up(S, 256), 256, 0, S, g.adj_pay.p, tmp32, tmp8);
        d2h(c, m.adj_rec, tmp32, S); d2h(c, m.adj_side, tmp8, S);
        HERA_LAUNCH(c, widen_kernel, div_up(N + 1, 256), 256, 0, N + 1, g.adj_off.p, offs);
        d2h(c, m.adj_off, offs, N + 1);
        d2h(c, m.canon, g.canon.p, S);
        for (int k = 0; k < 4; k++) if (g.rank_ready[k]) d2h(c, m.rank[k], g.rank[k].p, S);
        HERA_CUDA(cudaStreamSynchronize(c.stream));
    }
    out->node_len = m.node_len.p; out->node_anchor = m.node_anchor.p; out->node_name_off = m.node_name_off.p; out->node_names = m.node_names.p;
    out->e_q = m.e_q.p; out->e_t = m.e_t.p; out->e_qs = m.e_qs.p; out->e_qe = m.e_qe.p; out->e_ts = m.e_ts.p; out->e_te = m.e_te.p;
    out->e_os = m.e_os.p; out->e_si = m.e_si.p; out->e_es = m.e_es.p; out->e_strand = m.e_strand.p;
    out->adj_off = m.adj_off.p; out->adj_rec = m.adj_rec.p; out->adj_side = m.adj_side.p; out->canon = m.canon.p;
    out->rank_es = g.rank_ready[0] ? m.rank[0].p : nullptr; out->rank_os = g.rank_ready[1] ? m.rank[1].p : nullptr;
    out->rank_es_sqrt = g.rank_ready[2] ? m.rank[2].p : nullptr; out->rank_os_sqrt = g.rank_ready[3] ? m.rank[3].p : nullptr;
}
}  // namespace hera

// ---- binary cache of the parsed graph (the reference has no such format) -----------
// Everything K1-K3 produce: kept overlaps in file order with their node ids, and the node table.  A later
// run with other walk parameters restores it and goes straight to K4 (CSR, ranks) -- the PAF text (42 GB at
// configs[4], bound by host->device copy) is not read again.
namespace hera {

namespace {
constexpr char CACHE_MAGIC[8] = {'H', 'E', 'R', 'A', 'K', 'E', 'P', '2'};
constexpr uint32_t CACHE_VERSION = 2;
struct CacheHeader {
    char magic[8];
    uint32_t version, header_bytes, filter_bytes, reserved;
    uint64_t E, N, name_bytes, payload_bytes, payload_sum;   // payload_sum: sum of the payload's bytes as 64-bit words (tail bytes zero-padded)
    hera_filter_params fp;
};

struct Checksum {
    uint64_t sum = 0, bytes = 0;
    void add(const char *p, size_t n) {
        size_t i = 0;
        for (; i + 8 <= n; i += 8) { uint64_t w; std::memcpy(&w, p + i, 8); sum += w; }
        if (i < n) { uint64_t w = 0; std::memcpy(&w, p + i, n - i); sum += w; }
        bytes += n;
    }
};

template <typename T>
void put_array(Ctx &c, FILE *f, const T *dev, uint64_t n, std::vector<char> &buf, Checksum &cs) {
    buf.resize(n * sizeof(T));
    if (n) {
        HERA_CUDA(cudaMemcpyAsync(buf.data(), dev, n * sizeof(T), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        HERA_REQUIRE(fwrite(buf.data(), sizeof(T), n, f) == n, 50, "cache: short write");
        cs.add(buf.data(), buf.size());
    }
}

template <typename T>
void get_array(Ctx &c, FILE *f, DBuf<T> &dev, uint64_t n, std::vector<char> &buf, Checksum &cs) {
    dev.reserve(n + 1);
    buf.resize(n * sizeof(T));
    if (n) {
        HERA_REQUIRE(fread(buf.data(), sizeof(T), n, f) == n, 51, "cache: file is truncated");
        cs.add(buf.data(), buf.size());
        HERA_CUDA(cudaMemcpyAsync(dev.p, buf.data(), n * sizeof(T), cudaMemcpyHostToDevice, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
    }
}

uint64_t cache_payload_bytes(uint64_t E, uint64_t N, uint64_t name_bytes) {
    return E * (12 * 4 + 1) + N * (4 + 1) + (N + 1) * 8 + name_bytes;
}
}  // namespace

// ids below N, name offsets monotone and inside the blob: a cache that passes its checksum but was written by something
// else must not make later kernels index out of bounds
__global__ void cache_check_kernel(uint64_t E, uint64_t N, uint64_t name_bytes, const uint32_t *__restrict__ e_q, const uint32_t *__restrict__ e_t,
                                   const uint64_t *__restrict__ name_off, unsigned int *bad) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < E && (e_q[i] >= N || e_t[i] >= N)) *bad = 1;
    if (i < N && (name_off[i] > name_off[i + 1] || name_off[i + 1] > name_bytes)) *bad = 1;
    if (i == 0 && (name_off[0] != 0 || name_off[N] != name_bytes)) *bad = 1;
}

void save_graph_cache(Ctx &c, const char *path) {
    GraphTables &g = c.g;
    HERA_REQUIRE(g.ready, 7, "hera_graph_save: call hera_graph_finalize first");
    const RowView k = rows_view(c);
    FILE *f = fopen(path, "wb");
    if (!f) throw Error{1, std::string("Cannot write file: ") + path};
    struct Closer { FILE *f; ~Closer() { fclose(f); } } closer{f};
    CacheHeader h{};
    std::memcpy(h.magic, CACHE_MAGIC, 8);
    h.version = CACHE_VERSION; h.header_bytes = sizeof(CacheHeader); h.filter_bytes = sizeof(hera_filter_params);
    h.E = g.E; h.N = g.N; h.name_bytes = g.name_bytes; h.fp = c.fp;
    h.payload_bytes = cache_payload_bytes(g.E, g.N, g.name_bytes);
    HERA_REQUIRE(fwrite(&h, sizeof(h), 1, f) == 1, 50, "cache: short write");
    std::vector<char> buf;
    Checksum cs;
    put_array(c, f, k.qs, g.E, buf, cs); put_array(c, f, k.qe, g.E, buf, cs); put_array(c, f, k.ts, g.E, buf, cs); put_array(c, f, k.te, g.E, buf, cs);
    put_array(c, f, k.qlen, g.E, buf, cs); put_array(c, f, k.tlen, g.E, buf, cs); put_array(c, f, k.rec, g.E, buf, cs);
    put_array(c, f, k.os, g.E, buf, cs); put_array(c, f, k.si, g.E, buf, cs); put_array(c, f, k.es, g.E, buf, cs); put_array(c, f, k.flags, g.E, buf, cs);
    put_array(c, f, k.e_q, g.E, buf, cs); put_array(c, f, k.e_t, g.E, buf, cs);
    put_array(c, f, g.node_len.p, g.N, buf, cs); put_array(c, f, g.node_anchor.p, g.N, buf, cs);
    put_array(c, f, g.node_name_off.p, g.N + 1, buf, cs); put_array(c, f, g.node_names.p, g.name_bytes, buf, cs);
    HERA_REQUIRE(cs.bytes == h.payload_bytes, 50, "cache: internal size mismatch");
    h.payload_sum = cs.sum;
    HERA_REQUIRE(fseek(f, 0, SEEK_SET) == 0 && fwrite(&h, sizeof(h), 1, f) == 1, 50, "cache: short write");
}

// The filter stored in the file REPLACES the context's filter parameters (the kept overlaps are the result of that
// filter); hera_graph_set_filter after the load has no effect on them.
void load_graph_cache(Ctx &c, const char *path) {
    HERA_REQUIRE(xch_size(c) == 1, 55, "hera_graph_load_cache: a cache is restored on a single rank");
    FILE *f = fopen(path, "rb");
    if (!f) throw Error{1, std::string("Cannot open file: ") + path};
    struct Closer { FILE *f; ~Closer() { fclose(f); } } closer{f};
    CacheHeader h{};
    HERA_REQUIRE(fread(&h, sizeof(h), 1, f) == 1 && std::memcmp(h.magic, CACHE_MAGIC, 8) == 0, 52, "not a hera_b200 graph cache");
    HERA_REQUIRE(h.version == CACHE_VERSION && h.header_bytes == sizeof(CacheHeader) && h.filter_bytes == sizeof(hera_filter_params), 52,
                 "graph cache written by another version of the library");
    HERA_REQUIRE(c.kept.n == 0 && c.texts.empty(), 53, "hera_graph_load_cache needs a fresh graph (hera_graph_reset)");
    HERA_REQUIRE(h.E < (1ull << 31) && h.N <= 2 * h.E && h.N < (1ull << 32), 52, "graph cache: implausible sizes");
    HERA_REQUIRE(fseek(f, 0, SEEK_END) == 0, 52, "graph cache: cannot seek");
    const long file_bytes = ftell(f);
    HERA_REQUIRE(file_bytes >= 0 && h.payload_bytes == cache_payload_bytes(h.E, h.N, h.name_bytes) &&
                 static_cast<uint64_t>(file_bytes) == sizeof(CacheHeader) + h.payload_bytes, 52, "graph cache: sizes in the header do not match the file");
    HERA_REQUIRE(fseek(f, sizeof(CacheHeader), SEEK_SET) == 0, 52, "graph cache: cannot seek");
    GraphTables &g = c.g;
    std::vector<char> buf;
    KeptTable &k = c.kept;
    const uint64_t E = h.E, N = h.N;
    Checksum cs;
    get_array(c, f, k.qs, E, buf, cs); get_array(c, f, k.qe, E, buf, cs); get_array(c, f, k.ts, E, buf, cs); get_array(c, f, k.te, E, buf, cs);
    get_array(c, f, k.qlen, E, buf, cs); get_array(c, f, k.tlen, E, buf, cs); get_array(c, f, k.rec, E, buf, cs);
    get_array(c, f, k.os, E, buf, cs); get_array(c, f, k.si, E, buf, cs); get_array(c, f, k.es, E, buf, cs); get_array(c, f, k.flags, E, buf, cs);
    get_array(c, f, g.e_q, E, buf, cs); get_array(c, f, g.e_t, E, buf, cs);
    get_array(c, f, g.node_len, N, buf, cs); get_array(c, f, g.node_anchor, N, buf, cs);
    get_array(c, f, g.node_name_off, N + 1, buf, cs); get_array(c, f, g.node_names, h.name_bytes, buf, cs);
    HERA_REQUIRE(cs.sum == h.payload_sum, 52, "graph cache: checksum mismatch (truncated or corrupt file)");
    TempPool tmp(c.stream);
    unsigned int *d_bad = tmp.zeroed<unsigned int>(1), h_bad = 0;
    HERA_LAUNCH(c, cache_check_kernel, div_up(std::max<uint64_t>(std::max(E, N), 1), 256), 256, 0, E, N, h.name_bytes, g.e_q.p, g.e_t.p, g.node_name_off.p, d_bad);
    HERA_CUDA(cudaMemcpyAsync(&h_bad, d_bad, sizeof(unsigned int), cudaMemcpyDeviceToHost, c.stream));
    HERA_CUDA(cudaStreamSynchronize(c.stream));
    HERA_REQUIRE(!h_bad, 52, "graph cache: node ids or name offsets out of range");
    k.n = E;
    c.rows_total = E;
    c.own_rows.assign(1, RowRange{0, E});
    LoadShape shape;
    shape.global0 = 0; shape.total = E; shape.counts.assign(1, E); shape.displs.assign(1, 0);
    c.load_shapes.assign(1, shape);
    g.N = N; g.name_bytes = h.name_bytes;
    c.fp = h.fp;
    c.cache_loaded = true;
    c.gathered.valid = false;
    g.ready = false;
}

}  // namespace hera
