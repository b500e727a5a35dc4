// K3 + K4: node interning, CSR adjacency, per-metric rank permutations, canonical slots.
//
// K3 replaces OverlapGraph::nodeIndex and the node creation in buildFrom (reference
// src/OverlapGraph.cpp:208-218, 251-258): node id = rank of the name's first appearance among
// kept records (query before target, rr load before cr load).  Names are interned through an
// open-addressing table keyed by the 64-bit name hash; every occurrence does an atomicMin of its
// position; positions that equal the table minimum are first appearances, and an exclusive scan
// over those flags is the node id.  A verification pass compares every occurrence's bytes with
// its node's name, so a 64-bit hash collision is detected, never silently merged.
//
// K4 replaces the adjacency pushes (src/OverlapGraph.cpp:241-248): each kept overlap e yields
// entry 2e (owner = target, stored as read) and 2e+1 (owner = query, q/t swapped).  A stable LSD
// radix sort by owner keeps record order inside each node's list, which is the reference's
// insertion order.  Rank permutations restate the per-step stable_sort of the greedy walk
// (src/PathManager.cpp:289-296); canon[] its first-parallel-edge lookup (:340-347).
#include <cstdio>
#include <cstring>
#include <vector>

#include "ctx.cuh"
#include "primitives.cuh"

namespace hera {

struct TextPtrs {
    const uint8_t *p[16];
};

__global__ void intern_insert_kernel(uint64_t occ, const uint64_t *__restrict__ qhash, const uint64_t *__restrict__ thash,
                                     unsigned long long *keys, uint32_t *minpos, uint64_t mask, uint32_t *__restrict__ occ_slot,
                                     unsigned int *distinct, unsigned int *overflow) {
    const uint64_t pos = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (pos >= occ) return;
    const uint64_t e = pos >> 1;
    const unsigned long long h = (pos & 1) ? thash[e] : qhash[e];
    uint64_t slot = h & mask;
    for (uint64_t probes = 0; probes <= mask; probes++) {
        unsigned long long cur = keys[slot];
        if (cur == 0) {
            cur = atomicCAS(&keys[slot], 0ull, h);
            if (cur == 0) { atomicAdd(distinct, 1u); cur = h; }
        }
        if (cur == h) {
            atomicMin(&minpos[slot], static_cast<uint32_t>(pos));
            occ_slot[pos] = static_cast<uint32_t>(slot);
            return;
        }
        slot = (slot + 1) & mask;
    }
    *overflow = 1;
}

__global__ void intern_first_flags_kernel(uint64_t occ, const uint32_t *__restrict__ occ_slot, const uint32_t *__restrict__ minpos, uint8_t *__restrict__ first) {
    const uint64_t pos = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (pos >= occ) return;
    first[pos] = minpos[occ_slot[pos]] == pos ? 1 : 0;
}

// First appearances create the node: length and anchor flag of THAT record (src/OverlapGraph.cpp:209-218).
__global__ void intern_create_nodes_kernel(uint64_t occ, const uint8_t *__restrict__ first, const uint32_t *__restrict__ first_scan,
                                           const uint32_t *__restrict__ occ_slot, uint32_t *__restrict__ slot_id, const uint32_t *__restrict__ qlen,
                                           const uint32_t *__restrict__ tlen, const uint8_t *__restrict__ flags, const uint64_t *__restrict__ qname,
                                           const uint64_t *__restrict__ tname, uint32_t *__restrict__ node_len, uint8_t *__restrict__ node_anchor,
                                           uint64_t *__restrict__ node_name, uint32_t *__restrict__ node_row) {
    const uint64_t pos = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (pos >= occ || !first[pos]) return;
    const uint64_t e = pos >> 1;
    const uint32_t id = first_scan[pos];
    slot_id[occ_slot[pos]] = id;
    const uint8_t f = flags[e];
    const bool cr = f & KF_CR, qctg = f & KF_QCTG;
    node_row[id] = static_cast<uint32_t>(e);
    if (pos & 1) {
        node_len[id] = tlen[e];
        node_anchor[id] = (cr && !qctg) ? 1 : 0;   // ContigPosition::TARGET
        node_name[id] = tname[e];
    } else {
        node_len[id] = qlen[e];
        node_anchor[id] = (cr && qctg) ? 1 : 0;    // ContigPosition::QUERY
        node_name[id] = qname[e];
    }
}

// Kept rows a rank parsed itself: only their name references point into texts it holds.
struct OwnRows {
    uint64_t begin[16], end[16];
    int n;
    __device__ bool has(uint64_t e) const {
        for (int i = 0; i < n; i++) if (e >= begin[i] && e < end[i]) return true;
        return false;
    }
};

__global__ void intern_assign_kernel(uint64_t E, const uint32_t *__restrict__ occ_slot, const uint32_t *__restrict__ slot_id, uint32_t *__restrict__ e_q,
                                     uint32_t *__restrict__ e_t) {
    const uint64_t e = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (e >= E) return;
    e_q[e] = slot_id[occ_slot[2 * e]];
    e_t[e] = slot_id[occ_slot[2 * e + 1]];
}

__global__ void name_len_kernel(uint64_t N, const uint64_t *__restrict__ node_name, uint32_t *__restrict__ lens) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < N) lens[i] = name_len(node_name[i]);
}

// Name bytes of the nodes whose creating row this rank parsed (the other ranks fill the rest).
__global__ void name_copy_kernel(uint64_t N, const uint64_t *__restrict__ node_name, const uint32_t *__restrict__ node_row, OwnRows own,
                                 const uint64_t *__restrict__ offs, TextPtrs texts, char *__restrict__ blob) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= N || !own.has(node_row[i])) return;
    const uint64_t r = node_name[i];
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
__global__ void intern_verify_kernel(OwnRows own, int range, const uint32_t *__restrict__ e_q, const uint32_t *__restrict__ e_t,
                                     const uint64_t *__restrict__ qname, const uint64_t *__restrict__ tname, const uint64_t *__restrict__ name_off_tab,
                                     const char *__restrict__ blob, TextPtrs texts, unsigned int *collision) {
    const uint64_t e = own.begin[range] + static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (e >= own.end[range]) return;
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

__global__ void csr_entries_kernel(uint64_t E, const uint32_t *__restrict__ e_q, const uint32_t *__restrict__ e_t, uint32_t *__restrict__ keys,
                                   uint32_t *__restrict__ vals) {
    const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= 2 * E) return;
    const uint64_t e = i >> 1;
    keys[i] = (i & 1) ? e_q[e] : e_t[e];
    vals[i] = static_cast<uint32_t>(i);
}

__global__ void csr_offsets_kernel(uint64_t S, uint64_t N, const uint32_t *__restrict__ keys, uint32_t *__restrict__ off) {
    const uint64_t s = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (s > S) return;
    if (s == S) { off[N] = static_cast<uint32_t>(S); return; }
    if (s == 0 || keys[s] != keys[s - 1]) off[keys[s]] = static_cast<uint32_t>(s);
}

__device__ __forceinline__ float metric_value(int metric, float es, float os);

__global__ void csr_gather_kernel(uint64_t S, const uint32_t *__restrict__ pay, const uint32_t *__restrict__ e_q, const uint32_t *__restrict__ e_t,
                                  const uint32_t *__restrict__ qs, const uint32_t *__restrict__ ts, const float *__restrict__ es,
                                  const float *__restrict__ os, const uint8_t *__restrict__ node_anchor, uint32_t *__restrict__ a_nbr,
                                  float *__restrict__ a_es, float *__restrict__ a_os, int32_t *__restrict__ a_delta, uint4 *__restrict__ adj4) {
    const uint64_t s = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const uint32_t p = pay[s], e = p >> 1, side = p & 1;
    const uint32_t nbr = side ? e_t[e] : e_q[e];
    a_nbr[s] = nbr | (node_anchor[nbr] ? 0x80000000u : 0u);
    a_es[s] = es[e];
    a_os[s] = os[e];
    const int64_t d = side ? static_cast<int64_t>(ts[e]) - static_cast<int64_t>(qs[e]) : static_cast<int64_t>(qs[e]) - static_cast<int64_t>(ts[e]);
    a_delta[s] = static_cast<int32_t>(d);
    adj4[s] = make_uint4(a_nbr[s], __float_as_uint(es[e]), __float_as_uint(os[e]), static_cast<uint32_t>(static_cast<int32_t>(d)));
}

// One warp per node.  The roulette of the Monte-Carlo walk needs the fp64 sum of the candidates'
// metrics in list order (src/PathManager.cpp:63-66,107-113).  If every value of the node's list is a
// non-negative finite fp32 number, hence an integer multiple of q = 2^(smallest exponent - 23), and the
// sum of ALL of them stays below 2^53 * q, then every partial sum over every subset is an integer
// multiple of q below 2^53 * q: exactly representable, so any summation order gives the sequential
// result bit for bit.  Bit m of node_exact[x] records that for metric m.
__global__ void __launch_bounds__(128) node_exact_kernel(uint32_t N, const uint32_t *__restrict__ off, const float *__restrict__ a_es,
                                                         const float *__restrict__ a_os, const uint32_t *__restrict__ a_nbr, uint8_t *__restrict__ node_exact, uint32_t *__restrict__ max_deg) {
    const uint32_t x = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (x >= N) return;
    const uint32_t base = off[x], deg = off[x + 1] - base;
    if (lane == 0) atomicMax(max_deg, deg);
    uint32_t flags = 0;
    bool anchor_nbr = false;   // bit 4: some neighbour is an anchor (only then can a walk step be forced, src/PathManager.cpp:57-60)
    for (uint32_t k = lane; k < deg; k += 32) anchor_nbr |= (a_nbr[base + k] >> 31) != 0;
    if (__any_sync(0xffffffffu, anchor_nbr)) flags |= 16u;
    for (int m = 0; m < 4; m++) {
        bool weird = false;
        int min_exp = 1 << 20;
        double part = 0.0;
        for (uint32_t k = lane; k < deg; k += 32) {
            const float v = metric_value(m, a_es[base + k], a_os[base + k]);
            const uint32_t bits = __float_as_uint(v), ex = (bits >> 23) & 0xffu;
            if ((bits >> 31) || ex == 0xffu) weird = true;
            else if (bits != 0) min_exp = min(min_exp, static_cast<int>(ex ? ex : 1u) - 150);
            part = __dadd_rn(part, static_cast<double>(v));
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            min_exp = min(min_exp, __shfl_xor_sync(0xffffffffu, min_exp, d));
            part = __dadd_rn(part, __shfl_xor_sync(0xffffffffu, part, d));
        }
        weird = __any_sync(0xffffffffu, weird);
        bool exact = !weird;
        if (exact && part > 0.0) {
            const int total_exp = static_cast<int>((__double_as_longlong(part) >> 52) & 0x7ff) - 1023;
            exact = total_exp < min_exp + 53;
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

// K4b, one CTA per node: canonical slots and rank permutations of the node's list, staged in shared memory.
//   canon[i] = first slot of the list with the same neighbour (shared-memory hash table: neighbour -> min index).
//   rank[m]  = stable descending order by metric m, the order the reference recomputes at every greedy step
//              (src/PathManager.cpp:289-296): a bitonic sort of 64-bit keys {orderable metric bits, ~index}; equal
//              metrics keep list order because the index is part of the key (-0.0 is folded into +0.0 first, as
//              `>` sees them).  A list with a NaN metric, or longer than CAP, is ranked by counting instead:
//              position of slot i = #{j : m[j] > m[i]} + #{j < i : m[j] == m[i]}  (what `>` on NaN gives).
// This kernel takes the lists longer than LOW entries (the short ones go to rank_small_kernel below).
struct RankPtrs { uint32_t *p[4]; };

__device__ __forceinline__ uint64_t rank_key(float v, uint32_t i, bool &nan) {
    uint32_t bits = __float_as_uint(v);
    if (v != v) nan = true;
    if (bits == 0x80000000u) bits = 0;
    const uint32_t u = (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
    return (static_cast<uint64_t>(u) << 32) | (0xffffffffu - i);
}

__device__ void rank_by_counting(const float *__restrict__ a_es, const float *__restrict__ a_os, int metric, uint32_t base, uint32_t deg,
                                 uint32_t *__restrict__ rank) {
    for (uint32_t i = threadIdx.x; i < deg; i += RK_THREADS) {
        const float mi = metric_value(metric, a_es[base + i], a_os[base + i]);
        uint32_t r = 0;
        for (uint32_t j = 0; j < deg; j++) { const float mj = metric_value(metric, a_es[base + j], a_os[base + j]); r += (mj > mi) || (mj == mi && j < i); }
        rank[base + r] = base + i;
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
constexpr int RS_SMALL = 512;
constexpr int RB_WARPS = 4, RB_THREADS = RB_WARPS * 32;

__device__ void rank_by_counting_warp(const float *__restrict__ a_es, const float *__restrict__ a_os, int metric, uint32_t base, uint32_t deg,
                                      uint32_t *__restrict__ rank) {
    for (uint32_t i = threadIdx.x & 31u; i < deg; i += 32) {
        const float mi = metric_value(metric, a_es[base + i], a_os[base + i]);
        uint32_t r = 0;
        for (uint32_t j = 0; j < deg; j++) { const float mj = metric_value(metric, a_es[base + j], a_os[base + j]); r += (mj > mi) || (mj == mi && j < i); }
        rank[base + r] = base + i;
    }
}

__global__ void __launch_bounds__(RB_THREADS) rank_small_kernel(const uint32_t *__restrict__ off, const uint32_t *__restrict__ a_nbr,
                                                                const float *__restrict__ a_es, const float *__restrict__ a_os, uint32_t metrics_mask,
                                                                RankPtrs rank, uint32_t *__restrict__ canon, uint32_t node0, uint32_t nodes) {
    __shared__ __align__(16) unsigned long long s_key[RB_WARPS][RS_SMALL];   // K[i]; first the canon hash table (keys)
    __shared__ __align__(16) unsigned long long s_bkt[RB_WARPS][RS_SMALL];   // keys in bucket order; first the canon hash table (values)
    __shared__ unsigned long long s_split[RB_WARPS][32];
    __shared__ uint32_t s_cnt[RB_WARPS][36];
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint32_t xi = blockIdx.x * RB_WARPS + w;
    if (xi >= nodes) return;
    const uint32_t x = node0 + xi;
    const uint32_t base = off[x], deg = off[x + 1] - base;
    if (deg == 0 || deg > RS_SMALL) return;   // longer lists: rank_canon_kernel
    uint32_t npad = 32;
    while (npad < deg) npad <<= 1;

    {   // canonical slots
        const uint32_t T = 2 * npad, shift = 32 - (31 - __clz(T));
        uint32_t *hkey = reinterpret_cast<uint32_t *>(s_key[w]), *hval = reinterpret_cast<uint32_t *>(s_bkt[w]);
        for (uint32_t i = lane; i < T; i += 32) { hkey[i] = 0xffffffffu; hval[i] = 0xffffffffu; }
        __syncwarp();
        for (uint32_t i = lane; i < deg; i += 32) {
            const uint32_t nbr = a_nbr[base + i];
            uint32_t h = (nbr * 2654435761u) >> shift;
            for (;;) {
                const uint32_t prev = atomicCAS(&hkey[h], 0xffffffffu, nbr);
                if (prev == 0xffffffffu || prev == nbr) { atomicMin(&hval[h], i); break; }
                h = (h + 1) & (T - 1);
            }
        }
        __syncwarp();
        for (uint32_t i = lane; i < deg; i += 32) {
            const uint32_t nbr = a_nbr[base + i];
            uint32_t h = (nbr * 2654435761u) >> shift;
            while (hkey[h] != nbr) h = (h + 1) & (T - 1);
            canon[base + i] = base + hval[h];
        }
    }

    unsigned long long *key = s_key[w], *bkt = s_bkt[w], *split = s_split[w];
    uint32_t *cnt = s_cnt[w];
    for (int m = 0; m < 4; m++) {
        if (!(metrics_mask & (1u << m))) continue;
        __syncwarp();   // the table / the previous metric's arrays are no longer read
        bool nan = false;
        for (uint32_t i = lane; i < deg; i += 32) key[i] = rank_key(metric_value(m, a_es[base + i], a_os[base + i]), i, nan);
        if (lane < 34) cnt[lane] = 0;
        if (lane < 2) cnt[32 + lane] = 0;
        __syncwarp();
        if (__any_sync(0xffffffffu, nan)) { rank_by_counting_warp(a_es, a_os, m, base, deg, rank.p[m]); continue; }
        uint32_t *out = reinterpret_cast<uint32_t *>(bkt);   // positions are staged here once the buckets have been read
        if (deg <= 64) {   // short list: compare with everything
            uint32_t r[2];
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const uint32_t i = lane + 32 * e;
                r[e] = 0;
                if (i < deg) {
                    const unsigned long long ki = key[i];
                    for (uint32_t j = 0; j < deg; j++) r[e] += key[j] > ki;
                }
            }
            __syncwarp();
#pragma unroll
            for (int e = 0; e < 2; e++) if (lane + 32 * e < deg) out[r[e]] = lane + 32 * e;
        } else {
            // 32 samples at equal distances, sorted descending in registers (bitonic network over the lanes)
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
            __syncwarp();
            // bucket of every key, place inside the bucket in arrival order
            constexpr int PER = RS_SMALL / 32;
            uint32_t bk[PER], at[PER];
#pragma unroll
            for (int e = 0; e < PER; e++) {
                const uint32_t i = lane + 32 * e;
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
            __syncwarp();
            // exclusive scan of the 33 bucket sizes
            const uint32_t c0 = cnt[lane], c32 = cnt[32];
            uint32_t incl = c0;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= static_cast<uint32_t>(d)) incl += up; }
            __syncwarp();
            cnt[lane] = incl - c0;
            if (lane == 31) { cnt[32] = incl; cnt[33] = incl + c32; }
            __syncwarp();
#pragma unroll
            for (int e = 0; e < PER; e++) {
                const uint32_t i = lane + 32 * e;
                if (i < deg) bkt[cnt[bk[e]] + at[e]] = key[i];
            }
            __syncwarp();
            uint32_t r[PER];
#pragma unroll
            for (int e = 0; e < PER; e++) {
                const uint32_t i = lane + 32 * e;
                r[e] = 0;
                if (i < deg) {
                    const unsigned long long ki = key[i];
                    const uint32_t b0 = cnt[bk[e]], b1 = cnt[bk[e] + 1];
                    uint32_t g = b0;
                    for (uint32_t j = b0; j < b1; j++) g += bkt[j] > ki;
                    r[e] = g;
                }
            }
            __syncwarp();   // the buckets have been read: their memory takes the positions
#pragma unroll
            for (int e = 0; e < PER; e++) if (lane + 32 * e < deg) out[r[e]] = lane + 32 * e;
        }
        __syncwarp();
        for (uint32_t p = lane; p < deg; p += 32) rank.p[m][base + p] = base + out[p];
    }
}

template <int CAP, int LOW, bool LAST>
__global__ void __launch_bounds__(RK_THREADS) rank_canon_kernel(const uint32_t *__restrict__ off, const uint32_t *__restrict__ a_nbr,
                                                                const float *__restrict__ a_es, const float *__restrict__ a_os, uint32_t metrics_mask,
                                                                RankPtrs rank, uint32_t *__restrict__ canon, uint32_t node0) {
    __shared__ __align__(8) uint32_t s_words[CAP * 4];   // hash table (2 CAP keys + 2 CAP values), then CAP sort keys of 64 bits
    __shared__ int s_nan;
    const uint32_t x = blockIdx.x + node0, tid = threadIdx.x;
    const uint32_t base = off[x], deg = off[x + 1] - base;
    if (deg <= LOW || deg == 0 || (!LAST && deg > CAP)) return;   // the other instance's node
    const bool fits = deg <= CAP;
    uint32_t npad = 32;
    while (npad < deg) npad <<= 1;

    // ---- canonical slots
    if (fits) {
        const uint32_t T = 2 * npad, shift = 32 - (31 - __clz(T));
        uint32_t *hkey = s_words, *hval = s_words + 2 * CAP;
        for (uint32_t i = tid; i < T; i += RK_THREADS) { hkey[i] = 0xffffffffu; hval[i] = 0xffffffffu; }
        __syncthreads();
        for (uint32_t i = tid; i < deg; i += RK_THREADS) {
            const uint32_t nbr = a_nbr[base + i];
            uint32_t h = (nbr * 2654435761u) >> shift;
            for (;;) {
                const uint32_t prev = atomicCAS(&hkey[h], 0xffffffffu, nbr);
                if (prev == 0xffffffffu || prev == nbr) { atomicMin(&hval[h], i); break; }
                h = (h + 1) & (T - 1);
            }
        }
        __syncthreads();
        for (uint32_t i = tid; i < deg; i += RK_THREADS) {
            const uint32_t nbr = a_nbr[base + i];
            uint32_t h = (nbr * 2654435761u) >> shift;
            while (hkey[h] != nbr) h = (h + 1) & (T - 1);
            canon[base + i] = base + hval[h];
        }
    } else {
        for (uint32_t i = tid; i < deg; i += RK_THREADS) {
            const uint32_t ni = a_nbr[base + i];
            uint32_t j = 0;
            while (a_nbr[base + j] != ni) j++;
            canon[base + i] = base + j;
        }
    }

    // ---- rank permutations
    uint64_t *key = reinterpret_cast<uint64_t *>(s_words);
    for (int m = 0; m < 4; m++) {
        if (!(metrics_mask & (1u << m))) continue;
        if (!fits) { rank_by_counting(a_es, a_os, m, base, deg, rank.p[m]); continue; }
        __syncthreads();   // the table / the previous metric's keys are no longer read
        if (tid == 0) s_nan = 0;
        __syncthreads();
        bool nan = false;
        for (uint32_t i = tid; i < npad; i += RK_THREADS)
            key[i] = i < deg ? rank_key(metric_value(m, a_es[base + i], a_os[base + i]), i, nan) : 0ull;
        if (nan) s_nan = 1;
        __syncthreads();
        if (s_nan) { rank_by_counting(a_es, a_os, m, base, deg, rank.p[m]); continue; }
        // Pair t of a step with distance j <= 32 lies in [64 (t / 32), 64 (t / 32) + 64): such steps only need the
        // warp to agree, a step with a longer distance (and the step after it) needs the whole CTA.
        bool wide = false;
        for (uint32_t k = 2; k <= npad; k <<= 1) {
            for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                if (j >= 64 || wide) __syncthreads(); else __syncwarp();
                wide = j >= 64;
                for (uint32_t t = tid; t < npad / 2; t += RK_THREADS) {
                    const uint32_t idx = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                    const uint64_t ka = key[idx], kb = key[idx + j];
                    const bool first_larger = (idx & k) == 0;   // descending blocks alternate with ascending ones; the last merge is descending
                    if ((ka < kb) == first_larger) { key[idx] = kb; key[idx + j] = ka; }
                }
            }
        }
        __syncthreads();
        for (uint32_t p = tid; p < deg; p += RK_THREADS) rank.p[m][base + p] = base + (0xffffffffu - static_cast<uint32_t>(key[p]));
    }
}


static int bits_for(uint64_t n) {
    int b = 1;
    while ((1ull << b) < n) b++;
    return b;
}

void finalize_graph(Ctx &c, uint32_t metrics_mask, hera_graph_info *out) {
    GraphTables &g = c.g;
    const uint64_t E = c.kept.n, occ = 2 * E;
    HERA_REQUIRE(occ < (1ull << 32), 4, "too many kept overlaps for 32-bit slots");
    cudaEvent_t ev[4];
    for (auto &e : ev) HERA_CUDA(cudaEventCreate(&e));
    HERA_CUDA(cudaEventRecord(ev[0], c.stream));
    g.ready = false;
    g.E = E; g.S = occ; g.anchors = 0; g.max_deg = 0;
    if (!c.cache_loaded) g.N = 0;
    for (bool &r : g.rank_ready) r = false;
    float intern_ms = 0, csr_ms = 0, rank_ms = 0;
    if (E > 0) {
        uint64_t N = 0;
        uint32_t *d_late = nullptr;          // results that are only looked at when the build is over
        cudaEvent_t verify_done = nullptr;
        if (!c.cache_loaded) {
        // ---- K3
        uint32_t *occ_slot = nullptr, *first_scan = nullptr, *d_small = nullptr;
        uint8_t *first = nullptr;
        HERA_CUDA(cudaMallocAsync(&occ_slot, occ * sizeof(uint32_t), c.stream));
        HERA_CUDA(cudaMallocAsync(&first_scan, occ * sizeof(uint32_t), c.stream));
        HERA_CUDA(cudaMallocAsync(&first, occ, c.stream));
        HERA_CUDA(cudaMallocAsync(&d_small, 8 * sizeof(uint32_t), c.stream));
        uint64_t cap = 1024;
        while (cap < 4 * E && cap < (1ull << 22)) cap <<= 1;
        unsigned long long *keys = nullptr;
        uint32_t *minpos = nullptr, *slot_id = nullptr;
        uint32_t h_small[8];
        for (;;) {
            HERA_CUDA(cudaMallocAsync(&keys, cap * sizeof(unsigned long long), c.stream));
            HERA_CUDA(cudaMallocAsync(&minpos, cap * sizeof(uint32_t), c.stream));
            HERA_CUDA(cudaMemsetAsync(keys, 0, cap * sizeof(unsigned long long), c.stream));
            HERA_CUDA(cudaMemsetAsync(minpos, 0xff, cap * sizeof(uint32_t), c.stream));
            HERA_CUDA(cudaMemsetAsync(d_small, 0, 8 * sizeof(uint32_t), c.stream));
            HERA_LAUNCH(c, intern_insert_kernel, div_up(occ, 256), 256, 0, occ, c.kept.qhash.p, c.kept.thash.p, keys, minpos, cap - 1, occ_slot, d_small, d_small + 1);
            HERA_CUDA(cudaMemcpyAsync(h_small, d_small, sizeof(h_small), cudaMemcpyDeviceToHost, c.stream));
            HERA_CUDA(cudaStreamSynchronize(c.stream));
            if (!h_small[1] && static_cast<uint64_t>(h_small[0]) * 2 <= cap) break;
            HERA_CUDA(cudaFreeAsync(keys, c.stream));
            HERA_CUDA(cudaFreeAsync(minpos, c.stream));
            cap <<= 2;  // table too full: probe chains get long, rebuild four times larger
        }
        N = h_small[0];
        g.N = N;
        HERA_CUDA(cudaMallocAsync(&slot_id, cap * sizeof(uint32_t), c.stream));
        HERA_LAUNCH(c, intern_first_flags_kernel, div_up(occ, 256), 256, 0, occ, occ_slot, minpos, first);
        exclusive_scan_u8_to_u32(c, first, occ, first_scan, nullptr);
        g.node_len.reserve(N); g.node_anchor.reserve(N); g.node_name.reserve(N); g.node_row.reserve(N); g.node_name_off.reserve(N + 1);
        g.e_q.reserve(E); g.e_t.reserve(E);
        HERA_LAUNCH(c, intern_create_nodes_kernel, div_up(occ, 256), 256, 0, occ, first, first_scan, occ_slot, slot_id, c.kept.qlen.p, c.kept.tlen.p,
                    c.kept.flags.p, c.kept.qname.p, c.kept.tname.p, g.node_len.p, g.node_anchor.p, g.node_name.p, g.node_row.p);
        HERA_LAUNCH(c, intern_assign_kernel, div_up(E, 256), 256, 0, E, occ_slot, slot_id, g.e_q.p, g.e_t.p);
        // node names: one blob in node order.  A rank holds the text of the rows it parsed; the blob is
        // completed by gathering every rank's part (names are short: N * ~10 bytes).
        TextPtrs tp{};
        HERA_REQUIRE(c.texts.size() <= 16 && c.own_rows.size() <= 16, 5, "more than 16 loads per graph are not supported");
        for (size_t i = 0; i < c.texts.size(); i++) tp.p[i] = c.texts[i]->data();
        OwnRows own{};
        own.n = static_cast<int>(c.own_rows.size());
        for (int i = 0; i < own.n; i++) { own.begin[i] = c.own_rows[i].begin; own.end[i] = c.own_rows[i].end; }
        {
            uint32_t *lens = nullptr;
            HERA_CUDA(cudaMallocAsync(&lens, (N + 1) * sizeof(uint32_t), c.stream));
            HERA_LAUNCH(c, name_len_kernel, div_up(N, 256), 256, 0, N, g.node_name.p, lens);
            exclusive_scan_u32_to_u64(c, lens, N, g.node_name_off.p, g.node_name_off.p + N);
            uint64_t blob_bytes = 0;
            HERA_CUDA(cudaMemcpyAsync(&blob_bytes, g.node_name_off.p + N, sizeof(uint64_t), cudaMemcpyDeviceToHost, c.stream));
            HERA_CUDA(cudaStreamSynchronize(c.stream));
            HERA_CUDA(cudaFreeAsync(lens, c.stream));
            g.name_bytes = blob_bytes;
            g.node_names.reserve(blob_bytes + 1);
            const int ranks = xch_size(c);
            if (ranks > 1) HERA_CUDA(cudaMemsetAsync(g.node_names.p, 0, blob_bytes + 1, c.stream));
            HERA_LAUNCH(c, name_copy_kernel, div_up(N, 256), 256, 0, N, g.node_name.p, g.node_row.p, own, g.node_name_off.p, tp, g.node_names.p);
            if (ranks > 1 && blob_bytes) {
                char *parts = nullptr;
                HERA_CUDA(cudaMallocAsync(&parts, blob_bytes * ranks, c.stream));
                std::vector<uint64_t> counts(ranks, blob_bytes), displs(ranks);
                for (int r = 0; r < ranks; r++) displs[r] = static_cast<uint64_t>(r) * blob_bytes;
                c.xch->all_gatherv(c, g.node_names.p, parts, counts.data(), displs.data(), 1);
                HERA_LAUNCH(c, name_merge_kernel, div_up(blob_bytes, 256), 256, 0, blob_bytes, ranks, parts, g.node_names.p);
                HERA_CUDA(cudaFreeAsync(parts, c.stream));
            }
        }
        // Every occurrence's bytes against its node's name: only reads, so it runs beside the CSR build on a side
        // stream; the verdict is looked at together with the last results of this build.
        HERA_CUDA(cudaMallocAsync(&d_late, 2 * sizeof(uint32_t), c.stream));   // [0] collision flag, [1] anchors
        HERA_CUDA(cudaMemsetAsync(d_late, 0, 2 * sizeof(uint32_t), c.stream));
        if (c.aux_streams.empty()) {
            cudaStream_t st;
            HERA_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
            c.aux_streams.push_back(st);
        }
        {
            cudaEvent_t names_ready;
            HERA_CUDA(cudaEventCreateWithFlags(&names_ready, cudaEventDisableTiming));
            HERA_CUDA(cudaEventRecord(names_ready, c.stream));
            HERA_CUDA(cudaStreamWaitEvent(c.aux_streams[0], names_ready, 0));
            cudaEventDestroy(names_ready);
            for (int i = 0; i < own.n; i++) {
                const uint64_t rows = own.end[i] - own.begin[i];
                if (rows) {
                    intern_verify_kernel<<<div_up(rows, 256), 256, 0, c.aux_streams[0]>>>(own, i, g.e_q.p, g.e_t.p, c.kept.qname.p, c.kept.tname.p,
                                                                                        g.node_name_off.p, g.node_names.p, tp, d_late);
                    c.launches.count++;
                    HERA_CUDA(cudaGetLastError());
                }
            }
            HERA_CUDA(cudaEventCreateWithFlags(&verify_done, cudaEventDisableTiming));
            HERA_CUDA(cudaEventRecord(verify_done, c.aux_streams[0]));
        }
        // anchors in node order
        uint32_t *anchor_scan = nullptr;
        HERA_CUDA(cudaMallocAsync(&anchor_scan, N * sizeof(uint32_t), c.stream));
        exclusive_scan_u8_to_u32(c, g.node_anchor.p, N, anchor_scan, d_late + 1);
        g.anchor_ids.reserve(N + 1);
        HERA_LAUNCH(c, anchor_list_kernel, div_up(N, 256), 256, 0, N, g.node_anchor.p, anchor_scan, g.anchor_ids.p);
        HERA_CUDA(cudaFreeAsync(anchor_scan, c.stream));
        HERA_CUDA(cudaFreeAsync(keys, c.stream)); HERA_CUDA(cudaFreeAsync(minpos, c.stream)); HERA_CUDA(cudaFreeAsync(slot_id, c.stream));
        HERA_CUDA(cudaFreeAsync(occ_slot, c.stream)); HERA_CUDA(cudaFreeAsync(first_scan, c.stream)); HERA_CUDA(cudaFreeAsync(first, c.stream));
        HERA_CUDA(cudaFreeAsync(d_small, c.stream));
        } else {
            // node table, e_q / e_t restored from a cache file (hera_graph_load_cache): only the anchor list is derived
            N = g.N;
            uint32_t *anchor_scan = nullptr;
            HERA_CUDA(cudaMallocAsync(&anchor_scan, N * sizeof(uint32_t), c.stream));
            HERA_CUDA(cudaMallocAsync(&d_late, 2 * sizeof(uint32_t), c.stream));
            HERA_CUDA(cudaMemsetAsync(d_late, 0, 2 * sizeof(uint32_t), c.stream));
            exclusive_scan_u8_to_u32(c, g.node_anchor.p, N, anchor_scan, d_late + 1);
            g.anchor_ids.reserve(N + 1);
            HERA_LAUNCH(c, anchor_list_kernel, div_up(N, 256), 256, 0, N, g.node_anchor.p, anchor_scan, g.anchor_ids.p);
            HERA_CUDA(cudaFreeAsync(anchor_scan, c.stream));
        }
        HERA_CUDA(cudaEventRecord(ev[1], c.stream));

        // ---- K4
        uint32_t *keys32 = nullptr, *keys_alt = nullptr, *vals_alt = nullptr;
        HERA_CUDA(cudaMallocAsync(&keys32, occ * sizeof(uint32_t), c.stream));
        HERA_CUDA(cudaMallocAsync(&keys_alt, occ * sizeof(uint32_t), c.stream));
        HERA_CUDA(cudaMallocAsync(&vals_alt, occ * sizeof(uint32_t), c.stream));
        g.adj_pay.reserve(occ); g.adj_off.reserve(N + 1);
        g.a_nbr.reserve(occ); g.a_es.reserve(occ); g.a_os.reserve(occ); g.a_delta.reserve(occ); g.canon.reserve(occ); g.adj4.reserve(occ); g.node_exact.reserve(N);
        HERA_LAUNCH(c, csr_entries_kernel, div_up(occ, 256), 256, 0, E, g.e_q.p, g.e_t.p, keys32, g.adj_pay.p);
        radix_sort_pairs(c, keys32, g.adj_pay.p, keys_alt, vals_alt, occ, bits_for(N));
        HERA_LAUNCH(c, csr_offsets_kernel, div_up(occ + 1, 256), 256, 0, occ, N, keys32, g.adj_off.p);
        HERA_LAUNCH(c, csr_gather_kernel, div_up(occ, 256), 256, 0, occ, g.adj_pay.p, g.e_q.p, g.e_t.p, c.kept.qs.p, c.kept.ts.p, c.kept.es.p, c.kept.os.p,
                    g.node_anchor.p, g.a_nbr.p, g.a_es.p, g.a_os.p, g.a_delta.p, g.adj4.p);
        uint32_t *d_max_deg = nullptr;
        HERA_CUDA(cudaMallocAsync(&d_max_deg, sizeof(uint32_t), c.stream));
        HERA_CUDA(cudaMemsetAsync(d_max_deg, 0, sizeof(uint32_t), c.stream));
        HERA_LAUNCH(c, node_exact_kernel, div_up(N * 32, 128), 128, 0, static_cast<uint32_t>(N), g.adj_off.p, g.a_es.p, g.a_os.p, g.a_nbr.p, g.node_exact.p, d_max_deg);
        c.h_counters.reserve(32 + xch_size(c));   // [14] late results, [15] max degree, [16..] slice boundaries of the ranks
        HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 15, d_max_deg, sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));   // read after the final sync below
        HERA_CUDA(cudaFreeAsync(d_max_deg, c.stream));
        HERA_CUDA(cudaFreeAsync(keys32, c.stream)); HERA_CUDA(cudaFreeAsync(keys_alt, c.stream)); HERA_CUDA(cudaFreeAsync(vals_alt, c.stream));
        HERA_CUDA(cudaEventRecord(ev[2], c.stream));

        // ---- rank permutations + canonical slots: each rank does its share of the nodes, the slices are all-gathered
        const int ranks = xch_size(c), me = xch_rank(c);
        cudaEvent_t cuts_ready = nullptr;
        if (ranks > 1) {   // slot ranges of the ranks' node ranges: read now, needed (on the host) once the rank kernels are queued
            uint64_t *d_cut = nullptr;
            HERA_CUDA(cudaMallocAsync(&d_cut, (ranks + 1) * sizeof(uint64_t), c.stream));
            HERA_LAUNCH(c, node_cut_kernel, 1, 64, 0, g.adj_off.p, N, ranks, d_cut);
            HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 16, d_cut, (ranks + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, c.stream));
            HERA_CUDA(cudaFreeAsync(d_cut, c.stream));
            HERA_CUDA(cudaEventCreateWithFlags(&cuts_ready, cudaEventDisableTiming));
            HERA_CUDA(cudaEventRecord(cuts_ready, c.stream));
        }
        const uint64_t n_lo = N * me / ranks, n_hi = N * (me + 1) / ranks;
        const unsigned my_nodes = static_cast<unsigned>(n_hi - n_lo);
        RankPtrs rp{};
        for (int m = 0; m < 4; m++) {
            if (!(metrics_mask & (1u << m))) continue;
            g.rank[m].reserve(occ);
            rp.p[m] = g.rank[m].p;
            g.rank_ready[m] = true;
        }
        if (my_nodes) {
            HERA_LAUNCH(c, rank_small_kernel, div_up(my_nodes, RB_WARPS), RB_THREADS, 0, g.adj_off.p, g.a_nbr.p, g.a_es.p, g.a_os.p, metrics_mask, rp, g.canon.p,
                        static_cast<uint32_t>(n_lo), my_nodes);
            HERA_LAUNCH(c, (rank_canon_kernel<2048, 512, true>), my_nodes, RK_THREADS, 0, g.adj_off.p, g.a_nbr.p, g.a_es.p, g.a_os.p, metrics_mask, rp, g.canon.p, static_cast<uint32_t>(n_lo));
        }
        if (ranks > 1) {
            HERA_CUDA(cudaEventSynchronize(cuts_ready));   // the slice boundaries left the device before the rank kernels started
            cudaEventDestroy(cuts_ready);
            const uint64_t *cut = c.h_counters.p + 16;
            std::vector<uint64_t> counts(ranks), displs(ranks);
            for (int r = 0; r < ranks; r++) { displs[r] = cut[r]; counts[r] = cut[r + 1] - cut[r]; }
            Exchange::Column cols[5] = {{g.canon.p, 4}};
            int ncols = 1;
            for (int m = 0; m < 4; m++)
                if (g.rank_ready[m]) cols[ncols++] = {g.rank[m].p, 4};
            c.xch->gather_rows(c, cols, ncols, 0, displs[me], counts.data(), displs.data());
        }
        HERA_CUDA(cudaEventRecord(ev[3], c.stream));
        if (verify_done) {
            HERA_CUDA(cudaStreamWaitEvent(c.stream, verify_done, 0));
            cudaEventDestroy(verify_done);
        }
        HERA_CUDA(cudaMemcpyAsync(c.h_counters.p + 14, d_late, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaFreeAsync(d_late, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        g.max_deg = static_cast<uint32_t>(c.h_counters.p[15] & 0xffffffffu);
        g.anchors = c.h_counters.p[14] >> 32;
        uint64_t collided = c.h_counters.p[14] & 0xffffffffu;
        if (xch_size(c) > 1) {
            std::vector<uint64_t> all(xch_size(c));
            c.xch->all_gather_host(c, &collided, 1, all.data());
            for (uint64_t v : all) collided |= v;
        }
        HERA_REQUIRE(!collided, 6, "64-bit name hash collision detected while interning node names");
        HERA_CUDA(cudaEventElapsedTime(&intern_ms, ev[0], ev[1]));
        HERA_CUDA(cudaEventElapsedTime(&csr_ms, ev[1], ev[2]));
        HERA_CUDA(cudaEventElapsedTime(&rank_ms, ev[2], ev[3]));
    }
    if (E == 0) for (int m = 0; m < 4; m++) g.rank_ready[m] = (metrics_mask >> m) & 1u;  // empty permutations
    g.ready = true;
    for (auto &e : ev) cudaEventDestroy(e);
    if (out) {
        out->nodes = g.N; out->anchors = g.anchors; out->edges = g.E; out->slots = g.S;
        out->intern_ms = intern_ms; out->csr_ms = csr_ms; out->rank_ms = rank_ms;
    }
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
        uint64_t *offs = nullptr; uint8_t *tmp8 = nullptr; uint32_t *tmp32 = nullptr;
        HERA_CUDA(cudaMallocAsync(&offs, (N + 1) * sizeof(uint64_t), c.stream));
        d2h(c, m.node_name_off, g.node_name_off.p, N + 1);
        m.node_names.reserve(g.name_bytes + 1);
        if (g.name_bytes) HERA_CUDA(cudaMemcpyAsync(m.node_names.p, g.node_names.p, g.name_bytes, cudaMemcpyDeviceToHost, c.stream));
        // edges
        d2h(c, m.e_q, g.e_q.p, E); d2h(c, m.e_t, g.e_t.p, E);
        d2h(c, m.e_qs, c.kept.qs.p, E); d2h(c, m.e_qe, c.kept.qe.p, E); d2h(c, m.e_ts, c.kept.ts.p, E); d2h(c, m.e_te, c.kept.te.p, E);
        d2h(c, m.e_os, c.kept.os.p, E); d2h(c, m.e_si, c.kept.si.p, E); d2h(c, m.e_es, c.kept.es.p, E);
        HERA_CUDA(cudaMallocAsync(&tmp8, S + 1, c.stream));
        HERA_CUDA(cudaMallocAsync(&tmp32, (S + 1) * sizeof(uint32_t), c.stream));
        HERA_LAUNCH(c, export_strand_kernel, div_up(E, 256), 256, 0, E, c.kept.flags.p, tmp8);
        d2h(c, m.e_strand, tmp8, E);
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        HERA_LAUNCH(c, export_adj_kernel, div_up(S, 256), 256, 0, S, g.adj_pay.p, tmp32, tmp8);
        d2h(c, m.adj_rec, tmp32, S); d2h(c, m.adj_side, tmp8, S);
        HERA_LAUNCH(c, widen_kernel, div_up(N + 1, 256), 256, 0, N + 1, g.adj_off.p, offs);
        d2h(c, m.adj_off, offs, N + 1);
        d2h(c, m.canon, g.canon.p, S);
        for (int k = 0; k < 4; k++) if (g.rank_ready[k]) d2h(c, m.rank[k], g.rank[k].p, S);
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        HERA_CUDA(cudaFreeAsync(offs, c.stream));
        HERA_CUDA(cudaFreeAsync(tmp8, c.stream)); HERA_CUDA(cudaFreeAsync(tmp32, c.stream));
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
constexpr char CACHE_MAGIC[8] = {'H', 'E', 'R', 'A', 'K', 'E', 'P', '1'};
struct CacheHeader {
    char magic[8];
    uint64_t E, N, name_bytes;
    hera_filter_params fp;
};

template <typename T>
void put_array(Ctx &c, FILE *f, const T *dev, uint64_t n, std::vector<char> &buf) {
    buf.resize(n * sizeof(T));
    if (n) {
        HERA_CUDA(cudaMemcpyAsync(buf.data(), dev, n * sizeof(T), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        HERA_REQUIRE(fwrite(buf.data(), sizeof(T), n, f) == n, 50, "cache: short write");
    }
}

template <typename T>
void get_array(Ctx &c, FILE *f, DBuf<T> &dev, uint64_t n, std::vector<char> &buf) {
    dev.reserve(n + 1);
    buf.resize(n * sizeof(T));
    if (n) {
        HERA_REQUIRE(fread(buf.data(), sizeof(T), n, f) == n, 51, "cache: file is truncated");
        HERA_CUDA(cudaMemcpyAsync(dev.p, buf.data(), n * sizeof(T), cudaMemcpyHostToDevice, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
    }
}
}  // namespace

void save_graph_cache(Ctx &c, const char *path) {
    GraphTables &g = c.g;
    HERA_REQUIRE(g.ready, 7, "hera_graph_save: call hera_graph_finalize first");
    FILE *f = fopen(path, "wb");
    if (!f) throw Error{1, std::string("Cannot write file: ") + path};
    struct Closer { FILE *f; ~Closer() { fclose(f); } } closer{f};
    CacheHeader h{};
    std::memcpy(h.magic, CACHE_MAGIC, 8);
    h.E = g.E; h.N = g.N; h.name_bytes = g.name_bytes; h.fp = c.fp;
    HERA_REQUIRE(fwrite(&h, sizeof(h), 1, f) == 1, 50, "cache: short write");
    std::vector<char> buf;
    KeptTable &k = c.kept;
    put_array(c, f, k.qs.p, g.E, buf); put_array(c, f, k.qe.p, g.E, buf); put_array(c, f, k.ts.p, g.E, buf); put_array(c, f, k.te.p, g.E, buf);
    put_array(c, f, k.qlen.p, g.E, buf); put_array(c, f, k.tlen.p, g.E, buf); put_array(c, f, k.rec.p, g.E, buf);
    put_array(c, f, k.os.p, g.E, buf); put_array(c, f, k.si.p, g.E, buf); put_array(c, f, k.es.p, g.E, buf); put_array(c, f, k.flags.p, g.E, buf);
    put_array(c, f, g.e_q.p, g.E, buf); put_array(c, f, g.e_t.p, g.E, buf);
    put_array(c, f, g.node_len.p, g.N, buf); put_array(c, f, g.node_anchor.p, g.N, buf);
    put_array(c, f, g.node_name_off.p, g.N + 1, buf); put_array(c, f, g.node_names.p, g.name_bytes, buf);
}

void load_graph_cache(Ctx &c, const char *path) {
    FILE *f = fopen(path, "rb");
    if (!f) throw Error{1, std::string("Cannot open file: ") + path};
    struct Closer { FILE *f; ~Closer() { fclose(f); } } closer{f};
    CacheHeader h{};
    HERA_REQUIRE(fread(&h, sizeof(h), 1, f) == 1 && std::memcmp(h.magic, CACHE_MAGIC, 8) == 0, 52, "not a hera_b200 graph cache");
    HERA_REQUIRE(c.kept.n == 0 && c.texts.empty(), 53, "hera_graph_load_cache needs a fresh graph (hera_graph_reset)");
    GraphTables &g = c.g;
    std::vector<char> buf;
    KeptTable &k = c.kept;
    const uint64_t E = h.E, N = h.N;
    get_array(c, f, k.qs, E, buf); get_array(c, f, k.qe, E, buf); get_array(c, f, k.ts, E, buf); get_array(c, f, k.te, E, buf);
    get_array(c, f, k.qlen, E, buf); get_array(c, f, k.tlen, E, buf); get_array(c, f, k.rec, E, buf);
    get_array(c, f, k.os, E, buf); get_array(c, f, k.si, E, buf); get_array(c, f, k.es, E, buf); get_array(c, f, k.flags, E, buf);
    get_array(c, f, g.e_q, E, buf); get_array(c, f, g.e_t, E, buf);
    get_array(c, f, g.node_len, N, buf); get_array(c, f, g.node_anchor, N, buf);
    get_array(c, f, g.node_name_off, N + 1, buf); get_array(c, f, g.node_names, h.name_bytes, buf);
    k.n = E;
    g.N = N; g.name_bytes = h.name_bytes;
    c.fp = h.fp;
    c.cache_loaded = true;
    g.ready = false;
}

}  // namespace hera
