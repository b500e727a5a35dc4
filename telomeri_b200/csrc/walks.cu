// K5 + K6: path search, one warp per walk.
//
// K5 replaces PathManager::buildMonteCarlo (reference src/PathManager.cpp:12-177): from every
// anchor, rebuild_attempts roulette walks.  The reference shares one std::mt19937 stream over
// all walks; here draw number k of walk (anchor, attempt) is Philox4x32-10 with counter
// (anchor node id, attempt, k, metric) and key = seed, turned into [0,1) the way libstdc++'s
// generate_canonical does, so a walk does not depend on any other walk.
// K6 replaces PathManager::buildDeterministic (src/PathManager.cpp:179-430): one walk per
// (anchor, first edge), greedy by the precomputed rank permutation.
//
// The 32 lanes of a warp stream the current node's CSR slice (neighbour ids with the anchor flag
// in bit 31, fp32 metric) and probe the walk's visited bitmap.  The roulette needs the fp64 sum
// of the candidate metrics in list order (src/PathManager.cpp:50-66,104-114).  A warp reduction
// associates differently, so it is used only when every partial sum is provably exact (all
// addends non-negative finite multiples of 2^k and the total below 2^(k+53)): then any order
// gives the sequential result bit for bit.  Otherwise the warp accumulates in list order.
#include <algorithm>
#include <chrono>
#include <string>
#include <type_traits>
#include <vector>

#include "ctx.cuh"
#include "primitives.cuh"

namespace hera {

constexpr int WK_THREADS = 128;
constexpr int WK_WARPS = WK_THREADS / 32;
constexpr int WK_CACHE_MAX = 1024;  // list entries (metric values) a warp can keep in shared memory between the two roulette passes

struct WalkGraph {
    const uint32_t *off, *canon, *rank, *node_len, *anchor_ids;
    const uint4 *adj4;        // {nbr | anchor bit 31, es, os, delta} per slot
    const uint2 *a8;          // {nbr | anchor bit 31, this call's stored score} per slot: what the roulette scan reads
    const uint8_t *node_exact;
    uint32_t N, anchors;
    int sqrt_metric;
    int use_os;               // which field of adj4 is the metric
    uint32_t exact_bit;       // bit of node_exact for this metric
};

struct WalkScratch {
    uint32_t *bitmap;     // warps * words
    uint32_t *nodes, *slots, *vlog;   // warps * cap
    uint32_t words, cap;
    uint32_t cache;       // entries of the per-warp value cache (multiple of 32, covers the longest list up to WK_CACHE_MAX)
};

struct WalkOut {
    // per walk
    uint32_t *res_n;       // nodes of the accepted path, 0 = none
    uint64_t *res_off;     // offset into the pool
    int64_t *res_len;
    uint32_t *res_draws;
    // pool
    uint32_t *pool_nodes, *pool_slots;
    unsigned long long pool_cap;
    unsigned long long *pool_used;    // also keeps counting past the capacity
    unsigned long long *steps;        // adjacency entries examined
    unsigned int *next_walk;
    unsigned int *overflow;           // set when a walk outgrew the per-warp scratch (cap): the job is run again with room for N nodes
    // walks are dealt out by id: this rank runs walk ids part_rank, part_rank + part_size, ... ; results are indexed by the local ordinal
    uint32_t part_rank, part_size;
};

__device__ __forceinline__ void philox_round(uint32_t &c0, uint32_t &c1, uint32_t &c2, uint32_t &c3, uint32_t k0, uint32_t k1) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
}

__device__ __forceinline__ double mc_draw(uint64_t seed, uint32_t metric, uint32_t anchor, uint32_t attempt, uint32_t draw) {
    uint32_t c0 = anchor, c1 = attempt, c2 = draw, c3 = metric;
    uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; r++) {
        philox_round(c0, c1, c2, c3, k0, k1);
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    // generate_canonical<double,53> over a 32-bit engine: (x0 + x1 * 2^32) / 2^64, one rounding in the sum
    const double sum = __dadd_rn(__uint2double_rn(c0), __dmul_rn(__uint2double_rn(c1), 4294967296.0));
    double u = __dmul_rn(sum, 5.42101086242752217e-20);  // exact: 2^-64
    if (u >= 1.0) u = 0.99999999999999988897769753748434595763683319091796875;
    return u;
}

__device__ __forceinline__ float walk_metric(const WalkGraph &g, uint32_t s) {
    const uint4 e = __ldg(g.adj4 + s);
    const float v = __uint_as_float(g.use_os ? e.z : e.y);
    return g.sqrt_metric ? __fsqrt_rn(v) : v;
}
__device__ __forceinline__ uint32_t slot_nbr(const WalkGraph &g, uint32_t s) { return __ldg(&reinterpret_cast<const uint32_t *>(g.adj4 + s)[0]); }
__device__ __forceinline__ int32_t slot_delta(const WalkGraph &g, uint32_t s) { return static_cast<int32_t>(__ldg(&reinterpret_cast<const uint32_t *>(g.adj4 + s)[3])); }

__device__ __forceinline__ bool bit_get(const uint32_t *bm, uint32_t i) { return (bm[i >> 5] >> (i & 31)) & 1u; }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = __dadd_rn(v, __shfl_xor_sync(0xffffffffu, v, d));
    return v;
}

// canon != nullptr: slots[1..] still hold the slots the greedy walk chose; the path stores their canonical slots (first
// parallel edge, src/PathManager.cpp:340-347) -- looked up here, once per accepted path, instead of inside every step.
__device__ __forceinline__ void emit_path(const WalkOut &o, uint32_t local, const uint32_t *nodes, const uint32_t *slots, uint32_t n_nodes, int64_t len,
                                          const uint32_t *__restrict__ canon = nullptr) {
    const uint32_t lane = lane_id();
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(o.pool_used, static_cast<unsigned long long>(n_nodes));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (lane == 0) { o.res_n[local] = n_nodes; o.res_off[local] = base; o.res_len[local] = len; }
    if (base + n_nodes <= o.pool_cap) {
        for (uint32_t i = lane; i < n_nodes; i += 32) o.pool_nodes[base + i] = nodes[i];
        for (uint32_t i = lane; i + 1 < n_nodes; i += 32) o.pool_slots[base + i] = (canon && i) ? canon[slots[i]] : slots[i];
    }
}

__device__ __forceinline__ void clear_visited(uint32_t *bm, uint32_t words, const uint32_t *vlog, uint32_t nlog) {
    const uint32_t lane = lane_id();
    if (nlog * 4u < words) {
        for (uint32_t i = lane; i < nlog; i += 32) bm[vlog[i] >> 5] = 0;
    } else {
        for (uint32_t i = lane; i < words; i += 32) bm[i] = 0;
    }
    __syncwarp();
}

// Visited-set probe: the walk's bitmap lives in shared memory when the graph is small enough
// (SMEM_BM), else in the warp's slab of global scratch.  (A Bloom filter in shared memory in front of the global bitmap
// was measured on configs[3] and gave nothing: the probes of a step are independent loads that hit L2.)
template <bool SMEM_BM>
__device__ __forceinline__ bool bm_get(const uint32_t *s_bm, const uint32_t *g_bm, uint32_t i) {
    if (SMEM_BM) return (s_bm[i >> 5] >> (i & 31)) & 1u;
    return (g_bm[i >> 5] >> (i & 31)) & 1u;
}

template <bool SMEM_BM, int WK_MLP, int MIN_CTAS>
__global__ void __launch_bounds__(WK_THREADS, MIN_CTAS) walk_mc_kernel(WalkGraph g, WalkScratch sc, WalkOut o, uint32_t total_walks, uint32_t attempts,
                                                             int backtrack_attempts, long long len_threshold, uint64_t seed, uint32_t metric_id) {
    // dynamic shared memory: per-warp value cache of the current node's list, then the per-warp visited bitmaps
    extern __shared__ __align__(16) uint32_t wk_smem[];
    const uint32_t lane = lane_id(), wib = warp_id();
    const uint32_t WK_CACHE = sc.cache;
    float *s_cval = reinterpret_cast<float *>(wk_smem) + wib * WK_CACHE;
    uint32_t *s_bm = wk_smem + WK_WARPS * WK_CACHE + (SMEM_BM ? wib * sc.words : 0u);
    const uint32_t wslot = blockIdx.x * WK_WARPS + wib;
    uint32_t *g_bm = sc.bitmap + (SMEM_BM ? 0ull : static_cast<uint64_t>(wslot) * sc.words);
    uint32_t *bm = SMEM_BM ? s_bm : g_bm;
    // the probe address is bmw + 4 (nb >> 5): keep the warp's bitmap base in registers (the compiler otherwise rebuilds
    // slab index x words + base from the kernel parameters for every probe: 7 instructions)
    const uint32_t *bmw = bm;
    if (!SMEM_BM) asm volatile("" : "+l"(bmw));
    uint32_t *nodes = sc.nodes + static_cast<uint64_t>(wslot) * sc.cap;
    uint32_t *slots = sc.slots + static_cast<uint64_t>(wslot) * sc.cap;
    uint32_t *vlog = sc.vlog + static_cast<uint64_t>(wslot) * sc.cap;
    unsigned long long examined = 0;
    if (SMEM_BM) {
        for (uint32_t i = lane; i < sc.words; i += 32) s_bm[i] = 0;
        __syncwarp();
    }

    for (;;) {
        uint32_t local = 0;
        if (lane == 0) local = atomicAdd(o.next_walk, 1u);
        local = __shfl_sync(0xffffffffu, local, 0);
        if (local >= total_walks) break;   // total_walks = walks of this rank
        const uint32_t walk = local * o.part_size + o.part_rank;
        const uint32_t start = g.anchor_ids[walk / attempts], attempt = walk % attempts;

        uint32_t n_nodes = 1, nlog = 1, draws = 0;
        double u_bank = 0.0;
        int backtracks = 0;
        long long sum_all = g.node_len[start];   // first node's length + every edge delta on the path
        uint32_t cur = start;
        bool accepted = false;
        if (lane == 0) { nodes[0] = start; vlog[0] = start; bm[start >> 5] |= 1u << (start & 31); }
        __syncwarp();

        for (;;) {
            const uint32_t base = g.off[cur], deg = g.off[cur + 1] - base;
            const uint32_t node_flags = g.node_exact[cur];
            const bool order_free = (node_flags & g.exact_bit) != 0;
            const bool may_force = (node_flags & 16u) != 0;   // the node has an anchor among its neighbours
            // ---- pass 1: forced anchor edge, candidate set (src/PathManager.cpp:53-67)
            int forced = -1;
            bool any_cand = false;
            uint32_t forced_nbw = 0;
            // WK_MLP 32-entry chunks in flight per round trip.  The loop body is branch-free inside a chunk group (ncu on
            // configs[3]: the kernel is issue-bound, 94 warp instructions per 32 entries with the forced-edge test as a branch
            // per chunk and the bitmap address recomputed per probe): per-lane pointers are hoisted, the probe of an entry
            // behind the list's end reads nothing and sees "visited", the forced-edge ballots are looked at once per group.
            const uint2 *lp = g.a8 + base + lane;
            float *cp = s_cval + lane;
            // (two instances of the loop: only nodes next to an anchor -- few -- need the forced-edge ballots)
            auto scan_list = [&](auto with_force) {
                constexpr bool FORCE = decltype(with_force)::value;
                for (uint32_t c0 = 0; c0 < deg; c0 += 32 * WK_MLP) {
                    uint2 a[WK_MLP];
#pragma unroll
                    for (int j = 0; j < WK_MLP; j++) a[j] = c0 + 32 * j + lane < deg ? __ldg(lp + c0 + 32 * j) : make_uint2(0, 0);
                    uint32_t fb[WK_MLP];
#pragma unroll
                    for (int j = 0; j < WK_MLP; j++) {
                        const uint32_t k = c0 + 32 * j + lane;
                        const bool valid = k < deg;
                        const uint32_t nbw = a[j].x, nb = nbw & 0x7fffffffu;
                        const uint32_t word = valid ? bmw[nb >> 5] : 0xffffffffu;
                        const bool cand = ((word >> (nb & 31u)) & 1u) == 0u;
                        fb[j] = FORCE ? __ballot_sync(0xffffffffu, valid && (nbw >> 31) && nb != start) : 0u;
                        any_cand |= cand;
                        // entry k keeps its place in the cache: its metric if it is a candidate, else -0.0f
                        // (adding -0.0 to a sum changes nothing, and the sign bit tells the two apart)
                        float v = __uint_as_float(a[j].y);
                        if (g.sqrt_metric) v = __fsqrt_rn(v);
                        if (k < WK_CACHE) cp[c0 + 32 * j] = cand ? v : -0.0f;
                    }
                    if (FORCE) {   // first forced edge of the group, if any: entries behind it do not matter
#pragma unroll
                        for (int j = 0; j < WK_MLP; j++) {
                            if (forced < 0 && fb[j]) {
                                const int src = __ffs(fb[j]) - 1;
                                forced = static_cast<int>(c0) + 32 * j + src;
                                forced_nbw = __shfl_sync(0xffffffffu, a[j].x, src);
                            }
                        }
                        if (forced >= 0) break;
                    }
                }
            };
            if (may_force) scan_list(std::true_type{}); else scan_list(std::false_type{});
            const uint32_t ncand = __any_sync(0xffffffffu, any_cand) ? 1u : 0u;   // only "none / some" matters below
            examined += deg;
            int pick = -1;
            if (forced >= 0) {
                pick = forced;
            } else if (ncand == 0) {
                // ---- dead end: backtrack (src/PathManager.cpp:72-101)
                if (backtracks < backtrack_attempts && n_nodes > 2) {
                    bool open = false;
                    do {
                        n_nodes--;
                        sum_all -= slot_delta(g, slots[n_nodes - 1]);
                        const uint32_t back = nodes[n_nodes - 1];
                        const uint32_t bb = g.off[back], bd = g.off[back + 1] - bb;
                        for (uint32_t c0 = 0; c0 < bd && !open; c0 += 32) {
                            const uint32_t k = c0 + lane;
                            const bool un = k < bd && !bm_get<SMEM_BM>(s_bm, g_bm, slot_nbr(g, bb + k) & 0x7fffffffu);
                            open = __any_sync(0xffffffffu, un);
                        }
                        examined += bd;
                    } while (n_nodes > 1 && !open);
                    if (n_nodes == 1) break;
                    backtracks++;
                    cur = nodes[n_nodes - 1];
                    continue;
                }
                break;
            } else {
                // ---- roulette (src/PathManager.cpp:102-115)
                __syncwarp();
                // draws come 32 at a time: lane l computes draw (draws & ~31) + l, one Philox call per lane per 32 steps
                if ((draws & 31u) == 0) u_bank = mc_draw(seed, metric_id, start, attempt, draws + lane);
                const double u = __shfl_sync(0xffffffffu, u_bank, draws & 31u);
                draws++;
                if (order_free && deg <= WK_CACHE) {
                    // Every partial sum is exact in any order (node_exact): lane l sums entries [l*K, (l+1)*K),
                    // lane totals are scanned, and the lane whose range crosses u*sum walks it once more.
                    const uint32_t K = ((deg + 31) >> 5) | 1u;   // odd stride: no bank conflicts
                    const uint32_t lo = min(lane * K, deg), hi = min(lo + K, deg);
                    double mine = 0.0;
#pragma unroll 4
                    for (uint32_t i = lo; i < hi; i++) mine = __dadd_rn(mine, static_cast<double>(s_cval[i]));
                    double incl = mine;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const double up = __shfl_up_sync(0xffffffffu, incl, d);
                        if (lane >= static_cast<uint32_t>(d)) incl = __dadd_rn(incl, up);
                    }
                    const double total = __shfl_sync(0xffffffffu, incl, 31);
                    const double target = __dmul_rn(u, total);
                    // first lane whose range holds a candidate at which the running sum reaches the target
                    int found = -1;
                    if (hi > lo && incl >= target) {
                        double run = __dsub_rn(incl, mine);   // exact: both are multiples of the node's quantum
                        for (uint32_t i = lo; i < hi; i++) {
                            const float v = s_cval[i];
                            if (__float_as_uint(v) >> 31) continue;   // not a candidate
                            run = __dadd_rn(run, static_cast<double>(v));
                            if (run >= target) { found = static_cast<int>(i); break; }
                        }
                    }
                    const uint32_t hit = __ballot_sync(0xffffffffu, found >= 0);
                    if (hit) pick = __shfl_sync(0xffffffffu, found, __ffs(hit) - 1);
                } else {
                    // list-order accumulation: every lane replays the same sequence of additions
                    double total = 0.0;
                    for (uint32_t c0 = 0; c0 < deg; c0 += 32) {
                        const uint32_t k = c0 + lane;
                        const bool valid = k < deg;
                        const bool cand = valid && !bm_get<SMEM_BM>(s_bm, g_bm, slot_nbr(g, base + k) & 0x7fffffffu);
                        const float v = cand ? walk_metric(g, base + k) : 0.0f;
                        const uint32_t cb = __ballot_sync(0xffffffffu, cand);
                        for (uint32_t m = cb; m; m &= m - 1) total = __dadd_rn(total, static_cast<double>(__shfl_sync(0xffffffffu, v, __ffs(m) - 1)));
                    }
                    const double target = __dmul_rn(u, total);
                    double seq = 0.0;
                    for (uint32_t c0 = 0; c0 < deg && pick < 0; c0 += 32) {
                        const uint32_t k = c0 + lane;
                        const bool valid = k < deg;
                        const bool cand = valid && !bm_get<SMEM_BM>(s_bm, g_bm, slot_nbr(g, base + k) & 0x7fffffffu);
                        const float v = cand ? walk_metric(g, base + k) : 0.0f;
                        const uint32_t cb = __ballot_sync(0xffffffffu, cand);
                        for (uint32_t m = cb; m; m &= m - 1) {
                            const int src = __ffs(m) - 1;
                            seq = __dadd_rn(seq, static_cast<double>(__shfl_sync(0xffffffffu, v, src)));
                            if (seq >= target) { pick = static_cast<int>(c0) + src; break; }
                        }
                    }
                    examined += 2ull * deg;
                }
                if (pick < 0) break;  // only with NaN metrics; the reference dereferences null here
            }
            // ---- take the edge (src/PathManager.cpp:117-142)
            const uint32_t slot = base + static_cast<uint32_t>(pick);
            // the taken entry's coordinate delta (and, for a roulette pick, its neighbour) from the full 16-byte entry
            const uint4 taken = __ldg(g.adj4 + slot);
            const uint32_t nbw = forced >= 0 ? forced_nbw : taken.x;
            const int32_t dl = static_cast<int32_t>(taken.w);
            cur = nbw & 0x7fffffffu;
            if (nlog >= sc.cap) { if (lane == 0) *o.overflow = 1u; break; }   // scratch sized for typical walks: the host re-runs the job with cap = N + 1
            if (lane == 0) {
                nodes[n_nodes] = cur; slots[n_nodes - 1] = slot; vlog[nlog] = cur;
                bm[cur >> 5] |= 1u << (cur & 31);
            }
            __syncwarp();
            n_nodes++; nlog++;
            sum_all += dl;
            const long long len = sum_all - dl;   // last edge excluded (src/Path.cpp:12-15)
            if (len <= 0) break;
            if (nbw >> 31) { accepted = true; break; }
            if (len >= len_threshold) break;
        }
        if (lane == 0) o.res_draws[local] = draws;
        if (accepted) {
            const long long len = sum_all - slot_delta(g, slots[n_nodes - 2]);
            if (len >= 0) emit_path(o, local, nodes, slots, n_nodes, len);
        }
        clear_visited(bm, sc.words, vlog, nlog);
    }
    if (lane == 0 && examined) atomicAdd(o.steps, examined);
}

// Greedy walk ids: prefix over anchors of the number of first edges tried
// (src/PathManager.cpp:204,224-229: the first edge that lands on an anchor ends the loop).
__global__ void det_count_kernel(WalkGraph g, uint32_t *first_edges) {   // one warp per anchor
    const uint32_t a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (a >= g.anchors) return;
    const uint32_t x = g.anchor_ids[a], base = g.off[x], deg = g.off[x + 1] - base;
    uint32_t n = deg;
    for (uint32_t c0 = 0; c0 < deg; c0 += 32) {
        const uint32_t k = c0 + lane;
        const uint32_t hit = __ballot_sync(0xffffffffu, k < deg && (__ldg(&g.a8[base + k].x) >> 31));
        if (hit) { n = c0 + __ffs(hit); break; }
    }
    if (lane == 0) first_edges[a] = n;
}

__global__ void __launch_bounds__(WK_THREADS) walk_det_kernel(WalkGraph g, WalkScratch sc, WalkOut o, uint32_t total_walks, const uint32_t *__restrict__ walk_base) {
    const uint32_t lane = lane_id(), wib = warp_id();
    const uint32_t wslot = blockIdx.x * WK_WARPS + wib;
    uint32_t *bm = sc.bitmap + static_cast<uint64_t>(wslot) * sc.words;
    uint32_t *nodes = sc.nodes + static_cast<uint64_t>(wslot) * sc.cap;
    uint32_t *slots = sc.slots + static_cast<uint64_t>(wslot) * sc.cap;
    uint32_t *vlog = sc.vlog + static_cast<uint64_t>(wslot) * sc.cap;
    unsigned long long examined = 0;
    for (;;) {
        uint32_t local = 0;
        if (lane == 0) local = atomicAdd(o.next_walk, 1u);
        local = __shfl_sync(0xffffffffu, local, 0);
        if (local >= total_walks) break;   // total_walks = walks of this rank
        const uint32_t walk = local * o.part_size + o.part_rank;
        // anchor = last a with walk_base[a] <= walk
        uint32_t lo = 0, hi = g.anchors;
        while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (walk_base[mid] <= walk) lo = mid; else hi = mid; }
        const uint32_t start = g.anchor_ids[lo];
        const uint32_t fs = g.off[start] + (walk - walk_base[lo]);
        const uint32_t nbw0 = slot_nbr(g, fs);
        uint32_t cur = nbw0 & 0x7fffffffu;
        uint32_t n_nodes = 2, nlog = 2;
        if (lane == 0) {
            nodes[0] = start; nodes[1] = cur; slots[0] = fs; vlog[0] = start; vlog[1] = cur;
            bm[start >> 5] |= 1u << (start & 31);
            bm[cur >> 5] |= 1u << (cur & 31);
        }
        __syncwarp();
        bool reached = false;
        if (nbw0 >> 31) {
            emit_path(o, local, nodes, slots, 2, static_cast<int64_t>(g.node_len[start]));
        } else {
            int step = 0;
            uint32_t skip = 0;
            bool went_back = false;
            for (;;) {
                const uint32_t base = g.off[cur], deg = g.off[cur + 1] - base;
                int chosen = -1;
                if (deg != 0 && skip < deg - 1) {
                    for (uint32_t c0 = skip; c0 < deg - 1 && chosen < 0; c0 += 32) {
                        const uint32_t k = c0 + lane;
                        uint32_t s = 0;
                        bool un = false;
                        if (k < deg - 1) { s = g.rank[base + k]; un = !bit_get(bm, slot_nbr(g, s) & 0x7fffffffu); }
                        const uint32_t ub = __ballot_sync(0xffffffffu, un);
                        if (ub) chosen = static_cast<int>(__shfl_sync(0xffffffffu, s, __ffs(ub) - 1));
                        examined += 32;
                    }
                }
                if (chosen < 0) {
                    if (step == 0 || went_back) break;
                    went_back = true;
                    step--;
                    skip++;
                    n_nodes--;
                    cur = nodes[n_nodes - 1];
                    if (lane == 0) bm[cur >> 5] &= ~(1u << (cur & 31));
                    __syncwarp();
                    continue;
                }
                const uint32_t nbw = slot_nbr(g, chosen);
                const uint32_t next = nbw & 0x7fffffffu;
                if (nlog >= sc.cap) { if (lane == 0) *o.overflow = 1u; break; }
                if (lane == 0) {
                    slots[n_nodes - 1] = static_cast<uint32_t>(chosen);   // canonical slot: when the path is accepted (emit_path)
                    bm[cur >> 5] |= 1u << (cur & 31);
                    nodes[n_nodes] = next;
                    vlog[nlog] = next;
                }
                __syncwarp();
                // the node just left is in the log already (it was logged when it was entered)
                cur = next;
                n_nodes++; nlog++;
                if (nbw >> 31) { reached = true; break; }
                step++;
            }
            if (reached) {
                // src/Path.cpp:6-16 over the final edge list
                long long part = 0;
                for (uint32_t i = lane; i + 2 < n_nodes; i += 32) part += slot_delta(g, i ? g.canon[slots[i]] : slots[i]);
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
                const long long len = static_cast<long long>(g.node_len[start]) + part;
                if (len >= 0) emit_path(o, local, nodes, slots, n_nodes, len, g.canon);
            }
        }
        if (lane == 0) o.res_draws[local] = 0;
        clear_visited(bm, sc.words, vlog, nlog);
    }
    if (lane == 0 && examined) atomicAdd(o.steps, examined);
}

// Copies accepted paths from the pool to the path store in walk order.
__global__ void __launch_bounds__(WK_THREADS) gather_paths_kernel(uint32_t total_walks, const uint32_t *__restrict__ res_n, const uint64_t *__restrict__ res_off,
                                                                  const int64_t *__restrict__ res_len, const uint32_t *__restrict__ accepted_scan,
                                                                  const uint64_t *__restrict__ node_scan, const uint32_t *__restrict__ pool_nodes,
                                                                  const uint32_t *__restrict__ pool_slots, uint64_t path_base, uint64_t node_base,
                                                                  uint64_t *__restrict__ off, uint32_t *__restrict__ nodes, uint32_t *__restrict__ slots,
                                                                  int64_t *__restrict__ len) {
    const uint32_t walk = (blockIdx.x * WK_THREADS + threadIdx.x) >> 5, lane = lane_id();
    if (walk >= total_walks) return;
    const uint32_t n = res_n[walk];
    if (!n) return;
    const uint64_t p = path_base + accepted_scan[walk];
    const uint64_t dst = node_base + node_scan[walk];
    const uint64_t src = res_off[walk];
    if (lane == 0) { off[p + 1] = dst + n; len[p] = res_len[walk]; }
    for (uint32_t i = lane; i < n; i += 32) nodes[dst + i] = pool_nodes[src + i];
    for (uint32_t i = lane; i + 1 < n; i += 32) slots[dst - p + i] = pool_slots[src + i];
}

__global__ void flag_nonzero_kernel(uint32_t n, const uint32_t *__restrict__ in, uint8_t *__restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[i] ? 1 : 0;
}

namespace {

struct WalkRun {
    uint32_t *res_n = nullptr, *res_draws = nullptr, *pool_nodes = nullptr, *pool_slots = nullptr;
    uint64_t *res_off = nullptr;
    int64_t *res_len = nullptr;
    unsigned long long *d_ctr = nullptr;  // [0] pool_used [1] steps [2] next_walk (low 32 bits)
    uint32_t *bitmap = nullptr, *nodes = nullptr, *slots = nullptr, *vlog = nullptr;
};

WalkGraph make_walk_graph(Ctx &c, int metric) {
    GraphTables &g = c.g;
    WalkGraph w{};
    w.off = g.adj_off.p; w.canon = g.canon.p; w.rank = g.rank_ready[metric] ? g.rank[metric].p : nullptr;
    w.node_len = g.node_len.p; w.anchor_ids = g.anchor_ids.p;
    w.adj4 = g.adj4.p; w.node_exact = g.node_exact.p;
    w.a8 = g.a8[(metric == HERA_OVERLAP_SCORE || metric == HERA_OVERLAP_SCORE_SQRT) ? 1 : 0].p;
    w.use_os = (metric == HERA_OVERLAP_SCORE || metric == HERA_OVERLAP_SCORE_SQRT) ? 1 : 0;
    w.exact_bit = 1u << metric;
    w.N = static_cast<uint32_t>(g.N); w.anchors = static_cast<uint32_t>(g.anchors);
    w.sqrt_metric = (metric == HERA_EXTENSION_SCORE_SQRT || metric == HERA_OVERLAP_SCORE_SQRT) ? 1 : 0;
    return w;
}

constexpr int WK_MAX_RANKS = 16;
struct RankTable {
    uint64_t walk_displ[WK_MAX_RANKS];   // where rank r's per-walk results start in the gathered arrays
    uint64_t pool_displ[WK_MAX_RANKS];   // where rank r's path pool starts in the gathered pool
};

// Gathered per-walk results are rank-major (rank r ran walk ids r, r + size, ...): put them back in walk-id order.
__global__ void walk_order_kernel(uint32_t total_walks, uint32_t size, RankTable t, const uint32_t *__restrict__ g_n, const uint64_t *__restrict__ g_off,
                                  const int64_t *__restrict__ g_len, const uint32_t *__restrict__ g_draws, uint32_t *__restrict__ res_n,
                                  uint64_t *__restrict__ res_off, int64_t *__restrict__ res_len, uint32_t *__restrict__ res_draws) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= total_walks) return;
    const uint32_t r = w % size;
    const uint64_t src = t.walk_displ[r] + w / size;
    res_n[w] = g_n[src];
    res_off[w] = g_off[src] + t.pool_displ[r];
    res_len[w] = g_len[src];
    res_draws[w] = g_draws[src];
}

// One walk call (Monte Carlo or greedy, one metric): scratch, the kernel launch on its own stream, and
// -- once the kernel is done -- the gather of the accepted paths of ALL ranks into c.all in walk-id
// order.  Several jobs may be in flight at once (hera_paths_build): their kernels share the GPU, which
// hides the long tail every walk kernel has (its critical path is its longest walk).
struct WalkJob {
    int kind = 0, metric = 0;
    uint32_t total_walks = 0, my_walks = 0;
    hera_path_params pp{};
    uint64_t seed = 0;
    WalkGraph wg{};
    WalkRun r;
    uint32_t *first_edges = nullptr, *walk_base = nullptr;   // greedy walks: first edges per anchor, prefix
    uint64_t pool_cap = 0, warps = 0;
    uint32_t cap = 0, words = 0;
    int warps_per_sm = 1;
    unsigned blocks = 0;
    cudaStream_t stream = nullptr;   // the stream the job's kernel runs on
    cudaStream_t home = nullptr;     // the context's stream: the job's memory is allocated and freed in its order
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    unsigned long long *h_ctr = nullptr;   // pinned: [0] pool used [1] entries examined [2] ticket [3] scratch overflow
    uint64_t launches = 0;

    WalkJob() = default;
    WalkJob(const WalkJob &) = delete;
    WalkJob &operator=(const WalkJob &) = delete;
    // A job that goes while it still holds memory (an exception between job_prepare and the end of job_finish) waits for
    // its kernel and gives everything back.
    ~WalkJob() { release(true); }
    void release(bool wait_for_kernel) {
        void **held[] = {reinterpret_cast<void **>(&r.res_n), reinterpret_cast<void **>(&r.res_draws), reinterpret_cast<void **>(&r.res_off),
                         reinterpret_cast<void **>(&r.res_len), reinterpret_cast<void **>(&r.d_ctr), reinterpret_cast<void **>(&r.bitmap),
                         reinterpret_cast<void **>(&r.nodes), reinterpret_cast<void **>(&r.slots), reinterpret_cast<void **>(&r.vlog),
                         reinterpret_cast<void **>(&r.pool_nodes), reinterpret_cast<void **>(&r.pool_slots), reinterpret_cast<void **>(&first_edges),
                         reinterpret_cast<void **>(&walk_base)};
        bool any = false;
        for (void **q : held) any = any || *q != nullptr;
        if (any && wait_for_kernel && stream) cudaStreamSynchronize(stream);
        for (void **q : held) {
            if (*q) cudaFreeAsync(*q, home);
            *q = nullptr;
        }
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        ev0 = ev1 = nullptr;
    }
};

uint32_t walk_share(uint32_t total_walks, int ranks, int r) { return static_cast<uint32_t>((static_cast<uint64_t>(total_walks) + ranks - 1 - r) / ranks); }

void job_launch_kernel(Ctx &c, WalkJob &j) {
    const uint32_t cache = std::min<uint32_t>(WK_CACHE_MAX, std::max<uint32_t>(64, (c.g.max_deg + 31) / 32 * 32));
    WalkScratch sc{j.r.bitmap, j.r.nodes, j.r.slots, j.r.vlog, j.words, j.cap, cache};
    WalkOut wo{j.r.res_n, j.r.res_off, j.r.res_len, j.r.res_draws, j.r.pool_nodes, j.r.pool_slots, j.pool_cap, j.r.d_ctr, j.r.d_ctr + 1,
               reinterpret_cast<unsigned int *>(j.r.d_ctr + 2), reinterpret_cast<unsigned int *>(j.r.d_ctr + 3),
               static_cast<uint32_t>(xch_rank(c)), static_cast<uint32_t>(xch_size(c))};
    HERA_CUDA(cudaMemsetAsync(j.r.d_ctr, 0, 4 * sizeof(unsigned long long), j.stream));
    HERA_CUDA(cudaMemsetAsync(j.r.res_n, 0, std::max<uint32_t>(j.my_walks, 1) * sizeof(uint32_t), j.stream));
    HERA_CUDA(cudaMemsetAsync(j.r.res_draws, 0, std::max<uint32_t>(j.my_walks, 1) * sizeof(uint32_t), j.stream));
    HERA_CUDA(cudaEventRecord(j.ev0, j.stream));
    if (j.my_walks) {
        if (j.kind == 0) {
            // visited bitmaps in shared memory while four of them fit beside the value cache (N <= 65536 nodes)
            const bool smem_bm = c.opt_walk_smem_bitmap && j.words <= 2048;
            const size_t smem = (1ull * WK_WARPS * cache + (smem_bm ? static_cast<size_t>(WK_WARPS) * j.words : 0)) * sizeof(uint32_t);
            auto launch = [&](auto kernel) {
                kernel<<<j.blocks, WK_THREADS, smem, j.stream>>>(j.wg, sc, wo, j.my_walks, static_cast<uint32_t>(j.pp.rebuild_attempts), j.pp.backtrack_attempts,
                                                                static_cast<long long>(j.pp.len_threshold), j.seed, static_cast<uint32_t>(j.metric));
            };
            // variants: chunks in flight per round trip x resident CTAs (registers): 4 x 8 (<= 64 regs), 2 x 12 (<= 40), 1 x 16 (<= 32)
            if (smem_bm) {
                if (c.opt_walk_variant == 1) launch(walk_mc_kernel<true, 2, 12>);
                else if (c.opt_walk_variant == 2) launch(walk_mc_kernel<true, 1, 16>);
                else if (c.opt_walk_variant == 3) launch(walk_mc_kernel<true, 5, 8>);
                else if (c.opt_walk_variant == 4) launch(walk_mc_kernel<true, 6, 8>);
                else if (c.opt_walk_variant == 5) launch(walk_mc_kernel<true, 3, 8>);
                else launch(walk_mc_kernel<true, 4, 8>);
            } else {
                if (c.opt_walk_variant == 1) launch(walk_mc_kernel<false, 2, 12>);
                else if (c.opt_walk_variant == 2) launch(walk_mc_kernel<false, 1, 16>);
                else launch(walk_mc_kernel<false, 4, 8>);
            }
        } else {
            walk_det_kernel<<<j.blocks, WK_THREADS, 0, j.stream>>>(j.wg, sc, wo, j.my_walks, j.walk_base);
        }
        c.launches.count++;
        j.launches++;
        HERA_CUDA(cudaGetLastError());
    }
    HERA_CUDA(cudaEventRecord(j.ev1, j.stream));
    HERA_CUDA(cudaMemcpyAsync(j.h_ctr, j.r.d_ctr, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, j.stream));
}


// Per-warp scratch of a job: visited bitmap + node / slot / visited logs for `cap` nodes.
void job_alloc_scratch(Ctx &c, WalkJob &j, bool full, cudaStream_t stream) {
    GraphTables &g = c.g;
    int sms = 148;
    HERA_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c.device));
    j.cap = static_cast<uint32_t>(full ? g.N + 1 : std::min<uint64_t>(g.N + 1, static_cast<uint64_t>(std::max(2, c.opt_walk_cap_small))));
    const uint64_t per_warp = (3ull * j.cap + j.words) * sizeof(uint32_t);
    uint64_t warps = static_cast<uint64_t>(sms) * std::max(1, j.warps_per_sm);
    while (warps > WK_WARPS && warps * per_warp > (8ull << 30)) warps /= 2;
    if (warps > j.my_walks) warps = (static_cast<uint64_t>(j.my_walks) + WK_WARPS - 1) / WK_WARPS * WK_WARPS;
    if (warps == 0) warps = WK_WARPS;
    j.blocks = static_cast<unsigned>((warps + WK_WARPS - 1) / WK_WARPS);
    j.warps = static_cast<uint64_t>(j.blocks) * WK_WARPS;
    WalkRun &r = j.r;
    HERA_CUDA(cudaMallocAsync(&r.bitmap, j.warps * j.words * sizeof(uint32_t), stream));
    HERA_CUDA(cudaMallocAsync(&r.nodes, j.warps * j.cap * sizeof(uint32_t), stream));
    HERA_CUDA(cudaMallocAsync(&r.slots, j.warps * j.cap * sizeof(uint32_t), stream));
    HERA_CUDA(cudaMallocAsync(&r.vlog, j.warps * j.cap * sizeof(uint32_t), stream));
    HERA_CUDA(cudaMemsetAsync(r.bitmap, 0, j.warps * j.words * sizeof(uint32_t), stream));
}

// Sizes the job, allocates its scratch (stream-ordered on the context's stream) and fills the walk table
// of a greedy job.  warps_per_sm: this job's share of the resident walk warps.
void job_prepare(Ctx &c, WalkJob &j, int warps_per_sm, int slot) {
    GraphTables &g = c.g;
    HERA_REQUIRE(g.ready, 7, "walks: call hera_graph_finalize first");
    HERA_REQUIRE(j.metric >= 0 && j.metric < 4, 8, "unknown metric");
    const int ranks = xch_size(c), me = xch_rank(c);
    HERA_REQUIRE(ranks <= WK_MAX_RANKS, 37, "more than 16 ranks are not supported");
    j.wg = make_walk_graph(c, j.metric);
    j.home = c.stream;
    const uint64_t launches0 = c.launches.count;
    if (j.kind == 0) {
        const uint64_t total = g.anchors * static_cast<uint64_t>(j.pp.rebuild_attempts > 0 ? j.pp.rebuild_attempts : 0);
        HERA_REQUIRE(total < (1ull << 32), 9, "too many Monte-Carlo walks for one call");
        j.total_walks = static_cast<uint32_t>(total);
    } else {
        HERA_REQUIRE(g.E == 0 || g.rank_ready[j.metric], 10, "rank permutation for this metric was not built (metrics_mask of hera_graph_finalize)");
        j.total_walks = 0;
        if (g.anchors) {
            TempPool tmp(c.stream);
            uint32_t *d_total = tmp.get<uint32_t>(1);
            HERA_CUDA(cudaMallocAsync(&j.first_edges, g.anchors * sizeof(uint32_t), c.stream));
            HERA_CUDA(cudaMallocAsync(&j.walk_base, g.anchors * sizeof(uint32_t), c.stream));
            HERA_LAUNCH(c, det_count_kernel, div_up(g.anchors * 32, 128), 128, 0, j.wg, j.first_edges);
            exclusive_scan_u32(c, j.first_edges, g.anchors, j.walk_base, d_total);
            HERA_CUDA(cudaMemcpyAsync(&j.total_walks, d_total, sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
            HERA_CUDA(cudaStreamSynchronize(c.stream));
        }
    }
    j.my_walks = walk_share(j.total_walks, ranks, me);
    HERA_CUDA(cudaEventCreate(&j.ev0));
    HERA_CUDA(cudaEventCreate(&j.ev1));
    c.walk_h_ctr.reserve(64);   // pinned once per context (cudaMallocHost / cudaFreeHost per job cost more than the small kernels around them)
    j.h_ctr = c.walk_h_ctr.p + 4 * slot;
    for (int k = 0; k < 4; k++) j.h_ctr[k] = 0;
    if (j.total_walks == 0) { j.launches = c.launches.count - launches0; return; }
    // A walk may visit every node once, but a typical one visits a few hundred: the per-warp scratch (node / slot /
    // visited logs) is sized for opt_walk_cap_small (16384) nodes so that a large graph does not cost resident warps (N = 420 k
    // nodes would mean 5 MB per warp); a walk that outgrows it raises WalkOut::overflow and the job is run again
    // with room for N nodes (remembered per kind and metric).
    j.words = static_cast<uint32_t>((g.N + 31) / 32);
    j.warps_per_sm = warps_per_sm;
    job_alloc_scratch(c, j, c.walk_full_cap[j.kind][j.metric] != 0, c.stream);
    const uint64_t res_cap = std::max<uint32_t>(j.my_walks, 1);
    WalkRun &r = j.r;
    HERA_CUDA(cudaMallocAsync(&r.res_n, res_cap * sizeof(uint32_t), c.stream));
    HERA_CUDA(cudaMallocAsync(&r.res_draws, res_cap * sizeof(uint32_t), c.stream));
    HERA_CUDA(cudaMallocAsync(&r.res_off, res_cap * sizeof(uint64_t), c.stream));
    HERA_CUDA(cudaMallocAsync(&r.res_len, res_cap * sizeof(int64_t), c.stream));
    HERA_CUDA(cudaMallocAsync(&r.d_ctr, 4 * sizeof(unsigned long long), c.stream));
    j.pool_cap = std::max<uint64_t>(std::max<uint64_t>(1u << 16, c.walk_pool_hint[j.kind][j.metric]), std::min<uint64_t>(static_cast<uint64_t>(j.my_walks) * 64, 1ull << 28));
    HERA_CUDA(cudaMallocAsync(&r.pool_nodes, j.pool_cap * sizeof(uint32_t), c.stream));
    HERA_CUDA(cudaMallocAsync(&r.pool_slots, j.pool_cap * sizeof(uint32_t), c.stream));
    j.launches = c.launches.count - launches0;
}

// Waits for the job's kernel, re-runs it with a larger pool if the accepted paths did not fit (walks are
// deterministic), exchanges results between ranks and appends the accepted paths to c.all.
void job_finish(Ctx &c, WalkJob &j, hera_walk_stats *out) {
    const int ranks = xch_size(c);
    const uint64_t launches0 = c.launches.count;
    uint64_t found = 0, steps = 0;
    float ms = 0;
    const uint32_t total_walks = j.total_walks;
    if (total_walks > 0) {
        WalkRun &r = j.r;
        for (;;) {
            HERA_CUDA(cudaStreamSynchronize(j.stream));
            const bool scratch_overflow = j.h_ctr[3] != 0;
            if (j.h_ctr[0] <= j.pool_cap && !scratch_overflow) break;
            if (scratch_overflow) {   // some walk visited more nodes than the small scratch holds: again with room for every node
                HERA_REQUIRE(j.cap < c.g.N + 1, 38, "internal: walk scratch overflow at full capacity");
                c.walk_full_cap[j.kind][j.metric] = 1;
                for (uint32_t **q : {&r.bitmap, &r.nodes, &r.slots, &r.vlog}) {
                    HERA_CUDA(cudaFreeAsync(*q, j.stream));
                    *q = nullptr;
                }
                job_alloc_scratch(c, j, true, j.stream);
            }
            if (j.h_ctr[0] > j.pool_cap) {
                for (uint32_t **q : {&r.pool_nodes, &r.pool_slots}) {
                    HERA_CUDA(cudaFreeAsync(*q, j.stream));
                    *q = nullptr;
                }
                j.pool_cap = j.h_ctr[0] + 1024;
                HERA_CUDA(cudaMallocAsync(&r.pool_nodes, j.pool_cap * sizeof(uint32_t), j.stream));
                HERA_CUDA(cudaMallocAsync(&r.pool_slots, j.pool_cap * sizeof(uint32_t), j.stream));
            }
            job_launch_kernel(c, j);
        }
        float ms_now = 0;
        HERA_CUDA(cudaEventElapsedTime(&ms_now, j.ev0, j.ev1));
        c.walk_work_hint[j.kind][j.metric] = static_cast<uint64_t>(static_cast<double>(ms_now) * 1e3 * static_cast<double>(j.warps)) + 1;   // warp-microseconds
        HERA_CUDA(cudaEventElapsedTime(&ms, j.ev0, j.ev1));
        steps = j.h_ctr[1];
        uint64_t &hint = c.walk_pool_hint[j.kind][j.metric];
        hint = std::max<uint64_t>(hint, j.h_ctr[0] + j.h_ctr[0] / 8);
        uint64_t new_nodes = j.h_ctr[0];

        // per-walk results in walk-id order + one pool holding every accepted path
        uint32_t *res_n = r.res_n, *res_draws = r.res_draws, *pool_nodes = r.pool_nodes, *pool_slots = r.pool_slots;
        uint64_t *res_off = r.res_off;
        int64_t *res_len = r.res_len;
        TempPool tmp(c.stream);   // gathered copies, flags and scans: gone when the scope ends, exception or not
        if (ranks > 1) {
            // exchange step: all-gather of the ranks' accepted paths
            const uint64_t mine[2] = {j.h_ctr[0], j.h_ctr[1]};
            std::vector<uint64_t> all(2 * static_cast<size_t>(ranks));
            c.xch->all_gather_host(c, mine, 2, all.data());
            RankTable t{};
            std::vector<uint64_t> wcount(ranks), wdispl(ranks), pcount(ranks), pdispl(ranks);
            uint64_t wtot = 0, ptot = 0;
            steps = 0;
            for (int k = 0; k < ranks; k++) {
                wcount[k] = walk_share(total_walks, ranks, k); wdispl[k] = wtot; wtot += wcount[k];
                pcount[k] = all[2 * k]; pdispl[k] = ptot; ptot += pcount[k];
                steps += all[2 * k + 1];
                t.walk_displ[k] = wdispl[k]; t.pool_displ[k] = pdispl[k];
            }
            uint32_t *g_n, *g_draws, *g_pn, *g_ps, *o_n, *o_draws;
            uint64_t *g_off, *o_off;
            int64_t *g_len, *o_len;
            auto alloc = [&](auto **p, uint64_t n) { *p = tmp.get<std::remove_pointer_t<std::remove_pointer_t<decltype(p)>>>(n); };
            alloc(&g_n, wtot); alloc(&g_draws, wtot); alloc(&g_off, wtot); alloc(&g_len, wtot); alloc(&g_pn, ptot); alloc(&g_ps, ptot);
            alloc(&o_n, wtot); alloc(&o_draws, wtot); alloc(&o_off, wtot); alloc(&o_len, wtot);
            // two row gathers (per-walk results, path pools): over peer windows on the NCCL transport (pack, mailbox
            // rendezvous, pull over NVLink), else one batch of all_gatherv calls
            const Exchange::Column wcols[4] = {{g_n, 4, r.res_n}, {g_draws, 4, r.res_draws}, {g_off, 8, r.res_off}, {g_len, 8, r.res_len}};
            const Exchange::Column pcols[2] = {{g_pn, 4, r.pool_nodes}, {g_ps, 4, r.pool_slots}};
            c.xch->gather_rows(c, wcols, 4, 0, 0, wcount.data(), wdispl.data());
            c.xch->gather_rows(c, pcols, 2, 0, 0, pcount.data(), pdispl.data());
            HERA_LAUNCH(c, walk_order_kernel, div_up(total_walks, 256), 256, 0, total_walks, static_cast<uint32_t>(ranks), t, g_n, g_off, g_len, g_draws,
                        o_n, o_off, o_len, o_draws);
            res_n = o_n; res_draws = o_draws; res_off = o_off; res_len = o_len; pool_nodes = g_pn; pool_slots = g_ps;
            new_nodes = ptot;
        }

        // order by walk id and append to the path store
        uint8_t *flags = tmp.get<uint8_t>(total_walks);
        uint32_t *acc_scan = tmp.get<uint32_t>(total_walks), *d_found = tmp.get<uint32_t>(1);
        uint64_t *node_scan = tmp.get<uint64_t>(total_walks);
        HERA_LAUNCH(c, flag_nonzero_kernel, div_up(total_walks, 256), 256, 0, total_walks, res_n, flags);
        exclusive_scan_u8_to_u32(c, flags, total_walks, acc_scan, d_found);
        exclusive_scan_u32_to_u64(c, res_n, total_walks, node_scan, nullptr);
        uint32_t h_found = 0;
        HERA_CUDA(cudaMemcpyAsync(&h_found, d_found, sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        found = h_found;
        PathStore &ps = c.all;
        ps.off.reserve(ps.paths + found + 1, true, ps.paths + 1, c.stream);
        ps.len.reserve(ps.paths + found, true, ps.paths, c.stream);
        ps.node.reserve(ps.nodes + new_nodes, true, ps.nodes, c.stream);
        ps.slot.reserve(ps.nodes + new_nodes, true, ps.nodes > ps.paths ? ps.nodes - ps.paths : 0, c.stream);
        if (ps.paths == 0) HERA_CUDA(cudaMemsetAsync(ps.off.p, 0, sizeof(uint64_t), c.stream));
        if (found) {
            HERA_LAUNCH(c, gather_paths_kernel, div_up(static_cast<uint64_t>(total_walks) * 32, WK_THREADS), WK_THREADS, 0, total_walks, res_n, res_off,
                        res_len, acc_scan, node_scan, pool_nodes, pool_slots, ps.paths, ps.nodes, ps.off.p, ps.node.p, ps.slot.p, ps.len.p);
        }
        ps.paths += found;
        ps.nodes += new_nodes;
        if (j.kind == 0) {
            const size_t before = c.mc_draws.size();
            c.mc_draws.resize(before + total_walks);
            HERA_CUDA(cudaMemcpyAsync(c.mc_draws.data() + before, res_draws, total_walks * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
        }
        HERA_CUDA(cudaStreamSynchronize(c.stream));
    }
    j.release(false);   // (the job's kernel has finished: the stream was synchronised above)
    j.h_ctr = nullptr;
    c.uniq_ready = false;
    c.groups_valid = false;
    c.call_kind.push_back(static_cast<uint32_t>(j.kind));
    c.call_metric.push_back(static_cast<uint32_t>(j.metric));
    c.call_found.push_back(found);
    if (out) {
        out->walks = total_walks; out->found = found; out->steps = steps; out->ms = ms;
        out->launches = static_cast<uint32_t>(j.launches + (c.launches.count - launches0));
    }
}

}  // namespace

// kinds[i]: 0 = Monte Carlo, 1 = greedy.  All n kernels are in flight together (one stream each); the
// accepted paths are appended in call order, exactly as n separate calls would leave them.
void walk_build(Ctx &c, int n, const int *kinds, const int *metrics, const hera_path_params &pp, uint64_t seed, hera_walk_stats *out) {
    HERA_REQUIRE(n >= 1 && n <= 8, 101, "hera_paths_build: between 1 and 8 calls");
    std::vector<WalkJob> jobs(n);
    // HERA_DEBUG_TIMING: host clock at the phase ends of the call (prepare = sizing + scratch, launch, finish i = wait + gather + append)
    const bool debug = getenv("HERA_DEBUG_TIMING") != nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    std::string phases;
    auto stamp = [&](const char *name) {
        if (!debug) return;
        char buf[64];
        snprintf(buf, sizeof buf, " %s +%.3f", name, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count());
        phases += buf;
    };
    const int budget = c.opt_walk_warps_per_sm;
    // Every call gets the whole budget of resident warps.  The kernels are persistent (walks are handed out by ticket), so
    // the CTAs of a kernel launched later start as the CTAs of an earlier one run out of walks: back to back, with the
    // tail of each kernel (its longest walks) hidden behind the next.  Measured on configs[3] (800 k walks per
    // Monte-Carlo call): the four kernels alone take 21 + 60 + 13 + 13 ms, side by side with split budgets 127 ms.
    // The budget is what fits at once: 64 registers per thread allow 32 warps per SM.
    for (int i = 0; i < n; i++) {
        WalkJob &j = jobs[i];
        j.kind = kinds[i] ? 1 : 0; j.metric = metrics[i]; j.pp = pp; j.seed = seed;
        job_prepare(c, j, budget, i);
    }
    stamp("prepared");
    if (n == 1) {
        jobs[0].stream = c.stream;
        if (jobs[0].total_walks) job_launch_kernel(c, jobs[0]);
    } else {
        while (c.aux_streams.size() < static_cast<size_t>(n)) {
            cudaStream_t s;
            HERA_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
            c.aux_streams.push_back(s);
        }
        Events<1> ready_ev(cudaEventDisableTiming);
        const cudaEvent_t ready = ready_ev[0];
        HERA_CUDA(cudaEventRecord(ready, c.stream));   // scratch allocation and the graph itself are ordered on the context's stream
        // longest first (warp-time of the same call the last time on this context; call order when nothing is known)
        std::vector<int> order(n);
        for (int i = 0; i < n; i++) order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](int x, int y) {
            return c.walk_work_hint[jobs[x].kind][jobs[x].metric & 3] > c.walk_work_hint[jobs[y].kind][jobs[y].metric & 3];
        });
        for (int i : order) {
            jobs[i].stream = c.aux_streams[i];
            HERA_CUDA(cudaStreamWaitEvent(jobs[i].stream, ready, 0));
            if (jobs[i].total_walks) job_launch_kernel(c, jobs[i]);
        }
    }
    stamp("launched");
    for (int i = 0; i < n; i++) {
        job_finish(c, jobs[i], out ? out + i : nullptr);
        stamp("finished");
    }
    if (debug) {
        std::string shares;
        for (int i = 0; i < n; i++) shares += " " + std::to_string(jobs[i].warps) + "w/cap" + std::to_string(jobs[i].cap);
        fprintf(stderr, "[hera rank %d] paths_build host clock (ms):%s; warps / scratch capacity per job:%s\n", xch_rank(c), phases.c_str(), shares.c_str());
    }
}

// ---- measured ceiling for list walking (north_star: "the pointer-chasing walks are reported against the measured
// random-gather roofline") -------------------------------------------------------------------------------------
// The same memory behaviour as a walk step and nothing else: a warp reads the whole adjacency list of its current
// node as 8-byte entries (what the roulette scan reads), then hops to a neighbour it found there (the next address depends on loaded data), with
// the walk kernels' number of resident warps.  No roulette, no visited set, no random numbers.
__global__ void __launch_bounds__(WK_THREADS) list_gather_probe_kernel(const uint32_t *__restrict__ off, const uint2 *__restrict__ adj, uint32_t N,
                                                                       uint32_t hops, unsigned long long *bytes_out, uint32_t *sink) {
    const uint32_t lane = threadIdx.x & 31u, w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint32_t x = (w * 2654435761u) % N, acc = 0;
    unsigned long long bytes = 0;
    for (uint32_t h = 0; h < hops; h++) {
        const uint32_t base = off[x], deg = off[x + 1] - base;
        const uint32_t pick = deg ? ((w + h) * 0x9E3779B1u >> 8) % deg : 0;
        uint32_t cand = x;
        for (uint32_t k = lane; k < deg; k += 32) {
            const uint2 e = adj[base + k];
            acc ^= e.y;
            if (k == pick) cand = e.x & 0x7fffffffu;
        }
        bytes += static_cast<unsigned long long>(deg) * 8;
        x = deg ? __shfl_sync(0xffffffffu, cand, pick & 31u) : (x + 1) % N;
    }
    if (lane == 0) atomicAdd(bytes_out, bytes);
    if (acc == 0x12345678u) *sink = acc;
}

void probe_list_gather(Ctx &c, uint32_t warps_per_sm, uint32_t hops, double *gbytes_per_s, float *ms_out) {
    GraphTables &g = c.g;
    HERA_REQUIRE(g.ready && g.N > 0, 102, "hera_probe_list_gather: finalize a graph first");
    int sms = 148;
    HERA_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c.device));
    const unsigned warps = static_cast<unsigned>(sms) * std::max(1u, std::min(64u, warps_per_sm));
    TempPool tmp(c.stream);
    unsigned long long *d_bytes = tmp.zeroed<unsigned long long>(2);
    Events<2> ev;
    const cudaEvent_t e0 = ev[0], e1 = ev[1];
    HERA_CUDA(cudaEventRecord(e0, c.stream));
    HERA_LAUNCH(c, list_gather_probe_kernel, div_up(static_cast<uint64_t>(warps) * 32, WK_THREADS), WK_THREADS, 0, g.adj_off.p, g.a8[0].p,
                static_cast<uint32_t>(g.N), hops, d_bytes, reinterpret_cast<uint32_t *>(d_bytes + 1));
    HERA_CUDA(cudaEventRecord(e1, c.stream));
    unsigned long long h_bytes = 0;
    HERA_CUDA(cudaMemcpyAsync(&h_bytes, d_bytes, sizeof(h_bytes), cudaMemcpyDeviceToHost, c.stream));
    HERA_CUDA(cudaStreamSynchronize(c.stream));
    float ms = 0;
    HERA_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    if (gbytes_per_s) *gbytes_per_s = ms > 0 ? static_cast<double>(h_bytes) / (ms * 1e-3) / 1e9 : 0.0;
    if (ms_out) *ms_out = ms;
}

void walk_monte_carlo(Ctx &c, int metric, const hera_path_params &pp, uint64_t seed, hera_walk_stats *out) {
    const int kind = 0;
    walk_build(c, 1, &kind, &metric, pp, seed, out);
}

void walk_deterministic(Ctx &c, int metric, hera_walk_stats *out) {
    const int kind = 1;
    hera_path_params pp{};
    walk_build(c, 1, &kind, &metric, pp, 0, out);
}

}  // namespace hera
