// K8: anchor-pair bucketing, length windows, groups, consensus.
//
// Replaces PathManager::getPathsBetweenAnchors / constructGroups / getBorderPathLengths
// (reference src/PathManager.cpp:508-645), PathWindow (src/PathWindow.cpp), PathGroup
// (src/PathGroup.cpp) and the inter-group pick of main() (src/Main.cpp:306-332).
//
// Paths are keyed by (first anchor, last anchor) and stably radix-sorted, which reproduces the
// std::map iteration order (node pointers order like node ids) with paths_ order inside a key.
// One CTA then owns one anchor pair: thread 0 replays libstdc++'s introsort (introsort.cuh) so
// that ties in length fall exactly as in the reference, and cuts the sorted list into groups;
// the CTA evaluates each group cooperatively -- length frequencies are run lengths of the
// sorted list, the most frequent length is a block max-reduction, discardNotFrequent an ordered
// block compaction, the int32-wrapping mean of calculateConsensusPath a block sum modulo 2^32,
// valid_path_number a block count of node-sequence matches.
#include <algorithm>

#include "ctx.cuh"
#include "introsort.cuh"
#include "primitives.cuh"

namespace hera {

constexpr int GP_THREADS = 128;

struct GroupArgs {
    uint32_t pairs;
    const uint32_t *pair_off;       // pairs + 1
    uint32_t *ids;                  // path ids, pair-major; sorted in place by length
    const int64_t *len;
    const uint64_t *off;
    const uint32_t *nodes;
    long long len_threshold;
    unsigned long long window_size;
    float ratio_threshold;
    uint32_t smem_items;            // SortItem capacity of the CTA's dynamic shared memory
    // scratch, all indexed like ids
    uint32_t *freq, *w_tot, *w_lowc, *w_highc;
    unsigned long long *w_lowl, *w_highl, *borders;
    uint2 *pick;
    // outputs
    uint32_t *grp_count;            // per pair
    uint32_t *grp_size0, *grp_kept; // group g of pair k at pair_off[k] + g
    long long *grp_cons;
    int *grp_valid;
    uint32_t *pig;                  // members after discardNotFrequent, stored from the group's own start
    long long *pair_cons;
};

// (length, path id) of one path while a pair's list is sorted in shared memory
struct SortItem {
    long long len;
    uint32_t id, pad;
};

// Thread 0: cut the sorted list into groups (src/PathManager.cpp:515-581, 586-624).  L(i) = length of
// the i-th path of the sorted list.
template <typename LenFn>
__device__ uint32_t split_groups(const GroupArgs &a, uint32_t b, uint32_t n, LenFn L) {
    uint32_t *gsize = a.grp_size0 + b;
    const unsigned long long min_len = L(0), max_len = L(n - 1);
    if (max_len - min_len < static_cast<unsigned long long>(a.len_threshold)) { gsize[0] = n; return 1; }
    uint32_t *tot = a.w_tot + b, *lowc = a.w_lowc + b, *highc = a.w_highc + b;
    unsigned long long *lowl = a.w_lowl + b, *highl = a.w_highl + b, *borders = a.borders + b;
    // non-empty windows [min + k*ws, min + (k+1)*ws) in ascending order (src/PathWindow.cpp:6-23)
    uint32_t nw = 0, i = 0;
    while (i < n) {
        const unsigned long long k = (L(i) - min_len) / a.window_size;
        const unsigned long long upper = min_len + (k + 1) * a.window_size;
        uint32_t t = 0, lc = 0, hc = 0;
        unsigned long long ll = 0, hl = 0;
        while (i < n && L(i) < upper) {
            const unsigned long long l = L(i);
            uint32_t c = 0;
            while (i < n && L(i) == l) { c++; i++; }
            if (t == 0 || c < lc) { lc = c; ll = l; }   // first minimum in ascending length order
            if (t == 0 || hc < c) { hc = c; hl = l; }   // first maximum
            t += c;
        }
        tot[nw] = t; lowc[nw] = lc; lowl[nw] = ll; highc[nw] = hc; highl[nw] = hl;
        nw++;
    }
    uint32_t nb = 0;
    for (uint32_t w = 1; w + 1 < nw; w++) {
        uint32_t valley = w - 1, peak = w - 1;
        for (uint32_t c = w; c <= w + 1; c++) {
            if (tot[c] < tot[valley]) valley = c;
            if (tot[peak] < tot[c]) peak = c;
        }
        if (__int2float_rn(static_cast<int>(lowc[valley])) < __fmul_rn(a.ratio_threshold, __int2float_rn(static_cast<int>(highc[peak])))) borders[nb++] = lowl[valley];
    }
    if (nb == 0) { gsize[0] = n; return 1; }
    // std::set: ascending, unique
    for (uint32_t x = 1; x < nb; x++) {
        const unsigned long long v = borders[x];
        uint32_t y = x;
        while (y > 0 && borders[y - 1] > v) { borders[y] = borders[y - 1]; y--; }
        borders[y] = v;
    }
    uint32_t nu = 0;
    for (uint32_t x = 0; x < nb; x++) if (nu == 0 || borders[nu - 1] != borders[x]) borders[nu++] = borders[x];
    uint32_t ng = 0, begin = 0;
    for (uint32_t x = 0; x < nu && begin < n; x++) {
        uint32_t end = begin + 1;
        while (end < n && L(end) < borders[x]) end++;
        gsize[ng++] = end - begin;
        begin = end;
    }
    if (begin < n) gsize[ng++] = n - begin;
    return ng;
}

__device__ __forceinline__ bool same_nodes(const GroupArgs &a, uint32_t p, uint32_t q) {
    const uint64_t pb = a.off[p], pn = a.off[p + 1] - pb, qb = a.off[q], qn = a.off[q + 1] - qb;
    if (pn != qn) return false;
    for (uint64_t i = 0; i < pn; i++) if (a.nodes[pb + i] != a.nodes[qb + i]) return false;
    return true;
}

template <typename T, typename Op>
__device__ __forceinline__ T block_reduce(T v, T *smem, Op op, T identity) {
    const uint32_t lane = lane_id(), warp = warp_id();
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, d));
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    T r = identity;
    for (uint32_t w = 0; w < (blockDim.x >> 5); w++) r = op(r, smem[w]);
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(GP_THREADS) group_pairs_kernel(const GroupArgs a) {
    __shared__ uint32_t s_u32[36];
    __shared__ uint32_t s_ng, s_run;
    const uint32_t k = blockIdx.x, tid = threadIdx.x;
    const uint32_t b = a.pair_off[k], n = a.pair_off[k + 1] - b;
    uint32_t *ids = a.ids + b;
    // The sort is libstdc++'s introsort run by one thread (tie order!); its working set -- (length, id) per
    // path -- is staged in shared memory when the pair fits, so that every compare and move is an LDS/STS.
    extern __shared__ __align__(16) unsigned char gp_smem[];
    SortItem *items = reinterpret_cast<SortItem *>(gp_smem);
    const bool staged = n <= a.smem_items;
    if (staged) {
        for (uint32_t i = tid; i < n; i += GP_THREADS) { const uint32_t id = ids[i]; items[i] = SortItem{a.len[id], id, 0u}; }
        __syncthreads();
        if (tid == 0) {
            libstdcxx_sort(items, static_cast<long>(n), [](const SortItem &x, const SortItem &y) { return x.len < y.len; });
            s_ng = split_groups(a, b, n, [&](uint32_t i) { return static_cast<unsigned long long>(items[i].len); });
            a.grp_count[k] = s_ng;
        }
        __syncthreads();
        for (uint32_t i = tid; i < n; i += GP_THREADS) ids[i] = items[i].id;
    } else if (tid == 0) {
        libstdcxx_sort(ids, static_cast<long>(n), [&](uint32_t x, uint32_t y) { return a.len[x] < a.len[y]; });
        s_ng = split_groups(a, b, n, [&](uint32_t i) { return static_cast<unsigned long long>(a.len[ids[i]]); });
        a.grp_count[k] = s_ng;
    }
    __syncthreads();
    const uint32_t ng = s_ng;
    uint32_t gb = 0;
    for (uint32_t g = 0; g < ng; g++) {
        const uint32_t gn = a.grp_size0[b + g];
        const uint32_t *gid = ids + gb;
        uint32_t *freq = a.freq + b + gb;
        uint32_t *pig = a.pig + b + gb;
        // length frequencies = run lengths (PathGroup ctor, src/PathGroup.cpp:17-22)
        uint32_t top = 0;
        for (uint32_t i = tid; i < gn; i += GP_THREADS) {
            const long long l = a.len[gid[i]];
            if (i == 0 || a.len[gid[i - 1]] != l) {
                uint32_t j = i + 1;
                while (j < gn && a.len[gid[j]] == l) j++;
                for (uint32_t t = i; t < j; t++) freq[t] = j - i;
                top = max(top, j - i);
            }
        }
        top = block_reduce<uint32_t>(top, s_u32, [](uint32_t x, uint32_t y) { return max(x, y); }, 0u);
        const uint32_t thr = top / 2;  // discardNotFrequent, src/PathGroup.cpp:40-55
        // ordered compaction of the survivors
        if (tid == 0) s_run = 0;
        __syncthreads();
        for (uint32_t c0 = 0; c0 < gn; c0 += GP_THREADS) {
            const uint32_t i = c0 + tid;
            const uint32_t keep = (i < gn && (thr == 0 || freq[i] >= thr)) ? 1u : 0u;
            uint32_t total;
            const uint32_t ex = block_exclusive_scan<uint32_t>(keep, s_u32, &total);
            if (keep) pig[s_run + ex] = gid[i];
            __syncthreads();
            if (tid == 0) s_run += total;
            __syncthreads();
        }
        const uint32_t kn = s_run;
        long long cons = -1;
        int valid = 0;
        if (kn > 0) {
            const long long front = a.len[pig[0]], back = a.len[pig[kn - 1]];
            if (!(static_cast<unsigned long long>(back - front) > 100000ull)) {  // include/PathGroup.hpp:55
                // std::accumulate with an int accumulator: the sum is cut to 32 bits at every step,
                // which equals the sum modulo 2^32 (src/PathGroup.cpp:66-68)
                uint32_t part = 0;
                for (uint32_t i = tid; i < kn; i += GP_THREADS) part += static_cast<uint32_t>(a.len[pig[i]]);
                const uint32_t sum32 = block_reduce<uint32_t>(part, s_u32, [](uint32_t x, uint32_t y) { return x + y; }, 0u);
                const unsigned long long avg = static_cast<unsigned long long>(static_cast<long long>(static_cast<int>(sum32))) / kn;
                uint32_t first = kn;
                for (uint32_t i = tid; i < kn; i += GP_THREADS)
                    if (static_cast<unsigned long long>(a.len[pig[i]]) >= avg) { first = i; break; }
                first = block_reduce<uint32_t>(first, s_u32, [](uint32_t x, uint32_t y) { return min(x, y); }, kn);
                cons = pig[first < kn ? first : kn - 1];
                uint32_t cnt = 0;  // calculateValidPathNumber, src/PathGroup.cpp:83-93
                for (uint32_t i = tid; i < kn; i += GP_THREADS) cnt += same_nodes(a, static_cast<uint32_t>(cons), pig[i]) ? 1u : 0u;
                valid = static_cast<int>(block_reduce<uint32_t>(cnt, s_u32, [](uint32_t x, uint32_t y) { return x + y; }, 0u));
            }
        }
        if (tid == 0) { a.grp_kept[b + g] = kn; a.grp_cons[b + g] = cons; a.grp_valid[b + g] = valid; }
        __syncthreads();
        gb += gn;
    }
    // inter-group pick (src/Main.cpp:306-332)
    if (tid == 0) {
        uint2 *pick = a.pick + b;
        uint32_t m = 0;
        for (uint32_t g = 0; g < ng; g++) if (a.grp_cons[b + g] >= 0) pick[m++] = make_uint2(static_cast<uint32_t>(a.grp_cons[b + g]), static_cast<uint32_t>(a.grp_valid[b + g]));
        long long chosen = -1;
        if (m == 1) chosen = pick[0].x;
        else if (m >= 2) {
            libstdcxx_sort(pick, static_cast<long>(m), [&](const uint2 &x, const uint2 &y) { return a.len[x.x] > a.len[y.x]; });
            uint32_t keep = 0;
            if (m > 2) for (uint32_t i = 1; i < m; i++) if (2 * static_cast<int>(pick[keep].y) <= static_cast<int>(pick[i].y)) keep = i;
            chosen = pick[keep].x;
        }
        a.pair_cons[k] = chosen;
    }
}

__global__ void anchor_rank_kernel(uint32_t anchors, const uint32_t *__restrict__ anchor_ids, uint32_t *__restrict__ rank_of_node) {
    const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a < anchors) rank_of_node[anchor_ids[a]] = a;
}

__global__ void pair_keys_kernel(uint32_t paths, uint32_t anchors, const uint64_t *__restrict__ off, const uint32_t *__restrict__ nodes,
                                 const uint32_t *__restrict__ rank_of_node, uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= paths) return;
    keys[i] = rank_of_node[nodes[off[i]]] * anchors + rank_of_node[nodes[off[i + 1] - 1]];
    vals[i] = i;
}

__global__ void pair_heads_kernel(uint32_t paths, const uint32_t *__restrict__ keys, uint8_t *__restrict__ head) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < paths) head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

__global__ void pair_offsets_kernel(uint32_t paths, const uint8_t *__restrict__ head, const uint32_t *__restrict__ head_scan, const uint32_t *__restrict__ keys,
                                    uint32_t *__restrict__ pair_off, uint32_t *__restrict__ pair_key, uint32_t pairs) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) pair_off[pairs] = paths;
    if (i < paths && head[i]) { pair_off[head_scan[i]] = i; pair_key[head_scan[i]] = keys[i]; }
}

// Longest run of equal keys (= paths of the largest anchor pair): sizes the shared memory of group_pairs_kernel.
__global__ void pair_max_kernel(uint32_t paths, const uint8_t *__restrict__ head, uint32_t *__restrict__ out_max) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= paths || !head[i]) return;
    uint32_t e = i + 1;
    while (e < paths && !head[e]) e++;
    atomicMax(out_max, e - i);
}

static int bits_for_u64(uint64_t n) {
    int b = 1;
    while ((1ull << b) < n) b++;
    return b;
}

void group_paths(Ctx &c, const hera_path_params &pp, hera_groups_view *out) {
    const PathStore &ps = c.uniq_ready ? c.uniq : c.all;
    auto &m = c.grm;
    // (cleared, not rebuilt: the vectors keep their capacity from call to call)
    m.pair_a1.clear(); m.pair_a2.clear(); m.pair_sorted.clear(); m.grp_pair.clear(); m.grp_size0.clear(); m.grp_pig.clear();
    m.pair_off.clear(); m.grp_pig_off.clear(); m.grp_consensus.clear(); m.pair_consensus.clear(); m.grp_valid.clear();
    *out = hera_groups_view{};
    HERA_REQUIRE(pp.window_size > 0, 11, "window_size must be positive");
    HERA_REQUIRE(ps.paths < (1ull << 31), 12, "too many paths");
    const uint32_t P = static_cast<uint32_t>(ps.paths), A = static_cast<uint32_t>(c.g.anchors);
    m.pair_off.push_back(0);
    m.grp_pig_off.push_back(0);
    float ms = 0;
    if (P > 0) {
        HERA_REQUIRE(static_cast<uint64_t>(A) * A < (1ull << 32), 13, "too many anchors for 32-bit pair keys");
        Events<2> ev;   // events and temporaries go when the scope ends, exception or not
        const cudaEvent_t ev0 = ev[0], ev1 = ev[1];
        HERA_CUDA(cudaEventRecord(ev0, c.stream));
        TempPool tmp(c.stream);
        auto alloc = [&](size_t bytes) { return static_cast<void *>(tmp.get<uint8_t>(bytes)); };
        uint32_t *rank_of_node = static_cast<uint32_t *>(alloc(c.g.N * sizeof(uint32_t)));
        uint32_t *keys = static_cast<uint32_t *>(alloc(P * sizeof(uint32_t))), *vals = static_cast<uint32_t *>(alloc(P * sizeof(uint32_t)));
        uint32_t *keys_alt = static_cast<uint32_t *>(alloc(P * sizeof(uint32_t))), *vals_alt = static_cast<uint32_t *>(alloc(P * sizeof(uint32_t)));
        uint8_t *head = static_cast<uint8_t *>(alloc(P));
        uint32_t *head_scan = static_cast<uint32_t *>(alloc(P * sizeof(uint32_t))), *d_pairs = static_cast<uint32_t *>(alloc(2 * sizeof(uint32_t)));   // [0] pairs [1] largest pair
        HERA_LAUNCH(c, anchor_rank_kernel, div_up(A, 128), 128, 0, A, c.g.anchor_ids.p, rank_of_node);
        HERA_LAUNCH(c, pair_keys_kernel, div_up(P, 256), 256, 0, P, A, ps.off.p, ps.node.p, rank_of_node, keys, vals);
        radix_sort_pairs(c, keys, vals, keys_alt, vals_alt, P, bits_for_u64(static_cast<uint64_t>(A) * A));
        HERA_LAUNCH(c, pair_heads_kernel, div_up(P, 256), 256, 0, P, keys, head);
        exclusive_scan_u8_to_u32(c, head, P, head_scan, d_pairs);
        HERA_CUDA(cudaMemsetAsync(d_pairs + 1, 0, sizeof(uint32_t), c.stream));
        HERA_LAUNCH(c, pair_max_kernel, div_up(P, 256), 256, 0, P, head, d_pairs + 1);
        uint32_t h_pairs[2] = {0, 0};
        HERA_CUDA(cudaMemcpyAsync(h_pairs, d_pairs, sizeof(h_pairs), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        const uint32_t K = h_pairs[0], max_pair = h_pairs[1];
        uint32_t *pair_off = static_cast<uint32_t *>(alloc((K + 1) * sizeof(uint32_t))), *pair_key = static_cast<uint32_t *>(alloc(K * sizeof(uint32_t)));
        HERA_LAUNCH(c, pair_offsets_kernel, div_up(P, 256), 256, 0, P, head, head_scan, keys, pair_off, pair_key, K);

        GroupArgs a{};
        a.pairs = K; a.pair_off = pair_off; a.ids = vals; a.len = reinterpret_cast<const int64_t *>(ps.len.p); a.off = ps.off.p; a.nodes = ps.node.p;
        a.len_threshold = pp.len_threshold; a.window_size = pp.window_size; a.ratio_threshold = pp.ratio_threshold;
        a.freq = static_cast<uint32_t *>(alloc(P * 4)); a.w_tot = static_cast<uint32_t *>(alloc(P * 4)); a.w_lowc = static_cast<uint32_t *>(alloc(P * 4));
        a.w_highc = static_cast<uint32_t *>(alloc(P * 4));
        a.w_lowl = static_cast<unsigned long long *>(alloc(P * 8)); a.w_highl = static_cast<unsigned long long *>(alloc(P * 8));
        a.borders = static_cast<unsigned long long *>(alloc(P * 8)); a.pick = static_cast<uint2 *>(alloc(P * 8));
        a.grp_count = static_cast<uint32_t *>(alloc(K * 4)); a.grp_size0 = static_cast<uint32_t *>(alloc(P * 4)); a.grp_kept = static_cast<uint32_t *>(alloc(P * 4));
        a.grp_cons = static_cast<long long *>(alloc(P * 8)); a.grp_valid = static_cast<int *>(alloc(P * 4)); a.pig = static_cast<uint32_t *>(alloc(P * 4));
        a.pair_cons = static_cast<long long *>(alloc(K * 8));
        // 16 B per path of the LARGEST pair, up to 160 KB per CTA (larger pairs sort in global memory): sized for all paths, the
        // 46 844 pairs of configs[3] (26 paths on average) ran one CTA per SM
        a.smem_items = std::min<uint32_t>(std::max<uint32_t>(max_pair, 32), 10240);
        const size_t smem = static_cast<size_t>(a.smem_items) * sizeof(SortItem);
        HERA_CUDA(cudaFuncSetAttribute(group_pairs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        HERA_LAUNCH(c, group_pairs_kernel, K, GP_THREADS, smem, a);
        HERA_CUDA(cudaEventRecord(ev1, c.stream));

        // flatten on the host (marshalling only)
        // one pinned staging block (pageable destinations would make every copy a synchronous staged one: 34 MB on configs[3])
        auto words = [](uint64_t n) { return (n + 1) / 2 * 2; };   // keep the 64-bit arrays aligned
        const uint64_t w_total = words(K + 1) + words(K) + words(P) + words(K) + 3 * words(P) + words(A) + 2 * P + 2 * K + words(P);
        c.grp_stage.reserve(w_total);
        uint32_t *w = c.grp_stage.p;
        auto take = [&](uint64_t n) { uint32_t *q = w; w += words(n); return q; };
        uint32_t *h_pair_off = take(K + 1), *h_pair_key = take(K), *h_ids = take(P), *h_cnt = take(K), *h_size0 = take(P), *h_kept = take(P), *h_pig = take(P),
                 *h_anchor = take(A);
        long long *h_cons = reinterpret_cast<long long *>(take(2 * P)), *h_pair_cons = reinterpret_cast<long long *>(take(2 * K));
        int *h_valid = reinterpret_cast<int *>(take(P));
        auto get = [&](void *dst, const void *src, size_t bytes) { if (bytes) HERA_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c.stream)); };
        get(h_pair_off, pair_off, (K + 1) * 4); get(h_pair_key, pair_key, K * 4); get(h_ids, vals, P * 4);
        get(h_cnt, a.grp_count, K * 4); get(h_size0, a.grp_size0, P * 4); get(h_kept, a.grp_kept, P * 4);
        get(h_pig, a.pig, P * 4); get(h_cons, a.grp_cons, P * 8); get(h_valid, a.grp_valid, P * 4);
        get(h_pair_cons, a.pair_cons, K * 8); get(h_anchor, c.g.anchor_ids.p, A * 4);
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        HERA_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));

        m.pair_sorted.assign(h_ids, h_ids + P);
        m.pair_a1.reserve(K); m.pair_a2.reserve(K); m.pair_off.reserve(K + 1); m.pair_consensus.reserve(K); m.grp_pig.reserve(P);
        for (uint32_t k = 0; k < K; k++) {
            m.pair_a1.push_back(h_anchor[h_pair_key[k] / A]);
            m.pair_a2.push_back(h_anchor[h_pair_key[k] % A]);
            m.pair_off.push_back(h_pair_off[k + 1]);
            m.pair_consensus.push_back(h_pair_cons[k]);
            const uint32_t b = h_pair_off[k];
            uint32_t gb = 0;
            for (uint32_t g = 0; g < h_cnt[k]; g++) {
                m.grp_pair.push_back(k);
                m.grp_size0.push_back(h_size0[b + g]);
                m.grp_pig.insert(m.grp_pig.end(), h_pig + b + gb, h_pig + b + gb + h_kept[b + g]);
                m.grp_pig_off.push_back(m.grp_pig.size());
                m.grp_consensus.push_back(h_cons[b + g]);
                m.grp_valid.push_back(h_valid[b + g]);
                gb += h_size0[b + g];
            }
        }
    }
    out->pairs = m.pair_a1.size(); out->groups = m.grp_pair.size(); out->sorted_total = m.pair_sorted.size(); out->pig_total = m.grp_pig.size();
    out->pair_a1 = m.pair_a1.data(); out->pair_a2 = m.pair_a2.data(); out->pair_off = m.pair_off.data(); out->pair_sorted = m.pair_sorted.data();
    out->grp_pair = m.grp_pair.data(); out->grp_size0 = m.grp_size0.data(); out->grp_pig_off = m.grp_pig_off.data(); out->grp_pig = m.grp_pig.data();
    out->grp_consensus = reinterpret_cast<const int64_t *>(m.grp_consensus.data()); out->grp_valid = m.grp_valid.data();
    out->pair_consensus = reinterpret_cast<const int64_t *>(m.pair_consensus.data());
    out->ms = ms;
}

}  // namespace hera
