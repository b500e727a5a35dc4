// Context of the hera_b200 library: every device-resident table of the hot path.
#pragma once

#include <memory>

#include "common.cuh"
#include "exchange.cuh"

namespace hera {

// Bits of KeptTable::flags
constexpr uint8_t KF_MINUS = 1;     // relative strand '-' (src/OverlapGraph.cpp:240)
constexpr uint8_t KF_CR = 2;        // record came from the contig-read load (anchors == true)
constexpr uint8_t KF_QCTG = 4;      // query name starts with "ctg" (case-insensitive), src/OverlapGraph.cpp:134

// Name reference: which loaded text, byte offset, length.
__host__ __device__ inline uint64_t pack_name(uint32_t file, uint64_t off, uint32_t len) {
    return (static_cast<uint64_t>(file) << 60) | (off << 16) | (len & 0xffffu);
}
__host__ __device__ inline uint32_t name_file(uint64_t r) { return static_cast<uint32_t>(r >> 60); }
__host__ __device__ inline uint64_t name_off(uint64_t r) { return (r >> 16) & ((1ull << 44) - 1); }
__host__ __device__ inline uint32_t name_len(uint64_t r) { return static_cast<uint32_t>(r & 0xffffu); }

// Overlaps that passed the filter, file order, rr load(s) first ("edges_" before node ids exist).
struct KeptTable {
    uint64_t n = 0;
    DBuf<uint32_t> qs, qe, ts, te, qlen, tlen, rec;
    DBuf<float> os, si, es;
    DBuf<uint8_t> flags;
    DBuf<uint64_t> qhash, thash, qname, tname;
    // keep = rows whose contents must survive a reallocation (default: the n committed rows)
    void reserve(uint64_t total, cudaStream_t s, uint64_t keep = ~0ull) {
        const uint64_t k = keep == ~0ull ? n : keep;
        qs.reserve(total, true, k, s); qe.reserve(total, true, k, s); ts.reserve(total, true, k, s); te.reserve(total, true, k, s);
        qlen.reserve(total, true, k, s); tlen.reserve(total, true, k, s); rec.reserve(total, true, k, s);
        os.reserve(total, true, k, s); si.reserve(total, true, k, s); es.reserve(total, true, k, s);
        flags.reserve(total, true, k, s);
        qhash.reserve(total, true, k, s); thash.reserve(total, true, k, s); qname.reserve(total, true, k, s); tname.reserve(total, true, k, s);
    }
};

// Raw pointers of KeptTable for kernels.
struct KeptPtrs {
    uint32_t *qs, *qe, *ts, *te, *qlen, *tlen, *rec;
    float *os, *si, *es;
    uint8_t *flags;
    uint64_t *qhash, *thash, *qname, *tname;
    uint64_t cap = ~0ull;   // rows the columns can hold
};

// A byte-range slice of a PAF file (sharded load): which of its tokens are this rank's.
struct ShardSpec {
    uint64_t owned;     // tokens starting at a byte >= owned belong to the next rank
    bool prev_nonws;    // the file byte before the slice is part of a token
    bool at_eof;        // the slice (with its look-ahead) ends at the end of the file
};

struct RowRange {
    uint64_t begin, end;
};

// Kept overlaps stay on the rank that parsed them.  One load contributes one piece: the local rows
// [local0, local0 + count) of this rank's table are the rows [global0, global0 + count) of the file-order table the
// reference calls edges_ (rr loads first, ranks in byte-range order inside a load).  A single rank has local == global.
constexpr int MAX_LOADS = 16;
struct RowPieces {
    int n = 0;
    uint64_t local0[MAX_LOADS], global0[MAX_LOADS], count[MAX_LOADS];
    __host__ __device__ uint64_t global_of(uint64_t l) const {
        for (int i = n - 1; i > 0; i--) if (l >= local0[i]) return global0[i] + (l - local0[i]);
        return global0[0] + l;
    }
    __host__ __device__ bool local_of(uint64_t g, uint64_t *l) const {
        for (int i = 0; i < n; i++) if (g >= global0[i] && g < global0[i] + count[i]) { *l = local0[i] + (g - global0[i]); return true; }
        return false;
    }
};

// What the ranks told each other about one load: kept rows per rank (global order = rank order inside the load).
struct LoadShape {
    uint64_t global0 = 0, total = 0;
    std::vector<uint64_t> counts, displs;
};

// Read-only view of the file-order table of kept overlaps with their node ids (what export, cache and scaffold read).
struct RowView {
    uint64_t n = 0;
    const uint32_t *qs, *qe, *ts, *te, *qlen, *tlen, *rec, *e_q, *e_t;
    const float *os, *si, *es;
    const uint8_t *flags;
};

struct LoadedText {
    DBuf<uint8_t> bytes;  // padded with '\n' up to a multiple of the tile size plus one tile
    uint64_t n = 0;
    const uint8_t *view = nullptr;   // hera_graph_load_device_view: the caller's buffer, parsed in place (nothing readable behind byte n)
    const uint8_t *data() const { return view ? view : bytes.p; }
};

struct GraphTables {
    bool ready = false;
    uint64_t N = 0, anchors = 0, E = 0, S = 0;
    uint32_t max_deg = 0;                    // longest adjacency list
    DBuf<uint32_t> e_q, e_t;                 // node ids per LOCAL kept overlap
    DBuf<uint32_t> node_len;
    DBuf<uint8_t> node_anchor;
    DBuf<uint32_t> node_pos;                 // first appearance of the node's name: 2 * global kept row + side (the row lives on one rank)
    DBuf<uint64_t> node_name_off;            // N + 1, into node_names
    DBuf<char> node_names;                   // every node's name, node order (complete on every rank)
    uint64_t name_bytes = 0;
    DBuf<uint32_t> adj_off;                  // N + 1
    DBuf<uint32_t> adj_pay;                  // slot -> 2 * kept + side
    // walk-facing view of a slot: {neighbour id | bit 31 = neighbour is an anchor, es, os, q_start - t_start (src/Path.cpp:13)},
    // one 128-bit load per entry.  Sharded runs: every rank writes its entries into every rank's array (peer memory).
    DBuf<uint4> adj4;
    DBuf<uint2> a8[2];                       // {neighbour | anchor bit, es} and {neighbour | anchor bit, os}: what the roulette scan streams (8 B / entry; derived from adj4 on every rank)
    DBuf<uint8_t> node_exact;                // bit m: the fp64 sum of metric m over any subset of the node's list is exact in any order
    DBuf<uint32_t> rank[4];                  // per metric; empty if not built
    bool rank_ready[4] = {false, false, false, false};
    DBuf<uint32_t> canon;
    DBuf<uint32_t> anchor_ids;               // anchor node ids in node order
};

struct PathStore {
    uint64_t paths = 0, nodes = 0;
    DBuf<uint64_t> off;      // paths + 1
    DBuf<uint32_t> node, slot;
    DBuf<int64_t> len;
};

// A text that is still in host memory (hera_graph_load_bytes): copied and parsed chunk by chunk.
constexpr uint64_t LOAD_CHUNK_GRAIN = 30720;      // the parse kernel's tile
constexpr uint64_t LOAD_CHUNK_MIN_GRAINS = 8;     // a record (two names below 64 KiB each) ends inside the chunk behind the one it starts in
struct HostSource {
    const uint8_t *text;
    uint64_t chunk_bytes;   // a multiple of LOAD_CHUNK_GRAIN, at least LOAD_CHUNK_MIN_GRAINS of them
};

struct Ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    Launches launches;
    std::string err;
    hera_filter_params fp{HERA_MODE_AVG, 0, 0.0f, UINT64_MAX, 1.0f};
    // measurement helpers / knobs
    cudaEvent_t timer0 = nullptr, timer1 = nullptr;
    int opt_flush_l2 = 0;
    int opt_walk_warps_per_sm = 32;   // resident walk warps per SM shared by the walk kernels in flight (8 CTAs of 4 warps at <= 64 registers)
    int opt_parse_variant = 0;     // 0 = CTA tiles (paf_parse_kernel), 1 = warp-autonomous sub-tiles (paf_parse_warp_kernel)
    int opt_walk_variant = 0;      // Monte-Carlo kernel variant (chunks in flight x resident CTAs), see walks.cu
    int opt_walk_cap_small = 16384;  // nodes a walk may visit before its job is re-run with scratch for N nodes (tests lower it)
    int opt_walk_smem_bitmap = 1;  // 0 = keep the walks' visited bitmaps in global memory whatever the graph size (tests)
    uint64_t walk_pool_hint[2][4] = {};   // per (kind, metric): largest path pool a walk call needed so far (avoids the overflow re-run)
    uint64_t walk_work_hint[2][4] = {};   // per (kind, metric): warp-microseconds the last call took (shares of the resident warps)
    int walk_full_cap[2][4] = {};         // per (kind, metric): a walk outgrew the small per-warp scratch once, size it for N nodes
    std::vector<cudaStream_t> aux_streams;   // hera_paths_build: one stream per walk kernel in flight
    cudaStream_t copy_stream = nullptr;      // hera_graph_load_bytes: host-to-device copies of the chunks of a text
    int64_t opt_load_chunk_bytes = 256ll << 20;   // chunk of a host text (rounded to whole tiles; 0 = one copy, then parse)
    DBuf<uint8_t> l2_flush_buf;
    std::unique_ptr<Exchange> xch;   // null = single rank

    std::vector<std::unique_ptr<LoadedText>> texts;
    std::vector<std::unique_ptr<LoadedText>> text_pool;   // buffers of texts dropped by hera_graph_reset, reused by the next loads
    KeptTable kept;                   // LOCAL rows: the kept overlaps this rank parsed itself (all of them on a single rank)
    int opt_verify_ctas_per_sm = 4;   // resident CTAs per SM of the name verification (side stream beside the CSR build)
    int opt_radix_digit_bits = 0;   // 0 = chosen from the key width; 8..11 forces the digit width of radix_sort_pairs (tests)
    int64_t opt_parse_count_rows = 64ll << 20;   // loads whose worst-case row count (bytes / 24) exceeds this size their rows by counting tokens / by what the context already holds
    bool parse_force_count = false;   // the optimistic row reservation of a load overflowed: count the tokens first this time
    bool cache_loaded = false;        // kept overlaps + node table came from hera_graph_load_cache: finalize skips K3
    std::vector<RowRange> own_rows;   // per load: the GLOBAL rows this rank parsed itself
    std::vector<LoadShape> load_shapes;   // per load: every rank's kept count
    uint64_t rows_total = 0;          // kept overlaps of all ranks (== kept.n on a single rank)
    struct GatheredRows {             // sharded runs, built on demand by rows_view(): the complete file-order table on this rank
        bool valid = false;
        DBuf<uint32_t> qs, qe, ts, te, qlen, tlen, rec, e_q, e_t;
        DBuf<float> os, si, es;
        DBuf<uint8_t> flags;
    } gathered;
    GraphTables g;

    PathStore all, uniq;
    bool uniq_ready = false;
    std::vector<uint32_t> call_kind, call_metric;
    std::vector<uint64_t> call_found;
    std::vector<uint32_t> mc_draws;  // host copy, canonical walk order

    // scratch
    DBuf<uint8_t> scratch_a, scratch_b, scratch_c, scratch_d;
    DBuf<uint64_t> counters;     // small device counter block
    HBuf<uint64_t> h_counters;
    HBuf<unsigned long long> walk_h_ctr;   // pinned result words of the walk jobs in flight (4 per job)

    // export mirrors (pinned)
    struct GraphMirror {
        HBuf<uint32_t> node_len, e_q, e_t, e_qs, e_qe, e_ts, e_te, adj_rec, rank[4], canon;
        HBuf<uint8_t> node_anchor, e_strand, adj_side;
        HBuf<uint64_t> node_name_off, adj_off;
        HBuf<char> node_names;
        HBuf<float> e_os, e_si, e_es;
    } gm;
    struct PathMirror {
        HBuf<uint64_t> off;
        HBuf<uint32_t> nodes, slots;
        HBuf<int64_t> len;
    } pm;
    struct GroupMirror {
        std::vector<uint32_t> pair_a1, pair_a2, pair_sorted, grp_pair, grp_size0, grp_pig;
        std::vector<uint64_t> pair_off, grp_pig_off;
        std::vector<int64_t> grp_consensus, pair_consensus;
        std::vector<int32_t> grp_valid;
    } grm;
    HBuf<uint32_t> grp_stage;    // pinned staging of hera_groups' device results
    bool groups_valid = false;   // grm describes the current paths
    struct ScaffoldMirror {
        std::vector<uint32_t> nodes, slots, logged;
        HBuf<char> fasta;
    } scm;
};

// ---- host-side helpers implemented in primitives.cu -----------------------------------------
// out[i] = sum of in[0..i), all on device; returns nothing, *d_total (device, may be null) gets the sum.
void exclusive_scan_u32(Ctx &c, const uint32_t *in, uint64_t n, uint32_t *out, uint32_t *d_total);
void exclusive_scan_u8_to_u32(Ctx &c, const uint8_t *in, uint64_t n, uint32_t *out, uint32_t *d_total);
void exclusive_scan_u32_to_u64(Ctx &c, const uint32_t *in, uint64_t n, uint64_t *out, uint64_t *d_total);
// Stable sort of (key, value) by the low `bits` bits of key.  Result ends up in keys/vals (the
// alt buffers are scratch of the same size).
void radix_sort_pairs(Ctx &c, uint32_t *keys, uint32_t *vals, uint32_t *keys_alt, uint32_t *vals_alt, uint64_t n, int bits);

// Writes a buffer larger than L2 when the flush_l2 option is on (call before an event-timed region).
void maybe_flush_l2(Ctx &c);

// ---- stages ------------------------------------------------------------------------------
// host != nullptr: the bytes are still in host memory; they are copied in chunks on the context's copy stream and every
// chunk is parsed as it lands (whole texts only; see chunk_carry_kernel).
void load_text_resident(Ctx &c, LoadedText &text, uint32_t file_id, bool anchors, uint32_t test_load_num, const ShardSpec *shard, hera_load_result *out,
                        const HostSource *host = nullptr);
void finalize_graph(Ctx &c, uint32_t metrics_mask, hera_graph_info *out);
RowPieces row_pieces(const Ctx &c);
RowView rows_view(Ctx &c);   // COLLECTIVE on a sharded context the first time after a finalize (all-gather of the rows)
void export_graph(Ctx &c, hera_graph_view *out);
void save_graph_cache(Ctx &c, const char *path);
void load_graph_cache(Ctx &c, const char *path);
void walk_monte_carlo(Ctx &c, int metric, const hera_path_params &pp, uint64_t seed, hera_walk_stats *out);
void walk_deterministic(Ctx &c, int metric, hera_walk_stats *out);
void probe_list_gather(Ctx &c, uint32_t warps_per_sm, uint32_t hops, double *gbytes_per_s, float *ms_out);
void walk_build(Ctx &c, int n, const int *kinds, const int *metrics, const hera_path_params &pp, uint64_t seed, hera_walk_stats *out);
void unique_paths(Ctx &c, uint64_t *n_unique);
void export_paths(Ctx &c, hera_paths_view *out);
void group_paths(Ctx &c, const hera_path_params &pp, hera_groups_view *out);
void build_scaffold(Ctx &c, const char *reads_path, const char *contigs_path, uint64_t min_path_num, hera_scaffold_view *out);

}  // namespace hera
