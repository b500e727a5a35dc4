/*
 * hera_b200.h -- C ABI of the B200-native HERA repeat-resolution hot path.
 *
 * The reference (lirfu/Telomeri) has no plugin/FFI layer: its hot path is reached only through
 * the C++ classes main() calls (src/Main.cpp:211-352).  This header is the boundary a binding
 * would use instead; each entry point names the reference member function(s) it replaces.
 * Plain pointers and sizes only.  Every function returns 0 on success; on failure it returns a
 * non-zero code and hera_last_error() describes it.  There is no CPU fallback: without a CUDA
 * device hera_create() fails.
 *
 * One host thread per context.  "View" structs hold HOST pointers owned by the context
 * (pinned memory), valid until the next call that mutates the same object or hera_destroy().
 */
#ifndef HERA_B200_H
#define HERA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct hera_ctx hera_ctx;

/* OverlapGraph::LengthCalculationMode, include/OverlapGraph.hpp:56-65 */
enum { HERA_MODE_MIN = 0, HERA_MODE_MAX = 1, HERA_MODE_AVG = 2, HERA_MODE_SUM = 3 };
/* Utils::Metrics, include/Utils.hpp:43-45 */
enum { HERA_EXTENSION_SCORE = 0, HERA_OVERLAP_SCORE = 1, HERA_EXTENSION_SCORE_SQRT = 2, HERA_OVERLAP_SCORE_SQRT = 3 };

/* OverlapGraph::FilterParameters, include/OverlapGraph.hpp:67-78 (defaults src/Main.cpp:56-62) */
typedef struct {
    int32_t mode;
    uint64_t min_overlap_length;
    float min_overlap_percentage;
    uint64_t max_overhang_length;
    float max_overhang_percentage;
} hera_filter_params;

/* PathManager::Parameters, include/PathManager.hpp:27-40 (defaults src/Main.cpp:63-69) */
typedef struct {
    int32_t rebuild_attempts;
    int32_t backtrack_attempts;
    int64_t len_threshold;
    uint64_t window_size;
    float ratio_threshold;
} hera_path_params;

typedef struct {
    uint64_t records;      /* "ctr": records consumed from this input (src/OverlapGraph.cpp:128) */
    uint64_t kept;         /* overlaps of this input that passed the filter */
    uint64_t kept_total;   /* cumulative, what "Loaded a/b edges" prints (src/OverlapGraph.cpp:150) */
    uint64_t bytes;
    float parse_ms;        /* device time of the tokenizer/filter/score kernels (CUDA events) */
    float h2d_ms;          /* host->device copy time, 0 for hera_graph_load_device */
    float exchange_ms;     /* sharded loads: device time of the all-gather of the kept overlaps (0 otherwise) */
} hera_load_result;

typedef struct {
    uint64_t nodes, anchors, edges, slots;   /* slots = 2 * edges adjacency entries */
    float intern_ms, csr_ms, rank_ms;
} hera_graph_info;

/* Host mirror of the graph.  Edge-like tables are structure-of-arrays; index e = kept overlap
 * in file order (OverlapGraph::edges_), index s = adjacency slot (nodes_[x].edges[k] has
 * s = adj_off[x] + k).  For slots the fields are as the owner's list stores them after the
 * q/t swap of src/OverlapGraph.cpp:245-248: q_* belong to the neighbour, t_* to the owner. */
typedef struct {
    uint64_t nodes, edges, slots;
    const uint32_t *node_len;
    const uint8_t *node_anchor;
    const uint64_t *node_name_off;   /* nodes + 1 */
    const char *node_names;
    const uint32_t *e_q, *e_t, *e_qs, *e_qe, *e_ts, *e_te;
    const float *e_os, *e_si, *e_es;
    const uint8_t *e_strand;
    const uint64_t *adj_off;         /* nodes + 1 */
    const uint32_t *adj_rec;         /* slot -> kept overlap */
    const uint8_t *adj_side;         /* 0 = owner is the PAF target, 1 = owner is the PAF query (swapped copy) */
    const uint32_t *rank_es, *rank_os, *rank_es_sqrt, *rank_os_sqrt;  /* per-node stable descending order, absolute slots; NULL if not built */
    const uint32_t *canon;           /* first slot of the owner with the same neighbour */
} hera_graph_view;

/* Paths in the order of PathManager::paths_.  Path i owns nodes [off[i], off[i+1]) and slots
 * [off[i] - i, off[i+1] - (i+1)). */
typedef struct {
    uint64_t paths, total_nodes;
    const uint64_t *off;
    const uint32_t *nodes;
    const uint32_t *slots;
    const int64_t *len;              /* Path::length(), src/Path.cpp:6-16 */
    uint64_t calls;                  /* build calls so far */
    const uint32_t *call_kind;       /* 0 = Monte Carlo, 1 = deterministic */
    const uint32_t *call_metric;
    const uint64_t *call_found;      /* "Found N paths." per call */
    uint64_t mc_walks;
    const uint32_t *mc_draws_per_walk;  /* random draws each Monte-Carlo walk consumed, canonical walk order */
} hera_paths_view;

typedef struct {
    uint64_t pairs, groups, sorted_total, pig_total;
    const uint32_t *pair_a1, *pair_a2;   /* anchor node ids, map order of getPathsBetweenAnchors */
    const uint64_t *pair_off;            /* pairs + 1, into pair_sorted */
    const uint32_t *pair_sorted;         /* unique-path ids per pair in the order constructGroups leaves them */
    const uint32_t *grp_pair;
    const uint32_t *grp_size0;           /* members before discardNotFrequent: consecutive ranges of pair_sorted */
    const uint64_t *grp_pig_off;         /* groups + 1 */
    const uint32_t *grp_pig;             /* members after discardNotFrequent */
    const int64_t *grp_consensus;        /* path id or -1 */
    const int32_t *grp_valid;            /* valid_path_number */
    const int64_t *pair_consensus;       /* inter-group pick (src/Main.cpp:306-332), path id or -1 */
    float ms;
} hera_groups_view;

/* Timing of the last walk call, for bench.py. */
typedef struct {
    uint64_t walks;        /* walks started by the call */
    uint64_t found;
    uint64_t steps;        /* adjacency entries examined (sum over walks) */
    float ms;
    uint32_t launches;     /* kernels launched by the call */
} hera_walk_stats;

/* ---- lifetime ---------------------------------------------------------------------------- */
hera_ctx *hera_create(int device);               /* NULL if the device cannot be used */
void hera_destroy(hera_ctx *ctx);
const char *hera_last_error(hera_ctx *ctx);      /* ctx may be NULL: error of the last failed hera_create */
uint64_t hera_kernel_launches(hera_ctx *ctx);    /* kernels launched by this context so far */
/* Measurement helpers (no reference counterpart; the reference only has Stopwatch, src/Stopwatch.cpp):
 * a CUDA-event pair on the context's own stream, and integer knobs.
 * Options: "flush_l2" (1 = write a buffer larger than L2 before every event-timed kernel region,
 * outside the timed region), "walk_warps_per_sm" (resident walk warps per SM shared by the walk kernels in flight, default 32), "walk_smem_bitmap" (0 = visited
 * bitmaps of the Monte-Carlo walks stay in global memory even when they would fit in shared memory),
 * "walk_variant" (chunks in flight x resident CTAs of the Monte-Carlo kernel, tuning), "parse_count_rows" (a load whose worst case, bytes / 24 rows, exceeds this many rows does not reserve that worst case: it
 * counts the tokens first or, on a context that already holds enough rows, tries with those and repeats the load if they
 * do not suffice; default 64 Mi), "walk_cap_small" (nodes a walk may
 * visit before its call is repeated with per-warp scratch for every node of the graph, default 16384), "parse_variant" (0 = the
 * CTA-tile tokenizer, default; 1 = the warp-autonomous tokenizer -- same results, see DESIGN.md K1+K2),
 * "load_chunk_bytes" (hera_graph_load_bytes copies and parses a host text in chunks of this size, rounded to whole
 * 30 720-byte tiles, at least 8 of them; 0 = one copy, then one launch; default 256 Mi). */
int hera_timer_start(hera_ctx *ctx);
int hera_timer_stop(hera_ctx *ctx, float *ms);   /* synchronises the stream */
int hera_set_option(hera_ctx *ctx, const char *key, int64_t value);
/* Measured ceiling for the walks on the finalized graph: warps_per_sm resident warps per SM each make `hops`
 * dependent hops, reading the whole adjacency list of the current node (16-byte entries) and moving on to a
 * neighbour found there -- the memory behaviour of a walk step (src/PathManager.cpp:52-113) without roulette,
 * visited set or random numbers.  Reports list bytes per second. */
int hera_probe_list_gather(hera_ctx *ctx, uint32_t warps_per_sm, uint32_t hops, double *gbytes_per_s, float *ms);

/* Parity evidence without moving the tables: 64-bit digests of the device-resident results, computed on the
 * device (no reference counterpart).  digest(array) = sum over its 32-bit words w[i] of mix(i, w[i]) mod 2^64 with
 *   x = (i + 1) * 0x9E3779B97F4A7C15 ^ (w[i] + 1) * 0xC2B2AE3D27D4EB4F;  x ^= x >> 32;  x *= 0xD6E8FEB86659FD93;  x ^= x >> 32
 * (bytes of a uint8 array count as words; 64-bit elements as two words, low first).  out[] gets HERA_DIGEST_WORDS
 * values in this order: node_len, node_anchor, e_q, e_t, adj_off, adj_rec*2+adj_side, walk view {neighbour | anchor
 * bit 31, es, os, q_start - t_start} (4 words per slot), canon, rank_es, rank_os, rank_es_sqrt, rank_os_sqrt (0 when
 * not built), then the paths hera_paths_export would return (unique ones after hera_paths_unique): off, nodes, slots,
 * len, then the tables of the last hera_groups call: pair_a1, pair_a2, pair_sorted, grp_size0, grp_pig, grp_consensus,
 * grp_valid, pair_consensus (0 before hera_groups).  bench.py compares the digests of all ranks with each other and
 * with the digests of a run that was compared with the oracle table by table (tests/golden). */
#define HERA_DIGEST_WORDS 24
int hera_digest(hera_ctx *ctx, uint64_t *out);

/* ---- sharded runs: one rank per GPU ---------------------------------------------
 * No reference counterpart (the reference is single-threaded, src/Main.cpp).  A communicator is
 * attached to a context once; afterwards hera_graph_load_file / hera_graph_load_shard,
 * hera_graph_finalize, hera_paths_monte_carlo and hera_paths_deterministic are COLLECTIVE: every
 * rank must call them with the same arguments.  The PAF text is parsed by byte range (each rank
 * tokenizes, filters and scores its slice), node ids and adjacency lists are built from every rank's own
 * overlaps (the distinct names of the ranks are merged; every rank writes its part of each node's list
 * straight into every rank's tables over NVLink), walks are dealt out by walk id and the accepted paths
 * all-gathered in the reference's path order, so every rank ends up with the same graph and the same
 * paths_, bit for bit what one GPU produces.
 *   NCCL transport: one process per GPU.  Rank 0 calls hera_nccl_unique_id, the launcher hands the
 *   128 bytes to every rank (torch.distributed broadcast, MPI_Bcast, a file), each rank calls
 *   hera_comm_init_nccl.  libnccl.so.2 is loaded on demand.
 *   Local transport: several contexts of one process, one host thread per rank
 *   (hera_comm_local_create once, hera_comm_attach_local per context). */
typedef struct hera_local_gang hera_local_gang;
int hera_nccl_unique_id(void *out128, char *err, uint64_t err_cap);      /* 0 = ok; err gets a message otherwise */
int hera_comm_init_nccl(hera_ctx *ctx, int rank, int nranks, const void *id128);
hera_local_gang *hera_comm_local_create(int nranks);
void hera_comm_local_destroy(hera_local_gang *gang);
int hera_comm_attach_local(hera_ctx *ctx, hera_local_gang *gang, int rank);
int hera_comm_rank(hera_ctx *ctx);   /* 0 without a communicator */
int hera_comm_size(hera_ctx *ctx);   /* 1 without a communicator */
/* Waits until every rank's stream is idle and every rank has made this call (no-op without a communicator).
 * bench.py aligns the ranks with it before a step's timed stages, so that a stage time does not contain the
 * wait for a rank whose host thread was late. */
int hera_comm_barrier(hera_ctx *ctx);
/* The slice of a `file_bytes` long PAF file rank `rank` of `nranks` must hand to hera_graph_load_shard:
 * it OWNS the tokens starting in [*begin, *end) and must also provide the bytes up to *avail_end
 * (so that a record that starts in the slice can be finished) and the byte before *begin. */
void hera_shard_range(uint64_t file_bytes, int rank, int nranks, uint64_t *begin, uint64_t *end, uint64_t *avail_end);
/* OverlapGraph::load over a byte-range slice (collective).  text = bytes [begin, avail_end) of the
 * file in host (on_device = 0) or device memory (1 = copied into the library's own buffer, 2 = parsed in
 * place under the contract of hera_graph_load_device_view); owned = end - begin, avail = avail_end - begin;
 * prev_byte = the file byte before `begin`, or -1 at the start of the file; at_eof = avail_end is the
 * end of the file.  Results (records, kept, kept_total) are global.  Returns 20 (and changes nothing)
 * when some slice holds a record formatted extraction would fail on (non-numeric field, strand of
 * two characters, more than 16 columns ...): reload that file unsharded with
 * hera_graph_load_bytes / hera_graph_load_device on every rank -- hera_graph_load_file does so itself. */
int hera_graph_load_shard(hera_ctx *ctx, const void *text, int on_device, uint64_t owned, uint64_t avail, int prev_byte, int at_eof,
                          int anchors, uint32_t test_load_num, hera_load_result *out);

/* ---- OverlapGraph ------------------------------------------------------------------------ */
/* graph.filter_params_ = ... (src/Main.cpp:213-217) */
int hera_graph_set_filter(hera_ctx *ctx, const hera_filter_params *fp);
/* OverlapGraph::load(filepath, anchors) (src/OverlapGraph.cpp:97-154): reads the file, copies it
 * to the device, tokenizes / filters / scores there, appends to the graph.  test_load_num as
 * OverlapGraph::test_load_num_ (0 = whole file).  A file that cannot be opened returns 1. */
int hera_graph_load_file(hera_ctx *ctx, const char *path, int anchors, uint32_t test_load_num, hera_load_result *out);
/* Same, PAF text already in host memory.  The text is copied to the device in chunks ("load_chunk_bytes", default
 * 256 MiB) on a copy stream of the context and every chunk is parsed by its own launch as soon as it has landed: the
 * tokenizer hides behind the copy (out->h2d_ms = the copies; out->parse_ms = first launch to last, waits for the
 * copy included).  Pinned host memory makes the copies asynchronous; pageable memory works, without the overlap. ... */
int hera_graph_load_bytes(hera_ctx *ctx, const void *host_text, uint64_t n, int anchors, uint32_t test_load_num, hera_load_result *out);
/* ... or already resident in device memory (n bytes at a cudaMalloc'ed address; the library
 * copies it device-to-device into its own padded buffer). */
int hera_graph_load_device(hera_ctx *ctx, const void *device_text, uint64_t n, int anchors, uint32_t test_load_num, hera_load_result *out);
/* Same, parsed IN PLACE: no copy is made.  device_text must be 16-byte aligned and lie in a cudaMalloc'ed
 * allocation (the 16-byte chunk that holds byte n-1 is read whole, nothing behind it); the caller keeps the
 * buffer alive and unchanged until hera_graph_reset / hera_destroy -- node names are read from it by
 * hera_graph_finalize and hera_graph_save.  What a caller whose overlapper already left its PAF output in
 * device memory uses; results are identical to hera_graph_load_device. */
int hera_graph_load_device_view(hera_ctx *ctx, const void *device_text, uint64_t n, int anchors, uint32_t test_load_num, hera_load_result *out);
/* Drops all loaded overlaps (a fresh OverlapGraph). */
int hera_graph_reset(hera_ctx *ctx);
/* Node numbering (OverlapGraph::nodeIndex + node creation, src/OverlapGraph.cpp:208-218,251-258),
 * adjacency lists (src/OverlapGraph.cpp:241-248) as CSR, per-metric rank permutations and
 * canonical slots used by the greedy walks.  metrics_mask: bit m set = build ranks for metric m.
 * May be called again after further loads. */
int hera_graph_finalize(hera_ctx *ctx, uint32_t metrics_mask, hera_graph_info *out);
/* Binary cache of the parsed graph (no reference counterpart): what loading, filtering,
 * scoring and node numbering produced -- kept overlaps in file order with node ids, node table with
 * names, the filter used.  hera_graph_save after hera_graph_finalize; hera_graph_load_cache on a fresh
 * graph restores it, then hera_graph_finalize builds CSR and ranks without reading any PAF text. */
int hera_graph_save(hera_ctx *ctx, const char *path);
int hera_graph_load_cache(hera_ctx *ctx, const char *path);
/* nodes_ / edges_ / nodes_[i].edges as flat host arrays.  On a sharded context COLLECTIVE the first time after a
 * hera_graph_finalize: the kept overlaps stay on the rank that parsed them until somebody needs the whole edges_
 * table (this call, hera_graph_save, hera_scaffold, hera_digest), then every rank gathers them.  A rank that makes
 * none of these calls while another rank does must call hera_graph_gather_rows instead. */
int hera_graph_export(hera_ctx *ctx, hera_graph_view *out);
int hera_graph_gather_rows(hera_ctx *ctx);

/* ---- allocation-free variant of the tokenizer ------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * The tokenizer / filter / score kernel of hera_graph_load_* over a PAF text resident in device memory, with its
 * scratch (counter block + the two look-back chains, one 64-bit word per 30 KiB tile and chain) in a file-scope
 * device array instead of the context's growable buffer: the call makes NO cudaMalloc / cudaFree / cudaMallocAsync
 * and needs no hera_ctx.  Limit: n <= HERA_NOALLOC_MAX_TEXT_BYTES (480 MiB of text; the array is ~256 KB and every
 * GPU that loads the library holds it until the process exits); longer texts are rejected with code 60.
 * The caller passes everything else: device_text (16-byte aligned, in a cudaMalloc'ed allocation; the 16-byte chunk
 * holding byte n-1 is read whole, nothing behind it), the 15 output columns in device memory with room for
 * `capacity` >= n / 24 + 2 rows (code 61 otherwise; rows are written in file order: what edges_ holds before node
 * ids exist -- qname/tname are {file:4 = 0, byte offset:44, length:16} references into the text, qhash/thash the
 * 64-bit name hashes, flags bit 0 = strand '-', bit 1 = anchors load, bit 2 = query name starts with "ctg"),
 * d_result = 4 words in DEVICE memory: records consumed, kept overlaps, index of the first record formatted
 * extraction would fail or desynchronise on (UINT64_MAX = none; otherwise only rows with rec below it are
 * meaningful: load such a text with hera_graph_load_*), tokens; and the CUDA stream (cudaStream_t, may be NULL)
 * everything is queued on -- nothing is synchronised, results are valid once the stream has passed the call.
 * All calls on a GPU SHARE the scratch array, which keeps the previous call's contents: each call first clears
 * the part it uses, and two calls whose kernels can overlap on the GPU (different streams without a
 * dependency) race -- order them.  On failure returns a non-zero code and writes a message to err (may be NULL). */
#define HERA_NOALLOC_MAX_TEXT_BYTES (16384ull * 30720ull - 30720ull)
typedef struct {
    uint32_t *qs, *qe, *ts, *te, *qlen, *tlen, *rec;
    float *os, *si, *es;
    uint8_t *flags;
    uint64_t *qhash, *thash, *qname, *tname;
} hera_kept_columns;
int hera_paf_parse_noalloc(int device, const void *device_text, uint64_t n, int anchors, const hera_filter_params *fp, uint32_t test_load_num,
                           const hera_kept_columns *out, uint64_t capacity, uint64_t *d_result, void *stream, char *err, uint64_t err_cap);

/* ---- PathManager ------------------------------------------------------------------------- */
/* PathManager::buildMonteCarlo (src/PathManager.cpp:12-177).  Draws come from Philox4x32-10 with
 * counter (anchor node, attempt, draw number, metric) and key = seed instead of one shared
 * std::mt19937 stream, so walks are independent of execution order. */
int hera_paths_monte_carlo(hera_ctx *ctx, int metric, const hera_path_params *pp, uint64_t seed, hera_walk_stats *out);
/* PathManager::buildDeterministic (src/PathManager.cpp:179-430). */
int hera_paths_deterministic(hera_ctx *ctx, int metric, hera_walk_stats *out);
/* The n build calls of main() (src/Main.cpp:237-243) as ONE call: kinds[i] = 0 buildMonteCarlo /
 * 1 buildDeterministic with metrics[i].  The walk kernels of all calls run concurrently (one stream
 * each), which hides the tail every walk kernel has; paths_ ends up exactly as the n separate calls in
 * this order would leave it.  out (may be NULL) gets n entries.  n <= 8. */
int hera_paths_build(hera_ctx *ctx, int n, const int *kinds, const int *metrics, const hera_path_params *pp, uint64_t seed, hera_walk_stats *out);
/* PathManager::filterUnique (src/PathManager.cpp:432-438). */
int hera_paths_unique(hera_ctx *ctx, uint64_t *n_unique);
/* Forget all paths (a fresh PathManager). */
int hera_paths_reset(hera_ctx *ctx);
int hera_paths_export(hera_ctx *ctx, hera_paths_view *out);

/* getPathsBetweenAnchors + constructGroups + PathWindow + PathGroup::{discardNotFrequent,
 * calculateConsensusPath, calculateValidPathNumber} + the inter-group pick
 * (src/PathManager.cpp:508-645, src/PathWindow.cpp, src/PathGroup.cpp, src/Main.cpp:277-342). */
int hera_groups(hera_ctx *ctx, const hera_path_params *pp, hera_groups_view *out);

/* The scaffold: chain of the anchor pairs' consensus paths (PathManager::constructConsensusPath,
 * src/PathManager.cpp:664-767, as main() calls it at src/Main.cpp:344-353; min_path_num = 10 there) and
 * Scaffolder (src/Scaffolder.cpp:12-147): reads / contigs FASTA (two lines per record; four for a path
 * ending in 'q'), name lookup, substring / complement / reverse stitching.  The FASTA files are indexed
 * and the sequence is assembled on the GPU.  `fasta` is the exact content of HERA_result.fasta.
 * Needs hera_groups first.  Returns 40 ("No consensus found!") when no anchor pair qualifies. */
typedef struct {
    uint64_t nodes;                 /* scaffold path: nodes, and nodes - 1 slots */
    const uint32_t *node, *slot;
    uint64_t logged;                /* positions in node[] of the nodes Scaffolder::write prints ("a-b-c") */
    const uint32_t *logged_index;
    uint64_t fasta_bytes, sequence_length;
    const char *fasta;
    float ms;                       /* FASTA load + index + lookup + stitch, CUDA events */
} hera_scaffold_view;
int hera_scaffold(hera_ctx *ctx, const char *reads_fasta, const char *contigs_fasta, uint64_t min_path_num, hera_scaffold_view *out);

#ifdef __cplusplus
}
#endif
#endif
