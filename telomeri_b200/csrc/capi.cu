// extern "C" entry points of libhera_b200.so (declared in include/hera_b200.h).
#include <algorithm>
#include <cstring>
#include <ctime>
#include <cstdio>
#include <cstdlib>
#include <fstream>

#include "ctx.cuh"

namespace hera {
uint64_t padded_text_size(uint64_t n);  // paf_parse.cu
void paf_parse_noalloc(int device, const void *device_text, uint64_t n, int anchors, const hera_filter_params &fp, uint32_t test_load_num,
                       const hera_kept_columns &out, uint64_t capacity, uint64_t *d_result, cudaStream_t stream);   // paf_parse.cu

__global__ void l2_flush_kernel(uint4 *p, uint64_t n) {
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<uint64_t>(gridDim.x) * blockDim.x) p[i] = make_uint4(i, 0, 0, 0);
}

void maybe_flush_l2(Ctx &c) {
    if (!c.opt_flush_l2) return;
    const uint64_t bytes = 512ull << 20;
    c.l2_flush_buf.reserve(bytes);
    l2_flush_kernel<<<1184, 256, 0, c.stream>>>(reinterpret_cast<uint4 *>(c.l2_flush_buf.p), bytes / 16);
    HERA_CUDA(cudaGetLastError());
}
static std::string g_create_error;

// ---- hera_digest --------------------------------------------------------------------------------
__host__ __device__ inline unsigned long long digest_mix(unsigned long long i, unsigned long long w) {
    unsigned long long x = ((i + 1) * 0x9E3779B97F4A7C15ull) ^ ((w + 1) * 0xC2B2AE3D27D4EB4Full);
    x ^= x >> 32;
    x *= 0xD6E8FEB86659FD93ull;
    x ^= x >> 32;
    return x;
}

template <typename T>
__global__ void __launch_bounds__(256) digest_kernel(const T *__restrict__ a, uint64_t n, unsigned long long *out) {
    unsigned long long s = 0;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<uint64_t>(gridDim.x) * blockDim.x)
        s += digest_mix(i, static_cast<unsigned long long>(a[i]));
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    if ((threadIdx.x & 31u) == 0 && s) atomicAdd(out, s);
}

template <typename T>
static void digest_dev(Ctx &c, const T *a, uint64_t n, unsigned long long *out) {
    if (!n) return;
    const unsigned grid = std::min<uint64_t>(div_up(n, 256), 148 * 8);
    HERA_LAUNCH(c, digest_kernel<T>, grid, 256, 0, a, n, out);
}

template <typename T>
static uint64_t digest_host_words(const std::vector<T> &v) {
    const uint32_t *w = reinterpret_cast<const uint32_t *>(v.data());
    const uint64_t n = v.size() * sizeof(T) / 4;
    unsigned long long s = 0;
    for (uint64_t i = 0; i < n; i++) s += digest_mix(i, w[i]);
    return s;
}

static void digest_tables(Ctx &c, uint64_t *out) {
    for (int i = 0; i < HERA_DIGEST_WORDS; i++) out[i] = 0;
    TempPool tmp(c.stream);
    unsigned long long *d = tmp.zeroed<unsigned long long>(HERA_DIGEST_WORDS);
    const GraphTables &g = c.g;
    if (g.ready && g.E) {
        const RowView rows = rows_view(c);   // (collective on a sharded context)
        digest_dev(c, g.node_len.p, g.N, d + 0);
        digest_dev(c, g.node_anchor.p, g.N, d + 1);
        digest_dev(c, rows.e_q, g.E, d + 2);
        digest_dev(c, rows.e_t, g.E, d + 3);
        digest_dev(c, g.adj_off.p, g.N + 1, d + 4);
        digest_dev(c, g.adj_pay.p, g.S, d + 5);
        digest_dev(c, reinterpret_cast<const uint32_t *>(g.adj4.p), 4 * g.S, d + 6);
        digest_dev(c, g.canon.p, g.S, d + 7);
        for (int m = 0; m < 4; m++) if (g.rank_ready[m]) digest_dev(c, g.rank[m].p, g.S, d + 8 + m);
    }
    const PathStore &ps = c.uniq_ready ? c.uniq : c.all;
    if (ps.paths) {
        digest_dev(c, reinterpret_cast<const uint32_t *>(ps.off.p), 2 * (ps.paths + 1), d + 12);
        digest_dev(c, ps.node.p, ps.nodes, d + 13);
        digest_dev(c, ps.slot.p, ps.nodes - ps.paths, d + 14);
        digest_dev(c, reinterpret_cast<const uint32_t *>(ps.len.p), 2 * ps.paths, d + 15);
    }
    HERA_CUDA(cudaMemcpyAsync(out, d, HERA_DIGEST_WORDS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c.stream));
    HERA_CUDA(cudaStreamSynchronize(c.stream));
    if (c.groups_valid) {   // the group tables are marshalled on the host by hera_groups
        const auto &m = c.grm;
        out[16] = digest_host_words(m.pair_a1); out[17] = digest_host_words(m.pair_a2); out[18] = digest_host_words(m.pair_sorted);
        out[19] = digest_host_words(m.grp_size0); out[20] = digest_host_words(m.grp_pig); out[21] = digest_host_words(m.grp_consensus);
        out[22] = digest_host_words(m.grp_valid); out[23] = digest_host_words(m.pair_consensus);
    }
}
}  // namespace hera

struct hera_ctx {
    hera::Ctx c;
};

struct hera_local_gang {
    std::shared_ptr<hera::LocalHub> hub;
};

using hera::Ctx;

#define HERA_API_BEGIN(ctxp)                                   \
    if (!(ctxp)) return 100;                                   \
    Ctx &c = (ctxp)->c;                                        \
    try {                                                      \
        if (cudaSetDevice(c.device) != cudaSuccess) throw hera::Error{2, "cudaSetDevice failed"};

#define HERA_API_END                                           \
        return 0;                                              \
    } catch (const hera::Error &e) {                           \
        c.err = e.msg;                                         \
        cudaGetLastError();                                    \
        if (c.xch && e.code != 20) c.xch->abort();             \
        return e.code;                                         \
    } catch (const std::exception &e) {                        \
        c.err = e.what();                                      \
        if (c.xch) c.xch->abort();                             \
        return 99;                                             \
    }

extern "C" {

hera_ctx *hera_create(int device) {
    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);
    if (err != cudaSuccess || device < 0 || device >= count) {
        hera::g_create_error = err != cudaSuccess ? std::string("no usable CUDA device: ") + cudaGetErrorString(err)
                                                  : "device index out of range (" + std::to_string(count) + " devices)";
        cudaGetLastError();
        return nullptr;
    }
    if (cudaSetDevice(device) != cudaSuccess) { hera::g_create_error = "cudaSetDevice failed"; return nullptr; }
    hera_ctx *h = new hera_ctx();
    h->c.device = device;
    // (Tried: highest priority for this stream so that the name verification on its side stream only takes the SM slots the
    // build leaves free.  Interning went from 5.1 to 2.3 ms on configs[3], but the starved verification then ran at the end,
    // on top of the rank kernels: graph build 34.0 -> 39.4 ms.  Equal priorities stay.)
    if (cudaStreamCreateWithFlags(&h->c.stream, cudaStreamNonBlocking) != cudaSuccess) {
        hera::g_create_error = "cudaStreamCreate failed";
        delete h;
        return nullptr;
    }
    // keep freed stream-ordered allocations cached instead of returning them to the driver
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        uint64_t threshold = UINT64_MAX;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
    }
    if (const char *v = getenv("HERA_PARSE_VARIANT")) h->c.opt_parse_variant = std::atoi(v);   // tuning experiments: same as hera_set_option("parse_variant")
    return h;
}

void hera_destroy(hera_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->c.device);
    cudaStreamSynchronize(ctx->c.stream);
    cudaStream_t s = ctx->c.stream;
    for (cudaStream_t a : ctx->c.aux_streams) { cudaStreamSynchronize(a); cudaStreamDestroy(a); }
    if (ctx->c.copy_stream) { cudaStreamSynchronize(ctx->c.copy_stream); cudaStreamDestroy(ctx->c.copy_stream); }
    delete ctx;
    cudaStreamDestroy(s);
}

const char *hera_last_error(hera_ctx *ctx) { return ctx ? ctx->c.err.c_str() : hera::g_create_error.c_str(); }

uint64_t hera_kernel_launches(hera_ctx *ctx) { return ctx ? ctx->c.launches.count : 0; }

int hera_timer_start(hera_ctx *ctx) {
    HERA_API_BEGIN(ctx)
    if (!c.timer0) { HERA_CUDA(cudaEventCreate(&c.timer0)); HERA_CUDA(cudaEventCreate(&c.timer1)); }
    HERA_CUDA(cudaEventRecord(c.timer0, c.stream));
    HERA_API_END
}

int hera_timer_stop(hera_ctx *ctx, float *ms) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(c.timer0 != nullptr && ms != nullptr, 101, "hera_timer_stop without hera_timer_start");
    HERA_CUDA(cudaEventRecord(c.timer1, c.stream));
    HERA_CUDA(cudaEventSynchronize(c.timer1));
    HERA_CUDA(cudaEventElapsedTime(ms, c.timer0, c.timer1));
    HERA_API_END
}

int hera_probe_list_gather(hera_ctx *ctx, uint32_t warps_per_sm, uint32_t hops, double *gbytes_per_s, float *ms) {
    HERA_API_BEGIN(ctx)
    hera::probe_list_gather(c, warps_per_sm, hops, gbytes_per_s, ms);
    HERA_API_END
}

int hera_set_option(hera_ctx *ctx, const char *key, int64_t value) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(key != nullptr, 101, "null option name");
    const std::string k(key);
    if (k == "flush_l2") c.opt_flush_l2 = value != 0;
    else if (k == "walk_warps_per_sm") c.opt_walk_warps_per_sm = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(64, value)));
    else if (k == "walk_variant") c.opt_walk_variant = static_cast<int>(value);
    else if (k == "parse_variant") c.opt_parse_variant = static_cast<int>(value);
    else if (k == "walk_smem_bitmap") c.opt_walk_smem_bitmap = value != 0;
    else if (k == "verify_ctas_per_sm") c.opt_verify_ctas_per_sm = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(8, value)));
    else if (k == "radix_digit_bits") c.opt_radix_digit_bits = static_cast<int>(value);
    else if (k == "parse_count_rows") c.opt_parse_count_rows = std::max<int64_t>(1, value);
    else if (k == "load_chunk_bytes") c.opt_load_chunk_bytes = std::max<int64_t>(0, value);
    else if (k == "walk_hints_reset") { for (auto &a : c.walk_work_hint) for (uint64_t &w : a) w = 0; }
    else if (k == "walk_cap_small") { c.opt_walk_cap_small = static_cast<int>(std::max<int64_t>(2, value)); for (auto &a : c.walk_full_cap) for (int &f : a) f = 0; }
    else throw hera::Error{102, "unknown option: " + k};
    HERA_API_END
}

int hera_digest(hera_ctx *ctx, uint64_t *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(out != nullptr, 101, "null output");
    hera::digest_tables(c, out);
    HERA_API_END
}

int hera_paf_parse_noalloc(int device, const void *device_text, uint64_t n, int anchors, const hera_filter_params *fp, uint32_t test_load_num,
                           const hera_kept_columns *out, uint64_t capacity, uint64_t *d_result, void *stream, char *err, uint64_t err_cap) {
    auto fail = [&](int code, const std::string &msg) {
        if (err && err_cap) { std::strncpy(err, msg.c_str(), err_cap - 1); err[err_cap - 1] = 0; }
        cudaGetLastError();
        return code;
    };
    try {
        if (!fp || !out) throw hera::Error{101, "hera_paf_parse_noalloc: null filter parameters or output columns"};
        hera::paf_parse_noalloc(device, device_text, n, anchors, *fp, test_load_num, *out, capacity, d_result, static_cast<cudaStream_t>(stream));
        return 0;
    } catch (const hera::Error &e) {
        return fail(e.code, e.msg);
    } catch (const std::exception &e) {
        return fail(99, e.what());
    }
}

int hera_graph_set_filter(hera_ctx *ctx, const hera_filter_params *fp) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(fp != nullptr, 101, "null filter parameters");
    c.fp = *fp;
    HERA_API_END
}

// Text buffers are recycled across hera_graph_reset: no cudaMalloc / cudaFree in the steady state.
static std::unique_ptr<hera::LoadedText> acquire_text(Ctx &c, uint64_t n) {
    std::unique_ptr<hera::LoadedText> t;
    const uint64_t need = hera::padded_text_size(n);
    size_t best = c.text_pool.size();
    for (size_t i = 0; i < c.text_pool.size(); i++) {   // best fit
        if (c.text_pool[i]->bytes.cap >= need && (best == c.text_pool.size() || c.text_pool[i]->bytes.cap < c.text_pool[best]->bytes.cap)) best = i;
    }
    if (best < c.text_pool.size()) { t = std::move(c.text_pool[best]); c.text_pool.erase(c.text_pool.begin() + best); }
    if (!t) t.reset(new hera::LoadedText());
    t->n = n;
    t->bytes.reserve(need);
    return t;
}

// Parses a text whose bytes are (being copied) on the device, in stream order.
static void commit_text(Ctx &c, std::unique_ptr<hera::LoadedText> t, int anchors, uint32_t test_load_num, const hera::ShardSpec *shard, hera_load_result *out,
                        const hera::HostSource *host = nullptr) {
    if (c.cache_loaded) {
        cudaStreamSynchronize(c.stream);
        if (!t->view) c.text_pool.push_back(std::move(t));
        throw hera::Error{54, "this graph was restored from a cache: hera_graph_reset before loading PAF text"};
    }
    const uint32_t file_id = static_cast<uint32_t>(c.texts.size());
    if (file_id >= 16) {
        cudaStreamSynchronize(c.stream);
        if (!t->view) c.text_pool.push_back(std::move(t));
        throw hera::Error{5, "more than 16 loads per graph are not supported"};
    }
    c.texts.push_back(std::move(t));
    try {
        hera::load_text_resident(c, *c.texts.back(), file_id, anchors != 0, test_load_num, shard, out, host);
    } catch (...) {
        cudaStreamSynchronize(c.stream);
        if (c.copy_stream) cudaStreamSynchronize(c.copy_stream);
        if (!c.texts.back()->view) c.text_pool.push_back(std::move(c.texts.back()));
        c.texts.pop_back();   // the failed load leaves no trace
        throw;
    }
}

static void load_common(Ctx &c, const void *src, uint64_t n, cudaMemcpyKind kind, int anchors, uint32_t test_load_num, const hera::ShardSpec *shard,
                        hera_load_result *out) {
    std::unique_ptr<hera::LoadedText> t = acquire_text(c, n);
    hera::Events<2> ev;
    const cudaEvent_t e0 = ev[0], e1 = ev[1];
    // A whole text in host memory is copied in chunks on the copy stream and parsed chunk by chunk as it lands
    // (load_text_resident): the parse hides behind the copy.  Slices of a sharded load and device sources: one copy, then parse.
    const uint64_t grains = static_cast<uint64_t>(std::max<int64_t>(0, c.opt_load_chunk_bytes)) / hera::LOAD_CHUNK_GRAIN;
    const bool chunked = kind == cudaMemcpyHostToDevice && !shard && grains > 0 && n > 0;
    const hera::HostSource host{static_cast<const uint8_t *>(src), std::max(grains, hera::LOAD_CHUNK_MIN_GRAINS) * hera::LOAD_CHUNK_GRAIN};
    HERA_CUDA(cudaEventRecord(e0, c.stream));
    if (n && !chunked) HERA_CUDA(cudaMemcpyAsync(t->bytes.p, src, n, kind, c.stream));
    HERA_CUDA(cudaEventRecord(e1, c.stream));
    hera_load_result tmp{};
    commit_text(c, std::move(t), anchors, test_load_num, shard, &tmp, chunked ? &host : nullptr);
    if (!chunked) {
        float copy_ms = 0;
        HERA_CUDA(cudaEventElapsedTime(&copy_ms, e0, e1));
        tmp.h2d_ms = kind == cudaMemcpyHostToDevice ? copy_ms : 0.0f;
    }
    if (out) *out = tmp;
}

// File bytes [b, e) -> device buffer, through two pinned staging blocks used in turn: the read of block i + 1 from
// the file overlaps the host-to-device copy of block i, and the host never holds more than two blocks of the file
// (a 42 GB PAF file needs 128 MiB of pinned memory, not 42 GB).  Returns the milliseconds the copies took on the stream.
static float stream_file_to_device(Ctx &c, std::ifstream &f, const char *path, uint64_t b, uint64_t e, uint8_t *dst) {
    constexpr uint64_t BLOCK = 64ull << 20;
    struct Staging {
        char *p[2] = {nullptr, nullptr};
        cudaEvent_t done[2] = {nullptr, nullptr}, e0 = nullptr, e1 = nullptr;
        ~Staging() {
            for (int i = 0; i < 2; i++) { if (p[i]) cudaFreeHost(p[i]); if (done[i]) cudaEventDestroy(done[i]); }
            if (e0) cudaEventDestroy(e0);
            if (e1) cudaEventDestroy(e1);
        }
    } st;
    const uint64_t block = std::min<uint64_t>(BLOCK, std::max<uint64_t>(e - b, 1));
    for (int i = 0; i < 2; i++) {
        if (i == 0 || e - b > block) HERA_CUDA(cudaMallocHost(&st.p[i], block));
        HERA_CUDA(cudaEventCreateWithFlags(&st.done[i], cudaEventDisableTiming));
    }
    HERA_CUDA(cudaEventCreate(&st.e0));
    HERA_CUDA(cudaEventCreate(&st.e1));
    f.clear();
    f.seekg(static_cast<std::streamoff>(b));
    HERA_CUDA(cudaEventRecord(st.e0, c.stream));
    int k = 0;
    for (uint64_t at = b; at < e; at += block, k ^= 1) {
        const uint64_t len = std::min(block, e - at);
        HERA_CUDA(cudaEventSynchronize(st.done[k]));   // the copy that last used this block has finished (no-op the first time)
        f.read(st.p[k], static_cast<std::streamsize>(len));
        HERA_REQUIRE(static_cast<uint64_t>(f.gcount()) == len, 1, std::string("short read on ") + path);
        HERA_CUDA(cudaMemcpyAsync(dst + (at - b), st.p[k], len, cudaMemcpyHostToDevice, c.stream));
        HERA_CUDA(cudaEventRecord(st.done[k], c.stream));
    }
    HERA_CUDA(cudaEventRecord(st.e1, c.stream));
    HERA_CUDA(cudaEventSynchronize(st.e1));   // the staging blocks are freed on return
    float ms = 0;
    HERA_CUDA(cudaEventElapsedTime(&ms, st.e0, st.e1));
    return ms;
}

int hera_graph_load_bytes(hera_ctx *ctx, const void *host_text, uint64_t n, int anchors, uint32_t test_load_num, hera_load_result *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(host_text != nullptr || n == 0, 101, "null text");
    load_common(c, host_text, n, cudaMemcpyHostToDevice, anchors, test_load_num, nullptr, out);
    HERA_API_END
}

int hera_graph_load_device(hera_ctx *ctx, const void *device_text, uint64_t n, int anchors, uint32_t test_load_num, hera_load_result *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(device_text != nullptr || n == 0, 101, "null text");
    load_common(c, device_text, n, cudaMemcpyDeviceToDevice, anchors, test_load_num, nullptr, out);
    HERA_API_END
}

int hera_graph_load_device_view(hera_ctx *ctx, const void *device_text, uint64_t n, int anchors, uint32_t test_load_num, hera_load_result *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(device_text != nullptr || n == 0, 101, "null text");
    HERA_REQUIRE((reinterpret_cast<uintptr_t>(device_text) & 15u) == 0, 101, "hera_graph_load_device_view: the text must start at a 16-byte aligned address");
    if (n == 0) {   // nothing to look at: an empty text of our own keeps every later stage on its usual path
        load_common(c, device_text, 0, cudaMemcpyDeviceToDevice, anchors, test_load_num, nullptr, out);
    } else {
        std::unique_ptr<hera::LoadedText> t(new hera::LoadedText());
        t->n = n;
        t->view = static_cast<const uint8_t *>(device_text);
        hera_load_result tmp{};
        commit_text(c, std::move(t), anchors, test_load_num, nullptr, &tmp);
        if (out) *out = tmp;
    }
    HERA_API_END
}

void hera_shard_range(uint64_t file_bytes, int rank, int nranks, uint64_t *begin, uint64_t *end, uint64_t *avail_end) {
    const uint64_t look_ahead = 1ull << 20;   // a record that starts in the slice must end within this many bytes after it
    if (nranks < 1) nranks = 1;
    if (rank < 0) rank = 0;
    if (rank >= nranks) rank = nranks - 1;
    // 16-byte aligned cuts keep the 128-bit loads of the tokenizer aligned when slices are views into one buffer
    auto cut = [&](int r) { return r >= nranks ? file_bytes : (file_bytes / nranks * r) & ~15ull; };
    const uint64_t b = cut(rank), e = cut(rank + 1);
    if (begin) *begin = b;
    if (end) *end = e;
    if (avail_end) *avail_end = std::min(file_bytes, e + look_ahead);
}

int hera_graph_load_shard(hera_ctx *ctx, const void *text, int on_device, uint64_t owned, uint64_t avail, int prev_byte, int at_eof, int anchors,
                          uint32_t test_load_num, hera_load_result *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(text != nullptr || avail == 0, 101, "null text");
    HERA_REQUIRE(owned <= avail, 101, "shard: owned > avail");
    const bool prev_nonws = prev_byte >= 0 && !(prev_byte == ' ' || (prev_byte >= 9 && prev_byte <= 13));
    hera::ShardSpec sp{owned, prev_nonws, at_eof != 0};
    if (on_device == 2 && avail > 0) {   // parsed in place, like hera_graph_load_device_view
        HERA_REQUIRE((reinterpret_cast<uintptr_t>(text) & 15u) == 0, 101, "hera_graph_load_shard(on_device = 2): the slice must start at a 16-byte aligned address");
        std::unique_ptr<hera::LoadedText> t(new hera::LoadedText());
        t->n = avail;
        t->view = static_cast<const uint8_t *>(text);
        hera_load_result tmp{};
        commit_text(c, std::move(t), anchors, test_load_num, &sp, &tmp);
        if (out) *out = tmp;
    } else {
        load_common(c, text, avail, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, anchors, test_load_num, &sp, out);
    }
    HERA_API_END
}

int hera_graph_load_file(hera_ctx *ctx, const char *path, int anchors, uint32_t test_load_num, hera_load_result *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(path != nullptr, 101, "null path");
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) throw hera::Error{1, std::string("Cannot open file: ") + path};
    const uint64_t n = static_cast<uint64_t>(f.tellg());
    auto load_range = [&](uint64_t b, uint64_t e, const hera::ShardSpec *sp) {
        std::unique_ptr<hera::LoadedText> t = acquire_text(c, e - b);
        float h2d_ms = 0;
        try {
            h2d_ms = stream_file_to_device(c, f, path, b, e, t->bytes.p);
        } catch (...) {
            cudaStreamSynchronize(c.stream);
            c.text_pool.push_back(std::move(t));
            throw;
        }
        hera_load_result tmp{};
        commit_text(c, std::move(t), anchors, test_load_num, sp, &tmp);
        tmp.h2d_ms = h2d_ms;
        if (out) *out = tmp;
    };
    bool whole = hera::xch_size(c) == 1;
    if (!whole) {
        // every rank reads only its byte range (+ the byte before it and the look-ahead)
        uint64_t b, e, a;
        hera_shard_range(n, hera::xch_rank(c), hera::xch_size(c), &b, &e, &a);
        int prev = -1;
        if (b) {
            f.clear();
            f.seekg(static_cast<std::streamoff>(b - 1));
            prev = f.get();
            HERA_REQUIRE(prev >= 0, 1, std::string("short read on ") + path);
        }
        const bool prev_nonws = prev >= 0 && !(prev == ' ' || (prev >= 9 && prev <= 13));
        hera::ShardSpec sp{e - b, prev_nonws, a == n};
        try {
            load_range(b, a, &sp);
        } catch (const hera::Error &err) {
            if (err.code != 20) throw;
            whole = true;   // an irregular record: all ranks fall back to the exact whole-file reader (same decision on every rank)
        }
    }
    if (whole) load_range(0, n, nullptr);
    HERA_API_END
}

int hera_nccl_unique_id(void *out128, char *err, uint64_t err_cap) {
    try {
        if (!out128) throw hera::Error{101, "null output"};
        hera::nccl_unique_id(out128);
        return 0;
    } catch (const hera::Error &e) {
        if (err && err_cap) { std::strncpy(err, e.msg.c_str(), err_cap - 1); err[err_cap - 1] = 0; }
        return e.code;
    }
}

int hera_comm_init_nccl(hera_ctx *ctx, int rank, int nranks, const void *id128) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(!c.xch, 36, "a communicator is already attached to this context");
    c.xch = hera::make_nccl_exchange(c, rank, nranks, id128);
    HERA_API_END
}

hera_local_gang *hera_comm_local_create(int nranks) {
    if (nranks < 1) return nullptr;
    hera_local_gang *g = new hera_local_gang();
    g->hub = hera::make_local_hub(nranks);
    return g;
}

void hera_comm_local_destroy(hera_local_gang *gang) { delete gang; }

int hera_comm_attach_local(hera_ctx *ctx, hera_local_gang *gang, int rank) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(gang != nullptr, 101, "null gang");
    HERA_REQUIRE(!c.xch, 36, "a communicator is already attached to this context");
    c.xch = hera::make_local_exchange(gang->hub, rank);
    HERA_API_END
}

int hera_comm_rank(hera_ctx *ctx) { return ctx ? hera::xch_rank(ctx->c) : 0; }
int hera_comm_size(hera_ctx *ctx) { return ctx ? hera::xch_size(ctx->c) : 1; }

int hera_comm_barrier(hera_ctx *ctx) {
    HERA_API_BEGIN(ctx)
    HERA_CUDA(cudaStreamSynchronize(c.stream));
    if (c.xch) c.xch->barrier(c);
    if (c.xch && getenv("HERA_DEBUG_TIMING")) {
        timespec ts;
        clock_gettime(CLOCK_REALTIME, &ts);
        fprintf(stderr, "[hera rank %d] barrier left at host clock %.3f\n", hera::xch_rank(c), static_cast<double>(ts.tv_sec % 1000) * 1e3 + static_cast<double>(ts.tv_nsec) * 1e-6);
    }
    HERA_API_END
}

int hera_graph_reset(hera_ctx *ctx) {
    HERA_API_BEGIN(ctx)
    HERA_CUDA(cudaStreamSynchronize(c.stream));
    for (auto &t : c.texts)
        if (!t->view) c.text_pool.push_back(std::move(t));   // (a view was the caller's buffer: nothing to recycle)
    c.texts.clear();
    std::sort(c.text_pool.begin(), c.text_pool.end(), [](const auto &a, const auto &b) { return a->bytes.cap > b->bytes.cap; });
    if (c.text_pool.size() > 4) c.text_pool.resize(4);   // keep the largest few
    c.own_rows.clear();
    c.load_shapes.clear();
    c.rows_total = 0;
    c.gathered.valid = false;
    c.cache_loaded = false;
    c.kept.n = 0;
    c.g.ready = false;
    c.g.N = c.g.E = c.g.S = c.g.anchors = 0;
    HERA_API_END
}

int hera_graph_finalize(hera_ctx *ctx, uint32_t metrics_mask, hera_graph_info *out) {
    HERA_API_BEGIN(ctx)
    hera::finalize_graph(c, metrics_mask, out);
    HERA_API_END
}

int hera_graph_gather_rows(hera_ctx *ctx) {
    HERA_API_BEGIN(ctx)
    (void)hera::rows_view(c);
    HERA_API_END
}

int hera_graph_save(hera_ctx *ctx, const char *path) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(path != nullptr, 101, "null path");
    hera::save_graph_cache(c, path);
    HERA_API_END
}

int hera_graph_load_cache(hera_ctx *ctx, const char *path) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(path != nullptr, 101, "null path");
    hera::load_graph_cache(c, path);
    HERA_API_END
}

int hera_graph_export(hera_ctx *ctx, hera_graph_view *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(out != nullptr, 101, "null output");
    hera::export_graph(c, out);
    HERA_API_END
}

int hera_paths_monte_carlo(hera_ctx *ctx, int metric, const hera_path_params *pp, uint64_t seed, hera_walk_stats *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(pp != nullptr, 101, "null path parameters");
    hera::walk_monte_carlo(c, metric, *pp, seed, out);
    HERA_API_END
}

int hera_paths_deterministic(hera_ctx *ctx, int metric, hera_walk_stats *out) {
    HERA_API_BEGIN(ctx)
    hera::walk_deterministic(c, metric, out);
    HERA_API_END
}

int hera_paths_build(hera_ctx *ctx, int n, const int *kinds, const int *metrics, const hera_path_params *pp, uint64_t seed, hera_walk_stats *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(pp != nullptr && kinds != nullptr && metrics != nullptr, 101, "null argument");
    hera::walk_build(c, n, kinds, metrics, *pp, seed, out);
    HERA_API_END
}

int hera_paths_unique(hera_ctx *ctx, uint64_t *n_unique) {
    HERA_API_BEGIN(ctx)
    hera::unique_paths(c, n_unique);
    HERA_API_END
}

int hera_paths_reset(hera_ctx *ctx) {
    HERA_API_BEGIN(ctx)
    HERA_CUDA(cudaStreamSynchronize(c.stream));
    c.all.paths = c.all.nodes = 0;
    c.uniq.paths = c.uniq.nodes = 0;
    c.uniq_ready = false;
    c.groups_valid = false;
    c.call_kind.clear(); c.call_metric.clear(); c.call_found.clear(); c.mc_draws.clear();
    HERA_API_END
}

int hera_paths_export(hera_ctx *ctx, hera_paths_view *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(out != nullptr, 101, "null output");
    hera::export_paths(c, out);
    HERA_API_END
}

int hera_groups(hera_ctx *ctx, const hera_path_params *pp, hera_groups_view *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(pp != nullptr && out != nullptr, 101, "null argument");
    c.groups_valid = false;
    hera::group_paths(c, *pp, out);
    c.groups_valid = true;
    HERA_API_END
}

int hera_scaffold(hera_ctx *ctx, const char *reads_fasta, const char *contigs_fasta, uint64_t min_path_num, hera_scaffold_view *out) {
    HERA_API_BEGIN(ctx)
    HERA_REQUIRE(out != nullptr, 101, "null output");
    hera::build_scaffold(c, reads_fasta, contigs_fasta, min_path_num, out);
    HERA_API_END
}

}  // extern "C"
