// Transports of the exchange steps (see exchange.cuh).
#include "exchange.cuh"

#include <dlfcn.h>
#include <nccl.h>   // types only: every NCCL entry point is resolved with dlsym

#include <chrono>
#include <condition_variable>
#include <cstring>
#include <mutex>

#include "ctx.cuh"

namespace hera {

int xch_rank(const Ctx &c) { return c.xch ? c.xch->rank : 0; }
int xch_size(const Ctx &c) { return c.xch ? c.xch->size : 1; }

// ---- LocalExchange ----------------------------------------------------------------------------
struct LocalHub {
    int n;
    std::mutex m;
    std::condition_variable cv;
    int arrived = 0;
    uint64_t generation = 0;
    bool aborted = false;
    std::vector<const void *> ptrs;
    std::vector<std::vector<uint64_t>> vals;
    std::vector<int> devices;   // CUDA device of every rank (-1 until it attached)
    explicit LocalHub(int nranks) : n(nranks), ptrs(nranks, nullptr), vals(nranks), devices(nranks, -1) {}

    void barrier() {
        std::unique_lock<std::mutex> lk(m);
        if (aborted) throw Error{30, "exchange aborted: another rank of the gang failed"};
        const uint64_t g = generation;
        if (++arrived == n) {
            arrived = 0;
            generation++;
            cv.notify_all();
            return;
        }
        const bool ok = cv.wait_for(lk, std::chrono::seconds(300), [&] { return generation != g || aborted; });
        if (!ok) {
            aborted = true;
            cv.notify_all();
            throw Error{31, "exchange timed out waiting for the other ranks (every rank must make the same collective calls)"};
        }
        if (aborted && generation == g) throw Error{30, "exchange aborted: another rank of the gang failed"};
    }
    void abort() {
        std::lock_guard<std::mutex> lk(m);
        aborted = true;
        cv.notify_all();
    }
};

std::shared_ptr<LocalHub> make_local_hub(int nranks) { return std::make_shared<LocalHub>(nranks); }
int local_hub_size(const LocalHub &hub) { return hub.n; }

namespace {

struct LocalExchange final : Exchange {
    std::shared_ptr<LocalHub> hub;
    bool peers_ready = false;
    LocalExchange(std::shared_ptr<LocalHub> h, int r) : hub(std::move(h)) { rank = r; size = hub->n; }
    const char *transport() const override { return "local"; }
    // Ranks on other GPUs pull from my buffers with peer copies: open my memory pool to them and enable
    // peer access where the hardware has it (NVLink / PCIe P2P); otherwise the driver stages through the host.
    void prepare_peers(Ctx &c) {
        if (peers_ready) return;
        {
            std::lock_guard<std::mutex> lk(hub->m);
            hub->devices[rank] = c.device;
        }
        hub->barrier();
        cudaMemPool_t pool = nullptr;
        const bool have_pool = cudaDeviceGetDefaultMemPool(&pool, c.device) == cudaSuccess;
        for (int r = 0; r < size; r++) {
            const int d = hub->devices[r];
            if (d < 0 || d == c.device) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, c.device, d) != cudaSuccess || !can) continue;
            const cudaError_t e = cudaDeviceEnablePeerAccess(d, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); continue; }
            cudaGetLastError();
            if (have_pool) {
                cudaMemAccessDesc desc{};
                desc.location.type = cudaMemLocationTypeDevice;
                desc.location.id = d;
                desc.flags = cudaMemAccessFlagsProtReadWrite;
                cudaMemPoolSetAccess(pool, &desc, 1);
                cudaGetLastError();
            }
        }
        hub->barrier();
        peers_ready = true;
    }
    void abort() override { hub->abort(); }
    void barrier(Ctx &c) override {
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        hub->barrier();
    }
    // One process: a rank's cudaMalloc'ed buffer is addressable by the others as it is (same device, or peer access).
    struct Shared { void *p = nullptr; size_t cap = 0; };
    std::vector<Shared> shared;
    ~LocalExchange() override { for (Shared &b : shared) if (b.p) cudaFree(b.p); }
    bool shared_buffers(Ctx &c, int key, size_t bytes, void **ptrs) override {
        prepare_peers(c);
        if (static_cast<size_t>(key) >= shared.size()) shared.resize(key + 1);
        Shared &b = shared[key];
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        hub->barrier();                       // nobody still writes into the buffer that may be replaced
        if (bytes > b.cap) {
            if (b.p) HERA_CUDA(cudaFree(b.p));
            b.cap = bytes + bytes / 4 + 256;
            HERA_CUDA(cudaMalloc(&b.p, b.cap));
        }
        {
            std::lock_guard<std::mutex> lk(hub->m);
            hub->ptrs[rank] = b.p;
        }
        hub->barrier();
        bool ok = true;
        for (int r = 0; r < size; r++) {
            ptrs[r] = const_cast<void *>(hub->ptrs[r]);
            if (hub->devices[r] != c.device) {
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, c.device, hub->devices[r]) != cudaSuccess || !can) ok = false;
            }
        }
        hub->barrier();                       // everybody has read the table before it is reused
        HERA_REQUIRE(ok, 39, "local gang: the ranks' devices cannot address each other's memory (peer access is needed)");
        return true;
    }
    void all_reduce_sum_u32(Ctx &, uint32_t *, size_t) override { throw Error{39, "local gang: all_reduce is not needed (buffers are shared)"}; }
    void all_gather_host(Ctx &c, const uint64_t *in, int count, uint64_t *out) override {
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        {
            std::lock_guard<std::mutex> lk(hub->m);
            hub->vals[rank].assign(in, in + count);
        }
        hub->barrier();
        for (int r = 0; r < size; r++) {
            HERA_REQUIRE(static_cast<int>(hub->vals[r].size()) == count, 32, "exchange: ranks disagree on the size of a host gather");
            std::memcpy(out + static_cast<size_t>(r) * count, hub->vals[r].data(), sizeof(uint64_t) * count);
        }
        hub->barrier();
    }
    void all_gather_dev(Ctx &c, const uint64_t *d_in, uint64_t *d_out, size_t count) override {
        std::vector<uint64_t> counts(size, count), displs(size);
        for (int r = 0; r < size; r++) displs[r] = static_cast<uint64_t>(r) * count;
        all_gatherv(c, d_in, d_out, counts.data(), displs.data(), sizeof(uint64_t));
    }
    void all_gatherv(Ctx &c, const void *send, void *recv, const uint64_t *counts, const uint64_t *displs, size_t elem) override {
        prepare_peers(c);
        HERA_CUDA(cudaStreamSynchronize(c.stream));   // my contribution is complete
        {
            std::lock_guard<std::mutex> lk(hub->m);
            hub->ptrs[rank] = send;
        }
        hub->barrier();
        for (int r = 0; r < size; r++) {
            if (!counts[r]) continue;
            char *dst = static_cast<char *>(recv) + displs[r] * elem;
            const void *src = hub->ptrs[r];
            if (dst == src) continue;
            if (hub->devices[r] == c.device) HERA_CUDA(cudaMemcpyAsync(dst, src, counts[r] * elem, cudaMemcpyDeviceToDevice, c.stream));
            else HERA_CUDA(cudaMemcpyPeerAsync(dst, c.device, src, hub->devices[r], counts[r] * elem, c.stream));
        }
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        hub->barrier();                               // nobody reuses a send buffer before every rank has pulled it
    }
};

}  // namespace

// ---- gather_rows ------------------------------------------------------------------------------
void Exchange::gather_rows(Ctx &c, const Column *cols, int ncols, uint64_t n0, uint64_t mine0, const uint64_t *counts, const uint64_t *displs) {
    // my rows move up to their global place: stage them so that no receive window overlaps a source
    bool any_inplace = false;
    for (int k = 0; k < ncols; k++) any_inplace = any_inplace || cols[k].src == nullptr;
    const bool staged = any_inplace && mine0 != n0 + displs[rank] && counts[rank] != 0;
    TempPool pool(c.stream);
    char *tmp0 = nullptr;
    if (staged) {
        size_t bytes = 0;
        for (int k = 0; k < ncols; k++) bytes += (counts[rank] * cols[k].elem + 255) / 256 * 256;
        tmp0 = pool.get<char>(bytes);
    }
    char *tmp = tmp0;
    begin_batch();
    for (int k = 0; k < ncols; k++) {
        char *col = static_cast<char *>(cols[k].base) + n0 * cols[k].elem;
        const void *send = cols[k].src ? cols[k].src : static_cast<char *>(cols[k].base) + mine0 * cols[k].elem;
        if (staged && !cols[k].src) {   // every column gets its own staging area: the batch runs at end_batch
            HERA_CUDA(cudaMemcpyAsync(tmp, send, counts[rank] * cols[k].elem, cudaMemcpyDeviceToDevice, c.stream));
            send = tmp;
            tmp += (counts[rank] * cols[k].elem + 255) / 256 * 256;
        }
        all_gatherv(c, send, col, counts, displs, cols[k].elem);
    }
    end_batch();
}

namespace {

constexpr int GR_MAX_COLS = 16, GR_MAX_RANKS = 16;
struct RowGatherArgs {
    void *col[GR_MAX_COLS];
    const void *src[GR_MAX_COLS];       // this rank's rows of the column when they are not inside it (else null)
    uint32_t elem[GR_MAX_COLS];
    const char *window[GR_MAX_RANKS];   // every rank's packed rows (window[r] lives on GPU r)
    uint64_t count[GR_MAX_RANKS], displ[GR_MAX_RANKS];
    uint64_t n0, mine0;
    int ncols, ranks, skip;   // skip: the rank whose rows are already in place (-1: none)
    int me;                   // the pulling rank: sources are visited in the order me + 1, me + 2, ... (mod ranks)
};
// Packed layout of a rank's rows inside its window: column after column, each starting on a 16-byte boundary.
__host__ __device__ inline uint64_t window_col_offset(const uint32_t *elem, int k, uint64_t count) {
    uint64_t off = 0;
    for (int j = 0; j < k; j++) off += (count * elem[j] + 15) / 16 * 16;
    return off;
}

// blockIdx.y = column.  Copies this rank's rows of the column into its window (4-byte words, the tail byte-wise).
__global__ void __launch_bounds__(256) rows_pack_kernel(RowGatherArgs a, int me, char *window) {
    const int k = blockIdx.y;
    const uint64_t bytes = a.count[me] * a.elem[k];
    const char *src = a.src[k] ? static_cast<const char *>(a.src[k]) : static_cast<const char *>(a.col[k]) + a.mine0 * a.elem[k];
    char *dst = window + window_col_offset(a.elem, k, a.count[me]);
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x, t = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (a.elem[k] % 4 == 0) {
        for (uint64_t i = t; i < bytes / 4; i += stride) reinterpret_cast<uint32_t *>(dst)[i] = reinterpret_cast<const uint32_t *>(src)[i];
    } else {
        for (uint64_t i = t; i < bytes; i += stride) dst[i] = src[i];
    }
}

// blockIdx.y = column, blockIdx.z = source: rank (me + 1 + z) mod ranks.  Reads that rank's packed rows of the column
// out of ITS memory (peer loads over NVLink; 16 bytes per load, the window side is aligned) and writes them to their
// global place in this rank's column.  Blocks are scheduled in z order, so at any moment the ranks pull from
// DIFFERENT sources (a ring schedule): with the same order everywhere all ranks would read from one GPU at a time and
// share its NVLink egress.
__global__ void __launch_bounds__(256) rows_pull_kernel(RowGatherArgs a) {
    const int k = blockIdx.y, r = (a.me + 1 + static_cast<int>(blockIdx.z)) % a.ranks;
    const uint64_t bytes = a.count[r] * a.elem[k];
    if (!bytes || r == a.skip) return;
    const char *src = a.window[r] + window_col_offset(a.elem, k, a.count[r]);
    char *dst = static_cast<char *>(a.col[k]) + (a.n0 + a.displ[r]) * a.elem[k];
    const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x, t = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint32_t e = a.elem[k];
    if (e == 4) {
        for (uint64_t i = t; i < (bytes + 15) / 16; i += stride) {
            const uint4 v = reinterpret_cast<const uint4 *>(src)[i];
            const uint64_t w = i * 4, words = bytes / 4;
            uint32_t *d = reinterpret_cast<uint32_t *>(dst);
            if (w + 4 <= words) { d[w] = v.x; d[w + 1] = v.y; d[w + 2] = v.z; d[w + 3] = v.w; }
            else { if (w < words) d[w] = v.x; if (w + 1 < words) d[w + 1] = v.y; if (w + 2 < words) d[w + 2] = v.z; }
        }
    } else if (e == 8) {
        for (uint64_t i = t; i < (bytes + 15) / 16; i += stride) {
            const uint4 v = reinterpret_cast<const uint4 *>(src)[i];
            const uint64_t w = i * 2, words = bytes / 8;
            uint2 *d = reinterpret_cast<uint2 *>(dst);
            d[w] = make_uint2(v.x, v.y);
            if (w + 1 < words) d[w + 1] = make_uint2(v.z, v.w);
        }
    } else {
        for (uint64_t i = t; i < bytes; i += stride) dst[i] = src[i];
    }
}

// ---- mailbox all-gather: a few 64-bit words per rank, exchanged by peer stores over NVLink ------------------------------
// Every rank owns a small cudaMalloc'ed mailbox that all other ranks map through CUDA IPC.  For gather number `seq`,
// rank r stores its words into slot [seq & 1][r] of EVERY rank's mailbox (its own included), fences, then stores seq
// into the slot's tag with release semantics; it then polls the tags of its own mailbox until every rank's message of
// this gather has arrived.  One launch of 32 threads, no NCCL kernel, no proxy thread: the latency is one NVLink
// round trip.  Two parities are enough: a rank can only start gather seq + 1 after it has received everybody's seq
// message, and nobody can start seq + 2 before this rank has sent seq + 1, i.e. before it is done reading parity seq & 1.
constexpr int MB_MAX_RANKS = GR_MAX_RANKS, MB_MAX_WORDS = 7, MB_SLOT = 1 + MB_MAX_WORDS;   // slot = {tag, words}
constexpr size_t MB_BYTES = 2ull * MB_MAX_RANKS * MB_SLOT * sizeof(unsigned long long);
struct MailboxArgs {
    unsigned long long *box[MB_MAX_RANKS];   // every rank's mailbox as mapped here (box[me] = mine)
    const unsigned long long *in;
    unsigned long long *out;
    unsigned long long seq;
    int me, ranks, count;
};
__global__ void __launch_bounds__(32) mailbox_all_gather_kernel(MailboxArgs a) {
    const int r = static_cast<int>(threadIdx.x);
    const int par = static_cast<int>(a.seq & 1ull);
    __threadfence_system();   // what earlier kernels of this stream stored into peer memory is ordered before the tags below
    if (r < a.ranks) {
        unsigned long long *slot = a.box[r] + (static_cast<size_t>(par) * MB_MAX_RANKS + a.me) * MB_SLOT;
        for (int i = 0; i < a.count; i++) asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(slot + 1 + i), "l"(a.in[i]) : "memory");
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(slot), "l"(a.seq) : "memory");
    }
    if (r < a.ranks) {
        const unsigned long long *slot = a.box[a.me] + (static_cast<size_t>(par) * MB_MAX_RANKS + r) * MB_SLOT;
        unsigned long long tag, t0 = 0, spins = 0;
        for (;;) {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(tag) : "l"(slot) : "memory");
            if (tag == a.seq) break;
            if ((++spins & 0xfffull) == 0) {   // a rank that never arrives must not hang the GPU: give up after 30 s (the launch fails)
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (!t0) t0 = now;
                else if (now - t0 > 30000000000ull) __trap();
            }
        }
        for (int i = 0; i < a.count; i++) {
            unsigned long long v;
            asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(slot + 1 + i) : "memory");
            a.out[static_cast<size_t>(r) * a.count + i] = v;
        }
    }
}

// ---- NcclExchange -----------------------------------------------------------------------------
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
};

NcclApi &nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    static std::string failure;
    std::call_once(once, [] {
        // RTLD_NOLOAD first: inside a PyTorch process the NCCL that torch already mapped is reused
        for (int flags : {RTLD_NOW | RTLD_NOLOAD, RTLD_NOW}) {
            for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
                if ((api.lib = dlopen(name, flags | RTLD_GLOBAL))) break;
            }
            if (api.lib) break;
        }
        if (!api.lib) { failure = std::string("cannot load libnccl.so.2: ") + (dlerror() ? dlerror() : "not found"); return; }
        auto sym = [&](const char *n) {
            void *p = dlsym(api.lib, n);
            if (!p && failure.empty()) failure = std::string("libnccl lacks ") + n;
            return p;
        };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
        api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(sym("ncclBroadcast"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
        api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
    });
    if (!failure.empty()) throw Error{33, failure};
    return api;
}

#define HERA_NCCL(expr)                                                                                       \
    do {                                                                                                      \
        ncclResult_t r__ = (expr);                                                                            \
        if (r__ != ncclSuccess && r__ != ncclInProgress)                                                      \
            throw ::hera::Error{34, std::string(#expr) + ": " + nccl_api().GetErrorString(r__)};              \
    } while (0)

struct NcclExchange final : Exchange {
    ncclComm_t comm = nullptr;
    uint64_t *d_small = nullptr, *h_small = nullptr;   // device / pinned staging of host gathers
    size_t d_small_cap = 0, h_small_cap = 0;
    // Teardown in three steps (an exported allocation must not be freed while a peer still has it mapped): every rank
    // closes the handles it imported, the ranks meet, then each frees what it exported.  hera_destroy is collective on a
    // sharded context; a communicator that was aborted (a rank failed) skips the rendezvous.
    bool aborted = false;
    void abort() override { aborted = true; }
    ~NcclExchange() override {
        unmap_mailbox();
        unmap_windows();
        for (SharedBuf &b : shared) unmap_shared(b);
        if (comm && !aborted) {
            uint64_t one = 1, *d = nullptr;
            // a plain NCCL all-gather as the rendezvous (the mailbox is already unmapped)
            if (cudaMalloc(&d, (1 + size) * sizeof(uint64_t)) == cudaSuccess) {
                cudaMemcpy(d, &one, sizeof(one), cudaMemcpyHostToDevice);
                if (nccl_api().AllGather(d, d + 1, 1, ncclUint64, comm, nullptr) == ncclSuccess) cudaStreamSynchronize(nullptr);
                cudaFree(d);
            }
        }
        if (mailbox) cudaFree(mailbox);
        if (window) cudaFree(window);
        for (SharedBuf &b : shared) if (b.mine) cudaFree(b.mine);
        if (d_small) cudaFree(d_small);
        if (h_small) cudaFreeHost(h_small);
        if (d_token) cudaFree(d_token);
        if (comm) nccl_api().CommDestroy(comm);
        cudaGetLastError();
    }

    // ---- shared buffers: one cudaMalloc'ed buffer per rank and key, mapped by every other rank through CUDA IPC ----
    struct SharedBuf {
        void *mine = nullptr;
        size_t cap = 0;
        std::vector<void *> peer;   // [r] = rank r's buffer as mapped here (mine for r == rank); empty = not mapped
    };
    std::vector<SharedBuf> shared;
    void unmap_shared(SharedBuf &b) {
        for (int r = 0; r < static_cast<int>(b.peer.size()); r++)
            if (r != rank && b.peer[r]) cudaIpcCloseMemHandle(b.peer[r]);
        b.peer.clear();
    }
    bool shared_buffers(Ctx &c, int key, size_t bytes, void **ptrs) override {
        if (static_cast<size_t>(key) >= shared.size()) shared.resize(key + 1);
        SharedBuf &b = shared[key];
        const bool ipc = !(size > GR_MAX_RANKS || getenv("HERA_NO_PEER_WINDOWS"));
        // every rank asks for the same size and the buffers grow the same way, so this decision is the same everywhere
        if (bytes > b.cap || (ipc && b.peer.empty() && ipc_shared_state >= 0)) {
            HERA_CUDA(cudaStreamSynchronize(c.stream));
            unmap_shared(b);
            barrier(c);                         // every rank has closed its mappings of the buffers that are replaced
            if (bytes > b.cap) {
                if (b.mine) HERA_CUDA(cudaFree(b.mine));
                b.mine = nullptr;
                b.cap = (bytes + bytes / 4 + 4096) / 512 * 512;
                HERA_CUDA(cudaMalloc(&b.mine, b.cap));
            }
            if (ipc && ipc_shared_state >= 0) {
                uint64_t mine[9] = {0};
                cudaIpcMemHandle_t h;
                bool ok = cudaIpcGetMemHandle(&h, b.mine) == cudaSuccess;
                if (ok) { std::memcpy(mine, &h, 64); mine[8] = 1; } else cudaGetLastError();
                std::vector<uint64_t> all(static_cast<size_t>(size) * 9);
                all_gather_host(c, mine, 9, all.data());
                for (int r = 0; r < size; r++) ok = ok && all[static_cast<size_t>(r) * 9 + 8] == 1;
                b.peer.assign(size, nullptr);
                for (int r = 0; ok && r < size; r++) {
                    if (r == rank) { b.peer[r] = b.mine; continue; }
                    cudaIpcMemHandle_t hr;
                    std::memcpy(&hr, &all[static_cast<size_t>(r) * 9], 64);
                    void *q = nullptr;
                    if (cudaIpcOpenMemHandle(&q, hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
                    b.peer[r] = q;
                }
                uint64_t vote = ok ? 1 : 0;   // all ranks must take the same path
                std::vector<uint64_t> votes(size);
                all_gather_host(c, &vote, 1, votes.data());
                for (uint64_t v : votes) ok = ok && v == 1;
                if (!ok) { unmap_shared(b); ipc_shared_state = -1; }
            }
        }
        const bool mapped = !b.peer.empty();
        for (int r = 0; r < size; r++) ptrs[r] = mapped ? b.peer[r] : (r == rank ? b.mine : nullptr);
        return mapped;
    }
    int ipc_shared_state = 0;   // -1: mapping each other's buffers failed once, do not try again
    void all_reduce_sum_u32(Ctx &c, uint32_t *buf, size_t count) override {
        HERA_NCCL(nccl_api().AllReduce(buf, buf, count, ncclUint32, ncclSum, comm, c.stream));
    }
    const char *transport() const override { return "nccl"; }
    void all_gather_host(Ctx &c, const uint64_t *in, int count, uint64_t *out) override {
        NcclApi &n = nccl_api();
        const size_t need = static_cast<size_t>(count) * (size + 1);
        if (need > d_small_cap) {
            if (d_small) HERA_CUDA(cudaFree(d_small));
            HERA_CUDA(cudaMalloc(&d_small, need * 2 * sizeof(uint64_t)));
            d_small_cap = need * 2;
        }
        if (need > h_small_cap) {
            if (h_small) HERA_CUDA(cudaFreeHost(h_small));
            HERA_CUDA(cudaHostAlloc(&h_small, need * 2 * sizeof(uint64_t), cudaHostAllocDefault));
            h_small_cap = need * 2;
        }
        uint64_t *d_in = d_small, *d_out = d_small + count;
        std::memcpy(h_small, in, sizeof(uint64_t) * count);   // pinned staging: pageable copies would serialise with the stream
        HERA_CUDA(cudaMemcpyAsync(d_in, h_small, sizeof(uint64_t) * count, cudaMemcpyHostToDevice, c.stream));
        if (mailbox_state > 0 && count <= MB_MAX_WORDS) all_gather_dev(c, d_in, d_out, static_cast<size_t>(count));   // peer stores, no NCCL kernel
        else HERA_NCCL(n.AllGather(d_in, d_out, static_cast<size_t>(count), ncclUint64, comm, c.stream));
        HERA_CUDA(cudaMemcpyAsync(h_small + count, d_out, sizeof(uint64_t) * count * size, cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        std::memcpy(out, h_small + count, sizeof(uint64_t) * count * size);
    }
    void all_gather_dev(Ctx &c, const uint64_t *d_in, uint64_t *d_out, size_t count) override {
        if (count <= static_cast<size_t>(MB_MAX_WORDS) && ensure_mailbox(c)) {
            MailboxArgs a{};
            for (int r = 0; r < size; r++) a.box[r] = peer_mailbox[r];
            a.in = reinterpret_cast<const unsigned long long *>(d_in);
            a.out = reinterpret_cast<unsigned long long *>(d_out);
            a.seq = ++mailbox_seq;
            a.me = rank; a.ranks = size; a.count = static_cast<int>(count);
            mailbox_all_gather_kernel<<<1, 32, 0, c.stream>>>(a);
            c.launches.count++;
            HERA_CUDA(cudaGetLastError());
            return;
        }
        HERA_NCCL(nccl_api().AllGather(d_in, d_out, count, ncclUint64, comm, c.stream));
    }

    // ---- mailbox (see mailbox_all_gather_kernel): allocated and mapped once per communicator ----
    unsigned long long *mailbox = nullptr;
    std::vector<unsigned long long *> peer_mailbox;
    unsigned long long mailbox_seq = 0;
    int mailbox_state = 0;             // 0 untested, 1 usable, -1 not available (NCCL all-gathers instead)
    void unmap_mailbox() {
        for (int r = 0; r < static_cast<int>(peer_mailbox.size()); r++)
            if (r != rank && peer_mailbox[r]) cudaIpcCloseMemHandle(peer_mailbox[r]);
        peer_mailbox.clear();
    }
    void close_mailbox() {   // (only on the failure path of ensure_mailbox: no rank uses a mailbox that not every rank could map)
        unmap_mailbox();
        if (mailbox) cudaFree(mailbox);
        mailbox = nullptr;
    }
    // Collective the first time (every rank reaches its first all_gather_dev at the same point of the protocol).
    bool ensure_mailbox(Ctx &c) {
        if (mailbox_state) return mailbox_state > 0;
        if (size > MB_MAX_RANKS || getenv("HERA_NO_PEER_WINDOWS") || getenv("HERA_NO_MAILBOX")) { mailbox_state = -1; return false; }
        uint64_t mine[9] = {0};
        cudaIpcMemHandle_t h;
        // (the tags are zero before any rank can know the handle: the host gather below is the barrier)
        bool ok = cudaMalloc(&mailbox, MB_BYTES) == cudaSuccess && cudaMemset(mailbox, 0, MB_BYTES) == cudaSuccess &&
                  cudaDeviceSynchronize() == cudaSuccess && cudaIpcGetMemHandle(&h, mailbox) == cudaSuccess;
        if (ok) { std::memcpy(mine, &h, 64); mine[8] = 1; } else cudaGetLastError();
        std::vector<uint64_t> all(static_cast<size_t>(size) * 9);
        all_gather_host(c, mine, 9, all.data());
        for (int r = 0; r < size; r++) ok = ok && all[static_cast<size_t>(r) * 9 + 8] == 1;
        peer_mailbox.assign(size, nullptr);
        for (int r = 0; ok && r < size; r++) {
            if (r == rank) { peer_mailbox[r] = mailbox; continue; }
            cudaIpcMemHandle_t hr;
            std::memcpy(&hr, &all[static_cast<size_t>(r) * 9], 64);
            void *p = nullptr;
            if (cudaIpcOpenMemHandle(&p, hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
            peer_mailbox[r] = static_cast<unsigned long long *>(p);
        }
        uint64_t vote = ok ? 1 : 0;   // all ranks must take the same path
        std::vector<uint64_t> votes(size);
        all_gather_host(c, &vote, 1, votes.data());
        for (uint64_t v : votes) ok = ok && v == 1;
        if (!ok) { close_mailbox(); mailbox_state = -1; return false; }
        mailbox_state = 1;
        return true;
    }

    // ---- peer windows: one cudaMalloc'ed buffer per rank, mapped by every other rank through CUDA IPC ----
    char *window = nullptr;
    size_t window_cap = 0;
    std::vector<char *> peer_window;   // [r] = rank r's window as seen from this process (mine for r == rank)
    unsigned gather_seq = 0;
    int ipc_state = 0;                 // 0 untested, 1 usable, -1 not available (fall back to NCCL broadcasts)
    uint64_t *d_token = nullptr;       // two words for the stream-ordered rendezvous

    void unmap_windows() {
        for (int r = 0; r < static_cast<int>(peer_window.size()); r++)
            if (r != rank && peer_window[r]) cudaIpcCloseMemHandle(peer_window[r]);
        peer_window.clear();
    }
    void free_window() {
        if (window) cudaFree(window);
        window = nullptr;
        window_cap = 0;
    }
    void close_windows() { unmap_windows(); free_window(); }
    // Every rank calls this with the same `need` (it is computed from the gathered counts).
    bool ensure_windows(Ctx &c, size_t need) {
        if (ipc_state < 0) return false;
        if (ipc_state == 0 && (size > GR_MAX_RANKS || getenv("HERA_NO_PEER_WINDOWS"))) { ipc_state = -1; return false; }
        if (need <= window_cap) return true;
        HERA_CUDA(cudaStreamSynchronize(c.stream));   // nobody still reads the old windows: every pull is followed by a collective
        unmap_windows();
        barrier(c);                                   // every rank has closed its mappings before any window is freed
        free_window();
        const size_t cap = (need + need / 2 + (1u << 20)) / 512 * 512;   // two halves, each a multiple of 256 bytes
        uint64_t mine[9] = {0};
        cudaIpcMemHandle_t h;
        static_assert(sizeof(h) == 64, "IPC handle is 64 bytes");
        bool ok = cudaMalloc(&window, cap) == cudaSuccess && cudaIpcGetMemHandle(&h, window) == cudaSuccess;
        if (ok) { std::memcpy(mine, &h, 64); mine[8] = 1; } else cudaGetLastError();
        std::vector<uint64_t> all(static_cast<size_t>(size) * 9);
        all_gather_host(c, mine, 9, all.data());
        for (int r = 0; r < size; r++) ok = ok && all[static_cast<size_t>(r) * 9 + 8] == 1;
        peer_window.assign(size, nullptr);
        for (int r = 0; ok && r < size; r++) {
            if (r == rank) { peer_window[r] = window; continue; }
            cudaIpcMemHandle_t hr;
            std::memcpy(&hr, &all[static_cast<size_t>(r) * 9], 64);
            void *p = nullptr;
            if (cudaIpcOpenMemHandle(&p, hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
            peer_window[r] = static_cast<char *>(p);
        }
        // all ranks must take the same path
        uint64_t vote = ok ? 1 : 0;
        std::vector<uint64_t> votes(size);
        all_gather_host(c, &vote, 1, votes.data());
        for (uint64_t v : votes) ok = ok && v == 1;
        if (!ok) { close_windows(); ipc_state = -1; return false; }
        window_cap = cap;
        ipc_state = 1;
        if (!d_token) HERA_CUDA(cudaMalloc(&d_token, (1 + GR_MAX_RANKS) * sizeof(uint64_t)));
        return true;
    }

    void gather_rows(Ctx &c, const Column *cols, int ncols, uint64_t n0, uint64_t mine0, const uint64_t *counts, const uint64_t *displs) override {
        RowGatherArgs a{};
        HERA_REQUIRE(ncols <= GR_MAX_COLS, 36, "gather_rows: too many columns");
        uint64_t max_rows = 0;
        for (int r = 0; r < size; r++) max_rows = std::max(max_rows, counts[r]);
        bool all_inplace = true;
        for (int k = 0; k < ncols; k++) { a.col[k] = cols[k].base; a.elem[k] = cols[k].elem; a.src[k] = cols[k].src; all_inplace = all_inplace && !cols[k].src; }
        const size_t need = window_col_offset(a.elem, ncols, max_rows) + 16;
        if (!ensure_windows(c, 2 * ((need + 255) / 256 * 256))) { Exchange::gather_rows(c, cols, ncols, n0, mine0, counts, displs); return; }
        // The windows have two halves used in turn: a half is written again two gathers later, and no rank gets past
        // the rendezvous of the gather in between before every rank has finished pulling from it.
        const size_t half = (gather_seq++ & 1u) ? window_cap / 2 : 0;
        for (int r = 0; r < size; r++) { a.window[r] = peer_window[r] + half; a.count[r] = counts[r]; a.displ[r] = displs[r]; }
        a.n0 = n0; a.mine0 = mine0; a.ncols = ncols; a.ranks = size;
        a.skip = all_inplace && mine0 == n0 + displs[rank] ? rank : -1;
        a.me = rank;
        if (counts[rank]) rows_pack_kernel<<<dim3(148, ncols), 256, 0, c.stream>>>(a, rank, window + half);
        // stream-ordered rendezvous: an NCCL all-gather of one word completes on a rank only after every rank's
        // stream has passed its pack kernel
        all_gather_dev(c, d_token, d_token + 1, 1);
        rows_pull_kernel<<<dim3(74, ncols, size), 256, 0, c.stream>>>(a);
        c.launches.count += 2;
        HERA_CUDA(cudaGetLastError());
    }
    void all_gatherv(Ctx &c, const void *send, void *recv, const uint64_t *counts, const uint64_t *displs, size_t elem) override {
        NcclApi &n = nccl_api();
        HERA_NCCL(n.GroupStart());
        for (int r = 0; r < size; r++) {
            if (!counts[r]) continue;
            char *dst = static_cast<char *>(recv) + displs[r] * elem;
            HERA_NCCL(n.Broadcast(r == rank ? send : dst, dst, counts[r] * elem, ncclInt8, r, comm, c.stream));
        }
        HERA_NCCL(n.GroupEnd());
    }
    void begin_batch() override { HERA_NCCL(nccl_api().GroupStart()); }
    void end_batch() override { HERA_NCCL(nccl_api().GroupEnd()); }
    void barrier(Ctx &c) override {
        uint64_t one = 1;
        std::vector<uint64_t> all(size);
        all_gather_host(c, &one, 1, all.data());
    }
};

}  // namespace

std::unique_ptr<Exchange> make_local_exchange(std::shared_ptr<LocalHub> hub, int rank) {
    HERA_REQUIRE(hub && rank >= 0 && rank < hub->n, 35, "local exchange: rank out of range");
    return std::unique_ptr<Exchange>(new LocalExchange(std::move(hub), rank));
}

void nccl_unique_id(void *out128) {
    static_assert(sizeof(ncclUniqueId) == NCCL_ID_BYTES, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    HERA_NCCL(nccl_api().GetUniqueId(&id));
    std::memcpy(out128, &id, sizeof(id));
}

std::unique_ptr<Exchange> make_nccl_exchange(Ctx &c, int rank, int nranks, const void *id128) {
    HERA_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks && id128, 35, "nccl exchange: bad rank / size / id");
    NcclApi &n = nccl_api();
    std::unique_ptr<NcclExchange> x(new NcclExchange());
    x->rank = rank;
    x->size = nranks;
    ncclUniqueId id;
    std::memcpy(&id, id128, sizeof(id));
    HERA_CUDA(cudaSetDevice(c.device));
    HERA_NCCL(n.CommInitRank(&x->comm, nranks, id, rank));
    return x;
}

}  // namespace hera
