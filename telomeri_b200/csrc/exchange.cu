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

// ---- NcclExchange -----------------------------------------------------------------------------
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
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
    uint64_t *d_small = nullptr;   // staging of host gathers
    size_t d_small_cap = 0;
    ~NcclExchange() override {
        if (d_small) cudaFree(d_small);
        if (comm) nccl_api().CommDestroy(comm);
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
        uint64_t *d_in = d_small, *d_out = d_small + count;
        HERA_CUDA(cudaMemcpyAsync(d_in, in, sizeof(uint64_t) * count, cudaMemcpyHostToDevice, c.stream));
        HERA_NCCL(n.AllGather(d_in, d_out, static_cast<size_t>(count), ncclUint64, comm, c.stream));
        HERA_CUDA(cudaMemcpyAsync(out, d_out, sizeof(uint64_t) * count * size, cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
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
