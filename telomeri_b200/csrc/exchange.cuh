// Exchange steps of the sharded hot path: one rank per GPU.
//
// The reference is single-threaded and has no counterpart; what is exchanged is its own data:
// kept overlaps ("edges_", src/OverlapGraph.cpp:233-239) after the byte-range sharded load, and
// accepted paths ("paths_", src/PathManager.cpp:146-155, 372-384) after the walk-sharded search.
// Two transports implement the same three operations:
//   * NcclExchange  -- one process per GPU (torchrun / mpirun); NCCL over NVLink 5 / NVSwitch on the
//                      context's own stream.  libnccl.so.2 is dlopen'ed when the communicator is
//                      created, so the library itself loads on boxes without NCCL.
//   * LocalExchange -- several contexts of ONE process (one host thread per rank, same or different
//                      GPUs); host rendezvous + cudaMemcpyAsync between the ranks' buffers (peer copies
//                      over NVLink when the contexts sit on different GPUs).  No device-side waiting.
#pragma once

#include <cstdint>
#include <memory>
#include <vector>

#include "common.cuh"

namespace hera {

struct Ctx;

struct Exchange {
    int rank = 0, size = 1;
    virtual ~Exchange() = default;
    // Every rank contributes `count` values; out[r * count + k] = value k of rank r (host memory).
    // Synchronises the context's stream.
    virtual void all_gather_host(Ctx &c, const uint64_t *in, int count, uint64_t *out) = 0;
    // Variable-size all-gather of device memory: rank r contributes counts[r] elements of `elem`
    // bytes starting at `send`; every rank receives them at recv + displs[r] * elem.  `send` must not
    // overlap any rank's receive window except its own (send == recv + displs[rank] * elem is allowed).
    // Stream-ordered on the context's stream.
    virtual void all_gatherv(Ctx &c, const void *send, void *recv, const uint64_t *counts, const uint64_t *displs, size_t elem) = 0;
    // Equal-size all-gather of device memory: every rank contributes `count` 64-bit values at d_in and receives
    // all of them at d_out (rank r's at d_out + r * count; d_in may be d_out + rank * count).  Stream-ordered,
    // no host synchronisation on the NCCL transport.
    virtual void all_gather_dev(Ctx &c, const uint64_t *d_in, uint64_t *d_out, size_t count) = 0;
    // Several all_gatherv calls between begin_batch and end_batch may be issued as one operation (one NCCL
    // group = one launch); their send and receive buffers must stay untouched until end_batch returns.
    virtual void begin_batch() {}
    virtual void end_batch() {}
    // Rows [mine0, mine0 + counts[rank]) of every column (structure-of-arrays table) are this rank's; afterwards every
    // rank holds every rank's rows, rank r's at n0 + displs[r] (mine0 is n0, or already n0 + displs[rank]).  The base implementation issues one all_gatherv per
    // column in one batch; the NCCL transport pulls the rows straight out of the peers' memory over NVLink with
    // its own kernel when the ranks can map each other's buffers.
    // A column may name a separate source for this rank's rows (src != nullptr: counts[rank] elements there, mine0 is ignored
    // for it); the walk stage gathers per-walk results and path pools out of their scratch buffers that way.
    struct Column { void *base; uint32_t elem; const void *src = nullptr; };
    virtual void gather_rows(Ctx &c, const Column *cols, int ncols, uint64_t n0, uint64_t mine0, const uint64_t *counts, const uint64_t *displs);
    // COLLECTIVE.  A buffer of at least `bytes` bytes on every rank under `key`, owned by the exchange and kept until a
    // larger one is asked for; ptrs[r] = rank r's buffer as addressable from THIS rank's device (ptrs[rank] = its own), so
    // a kernel can write its results straight into the final tables of every GPU (peer stores over NVLink).  Returns
    // false when the ranks cannot address each other's memory: then only ptrs[rank] is valid and the caller merges with
    // all_reduce_sum_u32.  Every rank must ask for the same size.
    virtual bool shared_buffers(Ctx &c, int key, size_t bytes, void **ptrs) = 0;
    // In-place element-wise sum over the ranks (fallback of shared_buffers: the ranks filled disjoint parts of zeroed arrays).
    virtual void all_reduce_sum_u32(Ctx &c, uint32_t *buf, size_t count) = 0;
    // Host barrier (after synchronising the context's stream).
    virtual void barrier(Ctx &c) = 0;
    // Called when this rank fails inside a collective call: wakes the ranks waiting for it.
    virtual void abort() {}
    virtual const char *transport() const = 0;
};

// ---- in-process transport -------------------------------------------------------------------
struct LocalHub;   // shared by the ranks of one gang; created by hera_comm_local_create
std::shared_ptr<LocalHub> make_local_hub(int nranks);
std::unique_ptr<Exchange> make_local_exchange(std::shared_ptr<LocalHub> hub, int rank);
int local_hub_size(const LocalHub &hub);

// ---- NCCL transport ---------------------------------------------------------------------------
constexpr int NCCL_ID_BYTES = 128;
void nccl_unique_id(void *out128);   // throws Error when libnccl cannot be loaded
std::unique_ptr<Exchange> make_nccl_exchange(Ctx &c, int rank, int nranks, const void *id128);

// Helpers over the context's exchange (size 1 when none is attached).
int xch_rank(const Ctx &c);
int xch_size(const Ctx &c);

}  // namespace hera
