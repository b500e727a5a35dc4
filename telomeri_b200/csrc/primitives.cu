// Host launchers of the scan and radix-sort kernels in primitives.cuh.
#include "ctx.cuh"
#include "primitives.cuh"

namespace hera {

// hist[d * nblocks + b] = number of keys of tile b whose current digit is d.
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint32_t *__restrict__ keys, uint64_t n, int shift, uint32_t *__restrict__ hist) {
    __shared__ uint32_t h[RS_BINS];
    for (int i = threadIdx.x; i < RS_BINS; i += RS_THREADS) h[i] = 0;
    __syncthreads();
    const uint64_t base = static_cast<uint64_t>(blockIdx.x) * RS_TILE;
#pragma unroll 4
    for (int k = 0; k < RS_ITEMS; k++) {
        uint64_t i = base + static_cast<uint64_t>(k) * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & (RS_BINS - 1)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < RS_BINS; i += RS_THREADS) hist[static_cast<uint64_t>(i) * gridDim.x + blockIdx.x] = h[i];
}

// Stable scatter: elements of a tile keep their relative order inside each digit.  Every warp owns a CONTIGUOUS 512-key
// stretch of the tile and ranks it on its own: per round of 32 keys the lanes with the same digit find each other with
// match.any, the lowest of them advances the warp's private digit counter (shared memory, no atomics) and hands the
// old value back, so a key's place among the warp's keys of its digit is known without any CTA-wide barrier.  One pass
// over the digit table (thread d = digit d) then turns the per-warp counts into destinations -- tile base from the
// scanned histogram + keys of earlier warps -- and the keys, still in registers, go out.  Three barriers per tile.
__global__ void __launch_bounds__(RS_THREADS, 4) rs_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                                                                uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint64_t n,
                                                                int shift, const uint32_t *__restrict__ hist_scanned) {
    __shared__ uint32_t warp_cnt[RS_WARPS][RS_BINS];   // per warp: keys of each digit seen so far, then destination of the warp's first key of the digit
    static_assert(RS_BINS == RS_THREADS, "one thread per digit in the prefix pass");
    for (int i = threadIdx.x; i < RS_WARPS * RS_BINS; i += RS_THREADS) (&warp_cnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t lane = lane_id(), warp = warp_id();
    const uint64_t base = static_cast<uint64_t>(blockIdx.x) * RS_TILE + static_cast<uint64_t>(warp) * (RS_ITEMS * 32);
    uint32_t key[RS_ITEMS], val[RS_ITEMS];
    uint16_t place[RS_ITEMS];
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const uint64_t i = base + static_cast<uint64_t>(k) * 32 + lane;
        key[k] = i < n ? keys_in[i] : 0u;
        val[k] = i < n ? vals_in[i] : 0u;
    }
    uint32_t *mine = warp_cnt[warp];
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const uint64_t i = base + static_cast<uint64_t>(k) * 32 + lane;
        const uint32_t d = i < n ? ((key[k] >> shift) & (RS_BINS - 1)) : RS_BINS + lane;   // dead lanes match nobody
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (static_cast<int>(lane) == leader && i < n) { old = mine[d]; mine[d] = old + __popc(peers); }
        old = __shfl_sync(0xffffffffu, old, leader);
        place[k] = static_cast<uint16_t>(old + __popc(peers & ((1u << lane) - 1u)));
        __syncwarp();
    }
    __syncthreads();
    {   // digit d: destinations of the warps' first keys
        const uint32_t d = threadIdx.x;
        uint32_t run = hist_scanned[static_cast<uint64_t>(d) * gridDim.x + blockIdx.x];
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) { const uint32_t t = warp_cnt[w][d]; warp_cnt[w][d] = run; run += t; }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const uint64_t i = base + static_cast<uint64_t>(k) * 32 + lane;
        if (i < n) {
            const uint32_t dst = mine[(key[k] >> shift) & (RS_BINS - 1)] + place[k];
            keys_out[dst] = key[k];
            vals_out[dst] = val[k];
        }
    }
}

namespace {

template <typename TIn, typename TOut>
void scan_impl(Ctx &c, const TIn *in, uint64_t n, TOut *out, TOut *d_total, int depth = 0) {
    if (n == 0) {
        if (d_total) HERA_CUDA(cudaMemsetAsync(d_total, 0, sizeof(TOut), c.stream));
        return;
    }
    const unsigned tiles = div_up(n, SCAN_TILE);
    if (tiles == 1) {
        HERA_LAUNCH(c, (scan_apply_kernel<TIn, TOut>), 1, SCAN_THREADS, 0, in, n, static_cast<const TOut *>(nullptr), out, d_total);
        return;
    }
    // tile sums live in a scratch allocation private to this recursion level
    TOut *sums = nullptr;
    HERA_CUDA(cudaMallocAsync(&sums, sizeof(TOut) * tiles, c.stream));
    HERA_LAUNCH(c, (scan_tile_sums_kernel<TIn, TOut>), tiles, SCAN_THREADS, 0, in, n, sums);
    scan_impl<TOut, TOut>(c, sums, tiles, sums, static_cast<TOut *>(nullptr), depth + 1);
    HERA_LAUNCH(c, (scan_apply_kernel<TIn, TOut>), tiles, SCAN_THREADS, 0, in, n, static_cast<const TOut *>(sums), out, d_total);
    HERA_CUDA(cudaFreeAsync(sums, c.stream));
}

}  // namespace

void exclusive_scan_u32(Ctx &c, const uint32_t *in, uint64_t n, uint32_t *out, uint32_t *d_total) { scan_impl<uint32_t, uint32_t>(c, in, n, out, d_total); }
void exclusive_scan_u8_to_u32(Ctx &c, const uint8_t *in, uint64_t n, uint32_t *out, uint32_t *d_total) { scan_impl<uint8_t, uint32_t>(c, in, n, out, d_total); }
void exclusive_scan_u32_to_u64(Ctx &c, const uint32_t *in, uint64_t n, uint64_t *out, uint64_t *d_total) { scan_impl<uint32_t, uint64_t>(c, in, n, out, d_total); }

void radix_sort_pairs(Ctx &c, uint32_t *keys, uint32_t *vals, uint32_t *keys_alt, uint32_t *vals_alt, uint64_t n, int bits) {
    if (n == 0) return;
    const unsigned tiles = div_up(n, RS_TILE);
    uint32_t *hist = nullptr;
    HERA_CUDA(cudaMallocAsync(&hist, sizeof(uint32_t) * RS_BINS * static_cast<size_t>(tiles), c.stream));
    uint32_t *kin = keys, *vin = vals, *kout = keys_alt, *vout = vals_alt;
    int passes = 0;
    for (int shift = 0; shift < bits; shift += RS_BITS, passes++) {
        HERA_LAUNCH(c, rs_hist_kernel, tiles, RS_THREADS, 0, kin, n, shift, hist);
        exclusive_scan_u32(c, hist, static_cast<uint64_t>(RS_BINS) * tiles, hist, nullptr);
        HERA_LAUNCH(c, rs_scatter_kernel, tiles, RS_THREADS, 0, kin, vin, kout, vout, n, shift, hist);
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
    if (passes & 1) {  // result sits in the alt buffers
        HERA_CUDA(cudaMemcpyAsync(keys, keys_alt, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, c.stream));
        HERA_CUDA(cudaMemcpyAsync(vals, vals_alt, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, c.stream));
    }
    HERA_CUDA(cudaFreeAsync(hist, c.stream));
}

}  // namespace hera
