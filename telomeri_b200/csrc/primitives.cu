// Host launchers of the scan and radix-sort kernels in primitives.cuh.
#include <algorithm>

#include "ctx.cuh"
#include "primitives.cuh"

namespace hera {

// Digits are BITS wide (8..11): the number of passes is what a sort of 130 M pairs pays for (every pass reads and writes
// all keys and values), so a 20-bit key takes two 10-bit passes instead of three 8-bit ones.
// hist[d * nblocks + b] = number of keys of tile b whose current digit is d.
template <int BITS>
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint32_t *__restrict__ keys, uint64_t n, int shift, uint32_t *__restrict__ hist) {
    constexpr int BINS = 1 << BITS;
    __shared__ uint32_t h[BINS];
    for (int i = threadIdx.x; i < BINS; i += RS_THREADS) h[i] = 0;
    __syncthreads();
    const uint64_t base = static_cast<uint64_t>(blockIdx.x) * RS_TILE;
#pragma unroll 4
    for (int k = 0; k < RS_ITEMS; k++) {
        uint64_t i = base + static_cast<uint64_t>(k) * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & (BINS - 1)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < BINS; i += RS_THREADS) hist[static_cast<uint64_t>(i) * gridDim.x + blockIdx.x] = h[i];
}

// Stable scatter: elements of a tile keep their relative order inside each digit.  Every warp owns a CONTIGUOUS 512-key
// stretch of the tile and ranks it on its own: per round of 32 keys the lanes with the same digit find each other with
// match.any, the lowest of them advances the warp's private digit counter (shared memory, no atomics) and hands the
// old value back, so a key's place among the warp's keys of its digit is known without any CTA-wide barrier.  One pass
// over the digit table then turns the per-warp counts into destinations -- tile base from the scanned histogram + keys of
// earlier warps -- and the keys, still in registers, go out.  Three barriers per tile.
template <int BITS>
__global__ void __launch_bounds__(RS_THREADS, 4) rs_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                                                                   uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint64_t n,
                                                                   int shift, const uint32_t *__restrict__ hist_scanned) {
    constexpr int BINS = 1 << BITS;
    extern __shared__ uint32_t rs_warp_cnt[];   // [RS_WARPS][BINS] per warp: keys of each digit seen so far, then destination of the warp's first key of the digit
    for (int i = threadIdx.x; i < RS_WARPS * BINS; i += RS_THREADS) rs_warp_cnt[i] = 0;
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
    uint32_t *mine = rs_warp_cnt + warp * BINS;
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const uint64_t i = base + static_cast<uint64_t>(k) * 32 + lane;
        const uint32_t d = i < n ? ((key[k] >> shift) & (BINS - 1)) : BINS + lane;   // dead lanes match nobody
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (static_cast<int>(lane) == leader && i < n) { old = mine[d]; mine[d] = old + __popc(peers); }
        old = __shfl_sync(0xffffffffu, old, leader);
        place[k] = static_cast<uint16_t>(old + __popc(peers & ((1u << lane) - 1u)));
        __syncwarp();
    }
    __syncthreads();
    for (uint32_t d = threadIdx.x; d < static_cast<uint32_t>(BINS); d += RS_THREADS) {   // digit d: destinations of the warps' first keys
        uint32_t run = hist_scanned[static_cast<uint64_t>(d) * gridDim.x + blockIdx.x];
#pragma unroll
        for (int w = 0; w < RS_WARPS; w++) { const uint32_t t = rs_warp_cnt[w * BINS + d]; rs_warp_cnt[w * BINS + d] = run; run += t; }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RS_ITEMS; k++) {
        const uint64_t i = base + static_cast<uint64_t>(k) * 32 + lane;
        if (i < n) {
            const uint32_t dst = mine[(key[k] >> shift) & (BINS - 1)] + place[k];
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
    TempPool tmp(c.stream);
    TOut *sums = tmp.get<TOut>(tiles);
    HERA_LAUNCH(c, (scan_tile_sums_kernel<TIn, TOut>), tiles, SCAN_THREADS, 0, in, n, sums);
    scan_impl<TOut, TOut>(c, sums, tiles, sums, static_cast<TOut *>(nullptr), depth + 1);
    HERA_LAUNCH(c, (scan_apply_kernel<TIn, TOut>), tiles, SCAN_THREADS, 0, in, n, static_cast<const TOut *>(sums), out, d_total);
}

}  // namespace

void exclusive_scan_u32(Ctx &c, const uint32_t *in, uint64_t n, uint32_t *out, uint32_t *d_total) { scan_impl<uint32_t, uint32_t>(c, in, n, out, d_total); }
void exclusive_scan_u8_to_u32(Ctx &c, const uint8_t *in, uint64_t n, uint32_t *out, uint32_t *d_total) { scan_impl<uint8_t, uint32_t>(c, in, n, out, d_total); }
void exclusive_scan_u32_to_u64(Ctx &c, const uint32_t *in, uint64_t n, uint64_t *out, uint64_t *d_total) { scan_impl<uint32_t, uint64_t>(c, in, n, out, d_total); }

namespace {
template <int BITS>
void radix_pass(Ctx &c, const uint32_t *kin, const uint32_t *vin, uint32_t *kout, uint32_t *vout, uint64_t n, int shift, uint32_t *hist, unsigned tiles) {
    constexpr int BINS = 1 << BITS;
    constexpr size_t smem = sizeof(uint32_t) * RS_WARPS * BINS;
    if (smem > 48 * 1024) HERA_CUDA(cudaFuncSetAttribute(rs_scatter_kernel<BITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    HERA_LAUNCH(c, rs_hist_kernel<BITS>, tiles, RS_THREADS, 0, kin, n, shift, hist);
    exclusive_scan_u32(c, hist, static_cast<uint64_t>(BINS) * tiles, hist, nullptr);
    HERA_LAUNCH(c, rs_scatter_kernel<BITS>, tiles, RS_THREADS, smem, kin, vin, kout, vout, n, shift, hist);
}
}  // namespace

void radix_sort_pairs(Ctx &c, uint32_t *keys, uint32_t *vals, uint32_t *keys_alt, uint32_t *vals_alt, uint64_t n, int bits) {
    if (n == 0) return;
    const unsigned tiles = div_up(n, RS_TILE);
    int passes = std::max(1, (bits + 10) / 11);              // digits of up to 11 bits
    int digit = std::max(8, (bits + passes - 1) / passes);   // 8 .. 11
    if (c.opt_radix_digit_bits >= 8 && c.opt_radix_digit_bits <= 11) {   // (tests: small inputs would never meet the wide digits)
        digit = c.opt_radix_digit_bits;
        passes = std::max(1, (bits + digit - 1) / digit);
    }
    TempPool tmp(c.stream);
    uint32_t *hist = tmp.get<uint32_t>((static_cast<size_t>(1) << digit) * tiles);
    uint32_t *kin = keys, *vin = vals, *kout = keys_alt, *vout = vals_alt;
    for (int p = 0; p < passes; p++) {
        const int shift = p * digit;
        switch (digit) {
            case 8: radix_pass<8>(c, kin, vin, kout, vout, n, shift, hist, tiles); break;
            case 9: radix_pass<9>(c, kin, vin, kout, vout, n, shift, hist, tiles); break;
            case 10: radix_pass<10>(c, kin, vin, kout, vout, n, shift, hist, tiles); break;
            default: radix_pass<11>(c, kin, vin, kout, vout, n, shift, hist, tiles); break;
        }
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
    if (passes & 1) {  // result sits in the alt buffers
        HERA_CUDA(cudaMemcpyAsync(keys, keys_alt, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, c.stream));
        HERA_CUDA(cudaMemcpyAsync(vals, vals_alt, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, c.stream));
    }
}

}  // namespace hera
