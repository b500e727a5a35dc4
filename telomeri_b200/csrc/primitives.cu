// Host launchers of the scan and radix-sort kernels in primitives.cuh.
#include "ctx.cuh"
#include "primitives.cuh"

namespace hera {

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
