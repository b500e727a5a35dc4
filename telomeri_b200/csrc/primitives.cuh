// Device-wide building blocks written for this library: exclusive scan and a stable LSD radix
// sort of (u32 key, u32 value) pairs with shared-memory staged ranking.  All launches go through
// the context's stream and are counted.
#pragma once

#include "common.cuh"

namespace hera {

struct Ctx;  // defined in ctx.cuh

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ uint32_t warp_id() { return threadIdx.x >> 5; }

// Block-wide exclusive scan of one value per thread (blockDim multiple of 32, <= 1024).
// Returns the exclusive prefix; *total gets the block sum.  smem: 33 entries of T.
template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T *smem, T *total) {
    const uint32_t lane = lane_id(), warp = warp_id();
    T incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T up = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= static_cast<uint32_t>(d)) incl += up;
    }
    if (lane == 31) smem[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const uint32_t nw = blockDim.x >> 5;
        T w = lane < nw ? smem[lane] : T(0);
        T wi = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            T up = __shfl_up_sync(0xffffffffu, wi, d);
            if (lane >= static_cast<uint32_t>(d)) wi += up;
        }
        if (lane < nw) smem[lane] = wi - w;
        if (lane == nw - 1) smem[32] = wi;
    }
    __syncthreads();
    T res = smem[warp] + incl - v;
    if (total) *total = smem[32];
    __syncthreads();
    return res;
}

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_sums_kernel(const TIn *__restrict__ in, uint64_t n, TOut *__restrict__ tile_sums) {
    __shared__ TOut smem[33];
    const uint64_t base = static_cast<uint64_t>(blockIdx.x) * SCAN_TILE;
    TOut s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        uint64_t i = base + static_cast<uint64_t>(k) * SCAN_THREADS + threadIdx.x;
        if (i < n) s += static_cast<TOut>(in[i]);
    }
    TOut total;
    block_exclusive_scan<TOut>(s, smem, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// out[i] = tile_base[tile] + exclusive prefix inside the tile.  Each thread owns SCAN_ITEMS
// consecutive elements so the in-thread order is the array order.
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(const TIn *__restrict__ in, uint64_t n, const TOut *__restrict__ tile_base,
                                                                  TOut *__restrict__ out, TOut *__restrict__ grand_total) {
    __shared__ TOut smem[33];
    const uint64_t base = static_cast<uint64_t>(blockIdx.x) * SCAN_TILE + static_cast<uint64_t>(threadIdx.x) * SCAN_ITEMS;
    TOut v[SCAN_ITEMS];
    TOut s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = (base + k < n) ? static_cast<TOut>(in[base + k]) : TOut(0);
        s += v[k];
    }
    TOut total;
    TOut excl = block_exclusive_scan<TOut>(s, smem, &total);
    TOut run = excl + (tile_base ? tile_base[blockIdx.x] : TOut(0));
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k < n) out[base + k] = run;
        run += v[k];
    }
    if (grand_total && blockIdx.x == gridDim.x - 1 && threadIdx.x == SCAN_THREADS - 1) *grand_total = run;
}

// ---- radix sort -------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;                 // chunks of RS_THREADS elements per tile
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;

}  // namespace hera
