// Shared plumbing of the hera_b200 library: context, growable device/pinned buffers, error
// handling, launch accounting.  Kernels live in the other .cu files of this directory.
#pragma once

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../include/hera_b200.h"

namespace hera {

struct Error {
    int code;
    std::string msg;
};

#define HERA_CUDA(expr)                                                                                   \
    do {                                                                                                  \
        cudaError_t err__ = (expr);                                                                       \
        if (err__ != cudaSuccess)                                                                         \
            throw ::hera::Error{2, std::string(#expr) + ": " + cudaGetErrorString(err__) + " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")"}; \
    } while (0)

#define HERA_REQUIRE(cond, code, text)                 \
    do {                                               \
        if (!(cond)) throw ::hera::Error{(code), (text)}; \
    } while (0)

// Stream-ordered temporaries, freed (stream-ordered) when the scope ends, exception or not.
struct TempPool {
    cudaStream_t s;
    std::vector<void *> ptrs;
    explicit TempPool(cudaStream_t st) : s(st) {}
    ~TempPool() { for (void *p : ptrs) if (p) cudaFreeAsync(p, s); }
    TempPool(const TempPool &) = delete;
    TempPool &operator=(const TempPool &) = delete;
    template <typename T>
    T *get(uint64_t n) {
        void *p = nullptr;
        HERA_CUDA(cudaMallocAsync(&p, std::max<uint64_t>(n, 1) * sizeof(T), s));
        ptrs.push_back(p);
        return static_cast<T *>(p);
    }
    void release(void *p) {   // before the scope ends (stream-ordered like the rest)
        for (auto &q : ptrs)
            if (q == p && p) { cudaFreeAsync(p, s); q = nullptr; }
    }
    template <typename T>
    T *zeroed(uint64_t n) {
        T *p = get<T>(n);
        HERA_CUDA(cudaMemsetAsync(p, 0, std::max<uint64_t>(n, 1) * sizeof(T), s));
        return p;
    }
};

// N CUDA events, destroyed when the scope ends, exception or not.
template <int N>
struct Events {
    cudaEvent_t e[N] = {};
    explicit Events(unsigned flags = cudaEventDefault) {
        for (int i = 0; i < N; i++) {
            cudaError_t err = cudaEventCreateWithFlags(&e[i], flags);
            if (err != cudaSuccess) {
                destroy();
                throw ::hera::Error{2, std::string("cudaEventCreateWithFlags: ") + cudaGetErrorString(err)};
            }
        }
    }
    ~Events() { destroy(); }
    Events(const Events &) = delete;
    Events &operator=(const Events &) = delete;
    cudaEvent_t operator[](int i) const { return e[i]; }
    void destroy() {
        for (int i = 0; i < N; i++)
            if (e[i]) { cudaEventDestroy(e[i]); e[i] = nullptr; }
    }
};

// Growable device array.  grow() keeps the contents when asked to.
template <typename T>
struct DBuf {
    T *p = nullptr;
    size_t cap = 0;
    bool borrowed = false;   // p belongs to somebody else (a buffer the exchange shares between the ranks)
    ~DBuf() { release(); }
    DBuf() = default;
    DBuf(const DBuf &) = delete;
    DBuf &operator=(const DBuf &) = delete;
    void release() {
        if (p && !borrowed) cudaFree(p);
        p = nullptr;
        cap = 0;
        borrowed = false;
    }
    void borrow(T *ptr, size_t n) {
        release();
        p = ptr;
        cap = n;
        borrowed = true;
    }
    void reserve(size_t n, bool keep = false, size_t keep_n = 0, cudaStream_t s = 0) {
        if (n <= cap) return;
        if (borrowed) throw ::hera::Error{2, "internal: a borrowed buffer cannot grow"};
        size_t ncap = n + n / 4 + 64;
        T *np = nullptr;
        HERA_CUDA(cudaMalloc(&np, ncap * sizeof(T)));
        if (keep && p && keep_n) HERA_CUDA(cudaMemcpyAsync(np, p, keep_n * sizeof(T), cudaMemcpyDeviceToDevice, s));
        if (p) {
            HERA_CUDA(cudaStreamSynchronize(s));
            cudaFree(p);
        }
        p = np;
        cap = ncap;
    }
};

// Growable pinned host array (export views point into these).
template <typename T>
struct HBuf {
    T *p = nullptr;
    size_t cap = 0;
    ~HBuf() { if (p) cudaFreeHost(p); }
    HBuf() = default;
    HBuf(const HBuf &) = delete;
    HBuf &operator=(const HBuf &) = delete;
    void reserve(size_t n) {
        if (n <= cap) return;
        if (p) cudaFreeHost(p);
        cap = n + n / 4 + 64;
        HERA_CUDA(cudaMallocHost(&p, cap * sizeof(T)));
    }
};

struct Launches {
    uint64_t count = 0;
};

// Launch helper: counts the launch and checks the launch status.
#define HERA_LAUNCH(ctx, kernel, grid, block, smem, ...)                      \
    do {                                                                      \
        kernel<<<(grid), (block), (smem), (ctx).stream>>>(__VA_ARGS__);       \
        (ctx).launches.count++;                                               \
        HERA_CUDA(cudaGetLastError());                                        \
    } while (0)

inline unsigned div_up(uint64_t a, uint64_t b) { return static_cast<unsigned>((a + b - 1) / b); }

}  // namespace hera
