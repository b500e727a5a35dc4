// The reference sorts each anchor pair's paths with std::sort by length (src/PathManager.cpp:
// 510-513) and the groups with consensus by consensus length (src/Main.cpp:312-315).  std::sort
// is unstable: which of several equal-length paths comes first -- and therefore which path
// calculateConsensusPath returns (src/PathGroup.cpp:71-78) -- is a property of the algorithm
// libstdc++ runs.  For bit-exact consensus paths the device executes the same algorithm
// (GCC 13 introsort: median-of-three quicksort down to ranges of 16, heap sort once the depth
// budget 2*floor(log2 n) is used up, final insertion sort), one thread per anchor pair, written
// here with an explicit range stack instead of recursion.  Sub-ranges are disjoint, so the
// order in which they are processed does not change the result.
#pragma once

#include <cstdint>

namespace hera {

template <typename T, typename Less>
__host__ __device__ inline void is_sift(T *a, long hole, long len, T value, Less less) {
    const long top = hole;
    long child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (less(a[child], a[child - 1])) child--;
        a[hole] = a[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        a[hole] = a[child - 1];
        hole = child - 1;
    }
    long parent = (hole - 1) / 2;
    while (hole > top && less(a[parent], value)) {
        a[hole] = a[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    a[hole] = value;
}

template <typename T, typename Less>
__host__ __device__ inline void is_heap_sort(T *a, long len, Less less) {
    if (len >= 2) {
        for (long parent = (len - 2) / 2;; parent--) {
            T v = a[parent];
            is_sift(a, parent, len, v, less);
            if (parent == 0) break;
        }
    }
    for (long last = len; last > 1;) {
        --last;
        T v = a[last];
        a[last] = a[0];
        is_sift(a, 0L, last, v, less);
    }
}

template <typename T>
__host__ __device__ inline void is_swap(T &x, T &y) { T t = x; x = y; y = t; }

template <typename T, typename Less>
__host__ __device__ inline long is_partition(T *a, long first, long last, Less less) {
    const long mid = first + (last - first) / 2;
    const long x = first + 1, y = mid, z = last - 1;
    // median of (x, y, z) goes to `first`
    if (less(a[x], a[y])) {
        if (less(a[y], a[z])) is_swap(a[first], a[y]);
        else if (less(a[x], a[z])) is_swap(a[first], a[z]);
        else is_swap(a[first], a[x]);
    } else if (less(a[x], a[z])) is_swap(a[first], a[x]);
    else if (less(a[y], a[z])) is_swap(a[first], a[z]);
    else is_swap(a[first], a[y]);
    long lo = first + 1, hi = last;
    for (;;) {
        while (less(a[lo], a[first])) ++lo;
        --hi;
        while (less(a[first], a[hi])) --hi;
        if (!(lo < hi)) return lo;
        is_swap(a[lo], a[hi]);
        ++lo;
    }
}

template <typename T, typename Less>
__host__ __device__ inline void is_linear_insert(T *a, long pos, Less less) {
    T v = a[pos];
    long prev = pos - 1;
    while (less(v, a[prev])) {
        a[pos] = a[prev];
        pos = prev;
        --prev;
    }
    a[pos] = v;
}

template <typename T, typename Less>
__host__ __device__ inline void is_insertion(T *a, long first, long last, Less less) {
    if (first == last) return;
    for (long i = first + 1; i != last; ++i) {
        if (less(a[i], a[first])) {
            T v = a[i];
            for (long p = i; p != first; --p) a[p] = a[p - 1];
            a[first] = v;
        } else {
            is_linear_insert(a, i, less);
        }
    }
}

// Same element moves as std::sort(a, a + n, less) of libstdc++ 13.
template <typename T, typename Less>
__host__ __device__ inline void libstdcxx_sort(T *a, long n, Less less) {
    if (n <= 0) return;
    int lg = 0;
    for (long m = n; m > 1; m >>= 1) lg++;
    struct Range { long first, last; int depth; };
    Range stack[160];  // every pop pushes at most one range and the depth budget is <= 2 * 63
    int top = 0;
    stack[top++] = Range{0, n, 2 * lg};
    while (top > 0) {
        Range r = stack[--top];
        while (r.last - r.first > 16) {
            if (r.depth == 0) {
                is_heap_sort(a + r.first, r.last - r.first, less);
                break;
            }
            --r.depth;
            const long cut = is_partition(a, r.first, r.last, less);
            if (top < 160) stack[top++] = Range{cut, r.last, r.depth};
            r.last = cut;
        }
    }
    if (n > 16) {
        is_insertion(a, 0L, 16L, less);
        for (long i = 16; i != n; ++i) is_linear_insert(a, i, less);
    } else {
        is_insertion(a, 0L, n, less);
    }
}

}  // namespace hera
