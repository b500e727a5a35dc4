// TEST INFRASTRUCTURE.  Restatement of the introsort that libstdc++ (GCC 13, bits/stl_algo.h +
// bits/stl_heap.h) runs for std::sort: median-of-three quicksort while a range is longer than
// 16, heap sort once the depth budget 2*floor(log2 n) is spent, one final insertion sort.
// std::sort is unstable, so which of several equal-length paths ends up first -- and with it
// the reference's consensus path (src/PathGroup.cpp:71-78) -- is decided by exactly these
// element moves.  tests/test_oracle_sort.py checks permutation equality against the real
// std::sort on tie-heavy and adversarial inputs.
#ifndef HERA_ORACLE_SORT_HPP
#define HERA_ORACLE_SORT_HPP

#include <cstddef>
#include <utility>

namespace hera_oracle {
namespace sort_detail {

template <typename T, typename Less>
void sift_down(T *a, ptrdiff_t hole, ptrdiff_t len, T value, Less less) {
    const ptrdiff_t top = hole;
    ptrdiff_t child = hole;
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
    ptrdiff_t parent = (hole - 1) / 2;
    while (hole > top && less(a[parent], value)) {
        a[hole] = a[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    a[hole] = value;
}

template <typename T, typename Less>
void heap_sort(T *a, ptrdiff_t len, Less less) {
    if (len >= 2) {
        for (ptrdiff_t parent = (len - 2) / 2;; parent--) {
            T v = a[parent];
            sift_down(a, parent, len, v, less);
            if (parent == 0) break;
        }
    }
    for (ptrdiff_t last = len; last > 1;) {
        --last;
        T v = a[last];
        a[last] = a[0];
        sift_down(a, ptrdiff_t(0), last, v, less);
    }
}

template <typename T, typename Less>
void median_to_first(T *result, T *x, T *y, T *z, Less less) {
    if (less(*x, *y)) {
        if (less(*y, *z)) std::swap(*result, *y);
        else if (less(*x, *z)) std::swap(*result, *z);
        else std::swap(*result, *x);
    } else if (less(*x, *z)) std::swap(*result, *x);
    else if (less(*y, *z)) std::swap(*result, *z);
    else std::swap(*result, *y);
}

template <typename T, typename Less>
T *partition_around_first(T *first, T *last, Less less) {
    T *mid = first + (last - first) / 2;
    median_to_first(first, first + 1, mid, last - 1, less);
    T *pivot = first;
    T *lo = first + 1;
    T *hi = last;
    for (;;) {
        while (less(*lo, *pivot)) ++lo;
        --hi;
        while (less(*pivot, *hi)) --hi;
        if (!(lo < hi)) return lo;
        std::swap(*lo, *hi);
        ++lo;
    }
}

template <typename T, typename Less>
void quick_phase(T *first, T *last, int depth, Less less) {
    while (last - first > 16) {
        if (depth == 0) {
            heap_sort(first, last - first, less);
            return;
        }
        --depth;
        T *cut = partition_around_first(first, last, less);
        quick_phase(cut, last, depth, less);
        last = cut;
    }
}

template <typename T, typename Less>
void linear_insert_unguarded(T *pos, Less less) {
    T v = *pos;
    T *prev = pos - 1;
    while (less(v, *prev)) {
        *pos = *prev;
        pos = prev;
        --prev;
    }
    *pos = v;
}

template <typename T, typename Less>
void insertion_guarded(T *first, T *last, Less less) {
    if (first == last) return;
    for (T *i = first + 1; i != last; ++i) {
        if (less(*i, *first)) {
            T v = *i;
            for (T *p = i; p != first; --p) *p = *(p - 1);
            *first = v;
        } else {
            linear_insert_unguarded(i, less);
        }
    }
}

}  // namespace sort_detail

template <typename T, typename Less>
void libstdcxx_sort(T *first, T *last, Less less) {
    if (first == last) return;
    ptrdiff_t n = last - first;
    int lg = 0;
    for (ptrdiff_t m = n; m > 1; m >>= 1) lg++;
    sort_detail::quick_phase(first, last, 2 * lg, less);
    if (n > 16) {
        sort_detail::insertion_guarded(first, first + 16, less);
        for (T *i = first + 16; i != last; ++i) sort_detail::linear_insert_unguarded(i, less);
    } else {
        sort_detail::insertion_guarded(first, last, less);
    }
}

}  // namespace hera_oracle
#endif
