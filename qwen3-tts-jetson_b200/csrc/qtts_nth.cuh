// qtts_nth.cuh — the selection algorithm behind the reference's codebook-0 top-k threshold.
//
// src/tts_transformer.cpp:2458-2466 calls
//     std::nth_element(scored_idx.begin(), scored_idx.begin() + top_k, scored_idx.end(), value-descending)
// and then reads scored_vals[scored_idx[top_k - 1]] as the threshold.  nth_element only fixes position
// top_k; position top_k - 1 holds SOME member of the top-k set, which one depends on the library's
// algorithm.  The reference is built with libstdc++ (GCC), whose std::nth_element is an introselect:
// median-of-three pivot (first + 1, middle, last - 1) moved to the front, unguarded Hoare partition,
// recursion into the side that holds nth until at most 3 elements remain, insertion sort of those,
// heap-select fallback after 2 * floor(log2(n)) partition rounds (bits/stl_algo.h: __introselect,
// __unguarded_partition_pivot, __move_median_to_first, __heap_select; bits/stl_heap.h).  This file
// restates that algorithm for an index array with the comparator "value of a > value of b", so the
// device picks the threshold the reference picks (measured on the random-init fixtures: the reference
// keeps as few as 4 of the nominal 50 candidates).  It compiles for host and device: tests/test_nth.py
// checks it on the CPU against std::nth_element itself (oracle/nth_check.cpp).
#pragma once
#ifndef __CUDACC__
#define QTTS_HD
#else
#define QTTS_HD __host__ __device__ __forceinline__
#endif

namespace qtts { namespace nth {

template <typename I> struct Desc {
    const float * v;
    QTTS_HD bool operator()(I a, I b) const { return v[a] > v[b]; }
};

template <typename I> QTTS_HD void swp(I * a, I * b) { const I t = *a; *a = *b; *b = t; }

template <typename I, typename C>
QTTS_HD void median_to_first(I * result, I * a, I * b, I * c, C comp) {
    if (comp(*a, *b)) {
        if (comp(*b, *c)) swp(result, b);
        else if (comp(*a, *c)) swp(result, c);
        else swp(result, a);
    } else if (comp(*a, *c)) swp(result, a);
    else if (comp(*b, *c)) swp(result, c);
    else swp(result, b);
}

template <typename I, typename C>
QTTS_HD I * unguarded_partition(I * first, I * last, I * pivot, C comp) {
    for (;;) {
        while (comp(*first, *pivot)) ++first;
        --last;
        while (comp(*pivot, *last)) --last;
        if (!(first < last)) return first;
        swp(first, last);
        ++first;
    }
}

template <typename I, typename C>
QTTS_HD void push_heap(I * first, int hole, int top, I value, C comp) {
    int parent = (hole - 1) / 2;
    while (hole > top && comp(first[parent], value)) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}

template <typename I, typename C>
QTTS_HD void adjust_heap(I * first, int hole, int len, I value, C comp) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (comp(first[child], first[child - 1])) child--;
        first[hole] = first[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        first[hole] = first[child - 1];
        hole = child - 1;
    }
    push_heap(first, hole, top, value, comp);
}

template <typename I, typename C>
QTTS_HD void heap_select(I * first, I * middle, I * last, C comp) {
    const int len = (int) (middle - first);
    if (len >= 2) {
        for (int parent = (len - 2) / 2;; --parent) {
            adjust_heap(first, parent, len, first[parent], comp);
            if (parent == 0) break;
        }
    }
    for (I * i = middle; i < last; ++i)
        if (comp(*i, *first)) {
            const I value = *i;
            *i = *first;
            adjust_heap(first, 0, len, value, comp);
        }
}

template <typename I, typename C>
QTTS_HD void insertion_sort(I * first, I * last, C comp) {
    if (first == last) return;
    for (I * i = first + 1; i != last; ++i) {
        const I val = *i;
        if (comp(val, *first)) {
            for (I * j = i; j != first; --j) *j = *(j - 1);
            *first = val;
        } else {
            I * j = i;
            while (comp(val, *(j - 1))) { *j = *(j - 1); --j; }
            *j = val;
        }
    }
}

// idx[0..n): a permutation of element ids; vals indexed by id.  After the call idx[nth] is the
// element a descending sort would put there, ids before it have values >= it, ids after <= it —
// arranged exactly as libstdc++'s std::nth_element leaves them.
template <typename I>
QTTS_HD void nth_element_desc(I * idx, int nth_pos, int n, const float * vals) {
    if (n <= 0 || nth_pos >= n) return;
    const Desc<I> comp{vals};
    I * first = idx, * last = idx + n, * nth = idx + nth_pos;
    int depth = 0;
    for (int m = n; m > 1; m >>= 1) ++depth;      // floor(log2(n))
    depth *= 2;
    while (last - first > 3) {
        if (depth == 0) {
            heap_select(first, nth + 1, last, comp);
            swp(first, nth);
            return;
        }
        --depth;
        I * mid = first + (last - first) / 2;
        median_to_first(first, first + 1, mid, last - 1, comp);
        I * cut = unguarded_partition(first + 1, last, first, comp);
        if (cut <= nth) first = cut; else last = cut;
    }
    insertion_sort(first, last, comp);
}

}}  // namespace qtts::nth
