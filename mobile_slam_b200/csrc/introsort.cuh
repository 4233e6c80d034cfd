// libstdc++ std::sort, restated so the reference's UNSTABLE tie order can be reproduced on the device.
//
// FeatureTracker::setMask sorts (track_cnt, point, id) triples with std::sort and a comparator that
// looks at track_cnt only (/root/reference/src/frontend/feature_tracker.cpp:70-73); which of two
// equally old features survives the min-distance greedy therefore depends on the exact element moves
// of GCC's introsort (bits/stl_algo.h: __introsort_loop, threshold 16, median-of-3 to first,
// __unguarded_partition, heap-sort fallback at depth 2*floor(log2 n), __final_insertion_sort).
// This header re-implements that algorithm step for step on an index array `v` with keys `key[v[i]]`;
// tests/test_host_logic.py checks it against the real std::sort (oracle/stdsort.cpp) on tie-heavy input.
#pragma once
#ifndef __CUDACC__
#ifndef __host__
#define __host__
#define __device__
#endif
#endif

namespace vs {

struct DescByKey {
    const int* key;
    __host__ __device__ bool operator()(int a, int b) const { return key[a] > key[b]; }
};

template <class Comp>
__host__ __device__ inline void iso_swap(int* v, int a, int b, Comp) {
    const int t = v[a];
    v[a] = v[b];
    v[b] = t;
}

template <class Comp>
__host__ __device__ inline void iso_push_heap(int* v, int first, int hole, int top, int value, Comp comp) {
    int parent = (hole - 1) / 2;
    while (hole > top && comp(v[first + parent], value)) {
        v[first + hole] = v[first + parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    v[first + hole] = value;
}

template <class Comp>
__host__ __device__ inline void iso_adjust_heap(int* v, int first, int hole, int len, int value, Comp comp) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (comp(v[first + child], v[first + child - 1])) child--;
        v[first + hole] = v[first + child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        v[first + hole] = v[first + child - 1];
        hole = child - 1;
    }
    iso_push_heap(v, first, hole, top, value, comp);
}

template <class Comp>
__host__ __device__ inline void iso_heap_sort(int* v, int first, int last, Comp comp) {
    // __partial_sort(first, last, last): __heap_select == make_heap, then __sort_heap
    const int len = last - first;
    if (len >= 2) {
        int parent = (len - 2) / 2;
        while (true) {
            const int value = v[first + parent];
            iso_adjust_heap(v, first, parent, len, value, comp);
            if (parent == 0) break;
            parent--;
        }
    }
    int l = last;
    while (l - first > 1) {
        --l;
        const int value = v[l];
        v[l] = v[first];
        iso_adjust_heap(v, first, 0, l - first, value, comp);
    }
}

template <class Comp>
__host__ __device__ inline void iso_move_median_to_first(int* v, int result, int a, int b, int c, Comp comp) {
    if (comp(v[a], v[b])) {
        if (comp(v[b], v[c])) iso_swap(v, result, b, comp);
        else if (comp(v[a], v[c])) iso_swap(v, result, c, comp);
        else iso_swap(v, result, a, comp);
    } else if (comp(v[a], v[c])) iso_swap(v, result, a, comp);
    else if (comp(v[b], v[c])) iso_swap(v, result, c, comp);
    else iso_swap(v, result, b, comp);
}

template <class Comp>
__host__ __device__ inline int iso_unguarded_partition(int* v, int first, int last, int pivot, Comp comp) {
    while (true) {
        while (comp(v[first], v[pivot])) ++first;
        --last;
        while (comp(v[pivot], v[last])) --last;
        if (!(first < last)) return first;
        iso_swap(v, first, last, comp);
        ++first;
    }
}

template <class Comp>
__host__ __device__ inline void iso_unguarded_linear_insert(int* v, int last, Comp comp) {
    const int val = v[last];
    int next = last - 1;
    while (comp(val, v[next])) {
        v[last] = v[next];
        last = next;
        --next;
    }
    v[last] = val;
}

template <class Comp>
__host__ __device__ inline void iso_insertion_sort(int* v, int first, int last, Comp comp) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
        if (comp(v[i], v[first])) {
            const int val = v[i];
            for (int j = i; j > first; --j) v[j] = v[j - 1];   // move_backward(first, i, i + 1)
            v[first] = val;
        } else {
            iso_unguarded_linear_insert(v, i, comp);
        }
    }
}

// std::sort(v, v + n, comp) — same element moves as libstdc++.
template <class Comp>
__host__ __device__ inline void std_sort_indices(int* v, int n, Comp comp) {
    if (n <= 0) return;
    // __introsort_loop with an explicit stack (the recursion is on the right part)
    int lg = 0;
    for (int t = n; t > 1; t >>= 1) ++lg;
    int stack_first[64], stack_last[64], stack_depth[64];
    int sp = 0;
    stack_first[0] = 0; stack_last[0] = n; stack_depth[0] = 2 * lg; sp = 1;
    while (sp > 0) {
        --sp;
        int first = stack_first[sp], last = stack_last[sp], depth = stack_depth[sp];
        // NOTE: libstdc++ recurses on [cut,last) *before* continuing with [first,cut); the two parts
        // are disjoint, so processing order does not change the result.
        while (last - first > 16) {
            if (depth == 0) {
                iso_heap_sort(v, first, last, comp);
                last = first;
                break;
            }
            --depth;
            const int mid = first + (last - first) / 2;
            iso_move_median_to_first(v, first, first + 1, mid, last - 1, comp);
            const int cut = iso_unguarded_partition(v, first + 1, last, first, comp);
            stack_first[sp] = cut; stack_last[sp] = last; stack_depth[sp] = depth; ++sp;
            last = cut;
        }
    }
    // __final_insertion_sort
    if (n > 16) {
        iso_insertion_sort(v, 0, 16, comp);
        for (int i = 16; i != n; ++i) iso_unguarded_linear_insert(v, i, comp);
    } else {
        iso_insertion_sort(v, 0, n, comp);
    }
}


#ifdef __CUDACC__
// ---------------------------------------------------------------------------------------------
// Warp-cooperative version of the __introsort_loop phase of std_sort_indices (same element moves).
//
// __unguarded_partition only ever looks at each element's ORIGINAL key: the left cursor stops at
// the k-th element (ascending) with !comp(e, pivot), the right cursor at the k-th element
// (descending) with !comp(pivot, e), and pair k is swapped while left < right.  Both stop lists are
// built with ballots, the first crossing pair K is found, pairs 0..K-1 are swapped in parallel, and
//   cut = min(Lpos[K], Rpos[K-1])
// (the left cursor cannot pass the slot the previous swap filled with an element it stops at).
// After this phase libstdc++ runs __final_insertion_sort, which is a STABLE sort of what the loop
// left behind; the caller does that with a parallel sort on (key, position).
// Call with all 32 lanes of one warp; v / key / lpos / rpos live in shared memory.
// ---------------------------------------------------------------------------------------------
// one __unguarded_partition_pivot step of [first, last) on a warp; returns the cut.  lpos / rpos: scratch of at least
// (last - first) ints each, private to this call.
__device__ inline int warp_partition_desc(int* v, const int* key, int first, int last, int* lpos, int* rpos, int lane) {
    const unsigned full = 0xffffffffu;
    DescByKey comp{key};
    const int mid = first + (last - first) / 2;
    if (lane == 0) iso_move_median_to_first(v, first, first + 1, mid, last - 1, comp);
    __syncwarp();
    const int pk = key[v[first]];
    const int lo = first + 1;
    // left stops: key <= pivot key, ascending positions
    int nL = 0;
    for (int i0 = lo; i0 < last; i0 += 32) {
        const int i = i0 + lane;
        const bool a = i < last && key[v[i]] <= pk;
        const unsigned bal = __ballot_sync(full, a);
        if (a) lpos[nL + __popc(bal & ((1u << lane) - 1u))] = i;
        nL += __popc(bal);
    }
    // right stops: key >= pivot key, descending positions
    int nR = 0;
    for (int i0 = last - 1; i0 >= lo; i0 -= 32) {
        const int i = i0 - lane;
        const bool b = i >= lo && key[v[i]] >= pk;
        const unsigned bal = __ballot_sync(full, b);
        if (b) rpos[nR + __popc(bal & ((1u << lane) - 1u))] = i;
        nR += __popc(bal);
    }
    __syncwarp();
    const int nmin = nL < nR ? nL : nR;
    int K = nmin;
    for (int k0 = 0; k0 < nmin; k0 += 32) {
        const int k = k0 + lane;
        const bool cross = k < nmin && lpos[k] >= rpos[k];
        const unsigned bal = __ballot_sync(full, cross);
        if (bal) {
            K = k0 + __ffs(bal) - 1;
            break;
        }
    }
    int cut = last;
    if (K < nL) cut = lpos[K];
    if (K >= 1 && rpos[K - 1] < cut) cut = rpos[K - 1];
    for (int k = lane; k < K; k += 32) {
        const int a = lpos[k], b = rpos[k];
        const int t = v[a];
        v[a] = v[b];
        v[b] = t;
    }
    __syncwarp();
    return cut;
}

// __introsort_loop on the whole CTA (blockDim.x = 32 * nwarps): the ranges of one recursion level are disjoint, so
// the warps partition them concurrently, level by level (the element moves inside a range do not depend on when
// the range is processed, so the result equals the serial recursion).  queue_a / queue_b: scratch of n / 8 + 2 ints
// each; lpos / rpos: scratch of n ints each (a range uses its own [first, last) slice).  Must be called by all
// threads; v / key / scratch live in shared memory.
__device__ inline void block_introsort_loop_desc(int* v, const int* key, int n, int* lpos, int* rpos, int* queue_a,
                                                 int* queue_b, int* s_counts /* [2] shared */, int tid, int nwarps) {
    if (n <= 16) return;                      // uniform
    const int lane = tid & 31, warp = tid >> 5;
    int lg = 0;
    for (int t = n; t > 1; t >>= 1) ++lg;
    // entry = first | last << 12 | depth << 25   (first < 4096, last <= 4096, depth <= 24)
    if (tid == 0) {
        queue_a[0] = 0 | (n << 12) | ((2 * lg) << 25);
        s_counts[0] = 1;
        s_counts[1] = 0;
    }
    __syncthreads();
    DescByKey comp{key};
    int* qa = queue_a;
    int* qb = queue_b;
    for (;;) {
        const int na = s_counts[0];
        if (na == 0) break;
        for (int r = warp; r < na; r += nwarps) {
            const int e = qa[r];
            const int first = e & 0xfff, last = (e >> 12) & 0x1fff;
            int depth = e >> 25;
            if (depth == 0) {
                if (lane == 0) iso_heap_sort(v, first, last, comp);
                __syncwarp();
                continue;
            }
            --depth;
            const int cut = warp_partition_desc(v, key, first, last, lpos + first, rpos + first, lane);
            if (lane == 0) {
                if (cut - first > 16) qb[atomicAdd(&s_counts[1], 1)] = first | (cut << 12) | (depth << 25);
                if (last - cut > 16) qb[atomicAdd(&s_counts[1], 1)] = cut | (last << 12) | (depth << 25);
            }
        }
        __syncthreads();
        if (tid == 0) {
            s_counts[0] = s_counts[1];
            s_counts[1] = 0;
        }
        int* t = qa;
        qa = qb;
        qb = t;
        __syncthreads();
    }
}
#endif  // __CUDACC__

}  // namespace vs
