// Reproduces the PERMUTATION that libstdc++'s std::sort produces (GCC 13 bits/stl_algo.h,
// bits/stl_heap.h: median-of-3 introsort down to runs of 16, heap-sort fallback at depth
// 2*floor(log2 n), then the final insertion sort).  The reference sorts the quadtree's
// (key count, UL.x) pairs with std::sort and a comparator that ties
// (Features/ORBextractor.cpp:684-696, 912-913); the order of tied elements decides which nodes are
// split before DistributeOctTree's `break`, so a device sort "by key" is not parity-safe -- the
// algorithm itself has to be replayed.  Host+device so the CPU suite can check it against the
// real std::sort (tests/test_stdsort_emul.py).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define ORBX_HD __host__ __device__ __forceinline__
#else
#define ORBX_HD inline
#endif

namespace orbx {

struct SortElem {
  unsigned long long key;  // compared with <
  unsigned int id;         // payload
  unsigned int pad;
};

namespace ss {

struct KeyLess {
  ORBX_HD bool operator()(const SortElem& a, const SortElem& b) const { return a.key < b.key; }
};
template <class T>
ORBX_HD void swp(T& a, T& b) { T t = a; a = b; b = t; }

template <class T, class Less>
ORBX_HD void push_heap(T* f, long hole, long top, T v, Less lt) {
  long parent = (hole - 1) / 2;
  while (hole > top && lt(f[parent], v)) {
    f[hole] = f[parent];
    hole = parent;
    parent = (hole - 1) / 2;
  }
  f[hole] = v;
}

template <class T, class Less>
ORBX_HD void adjust_heap(T* f, long hole, long len, T v, Less lt) {
  const long top = hole;
  long child = hole;
  while (child < (len - 1) / 2) {
    child = 2 * (child + 1);
    if (lt(f[child], f[child - 1])) child--;
    f[hole] = f[child];
    hole = child;
  }
  if ((len & 1) == 0 && child == (len - 2) / 2) {
    child = 2 * (child + 1);
    f[hole] = f[child - 1];
    hole = child - 1;
  }
  push_heap(f, hole, top, v, lt);
}

// std::__partial_sort(first, last, last): make_heap + sort_heap (the heap_select loop is empty)
template <class T, class Less>
ORBX_HD void heap_sort(T* f, long len, Less lt) {
  if (len >= 2) {
    long parent = (len - 2) / 2;
    for (;;) {
      T v = f[parent];
      adjust_heap(f, parent, len, v, lt);
      if (parent == 0) break;
      parent--;
    }
  }
  long last = len;
  while (last > 1) {
    --last;
    T v = f[last];
    f[last] = f[0];
    adjust_heap(f, 0, last, v, lt);
  }
}

template <class T, class Less>
ORBX_HD void unguarded_linear_insert(T* f, long last, Less lt) {
  T v = f[last];
  long next = last - 1;
  while (lt(v, f[next])) {
    f[last] = f[next];
    last = next;
    --next;
  }
  f[last] = v;
}

template <class T, class Less>
ORBX_HD void insertion_sort(T* f, long first, long last, Less lt) {
  if (first == last) return;
  for (long i = first + 1; i != last; ++i) {
    if (lt(f[i], f[first])) {
      T v = f[i];
      for (long j = i; j > first; --j) f[j] = f[j - 1];
      f[first] = v;
    } else {
      unguarded_linear_insert(f, i, lt);
    }
  }
}

}  // namespace ss

// Sorts f[0..n) exactly as std::sort(f, f + n, lt) does.
template <class T, class Less>
ORBX_HD void stdsort_emul_t(T* f, long n, Less lt) {
  using namespace ss;
  if (n <= 0) return;
  // __introsort_loop with an explicit stack of (first, last, depth) for the right-hand recursions
  long stF[64], stL[64];
  int stD[64];
  int sp = 0;
  int lg = 0;
  for (long t = n; t > 1; t >>= 1) ++lg;
  long first = 0, last = n;
  int depth = 2 * lg;
  for (;;) {
    while (last - first > 16) {
      if (depth == 0) {
        heap_sort(f + first, last - first, lt);
        break;
      }
      --depth;
      // __move_median_to_first(first, first+1, mid, last-1)
      const long a = first + 1, b = first + (last - first) / 2, c = last - 1;
      if (lt(f[a], f[b])) {
        if (lt(f[b], f[c])) swp(f[first], f[b]);
        else if (lt(f[a], f[c])) swp(f[first], f[c]);
        else swp(f[first], f[a]);
      } else if (lt(f[a], f[c])) swp(f[first], f[a]);
      else if (lt(f[b], f[c])) swp(f[first], f[c]);
      else swp(f[first], f[b]);
      // __unguarded_partition(first+1, last, first)
      long lo = first + 1, hi = last;
      for (;;) {
        while (lt(f[lo], f[first])) ++lo;
        --hi;
        while (lt(f[first], f[hi])) --hi;
        if (!(lo < hi)) break;
        swp(f[lo], f[hi]);
        ++lo;
      }
      // recurse on [lo, last) first (depth-first, exactly as the recursive call would), then
      // continue with [first, lo): the right part must be fully processed before the left.
      // The recursion is tail-structured: right call runs to completion, then the loop
      // continues on the left.  Emulate by pushing the LEFT continuation and looping on the right.
      stF[sp] = first; stL[sp] = lo; stD[sp] = depth; ++sp;
      first = lo;
    }
    if (sp == 0) break;
    --sp;
    first = stF[sp]; last = stL[sp]; depth = stD[sp];
  }
  // __final_insertion_sort
  if (n > 16) {
    insertion_sort(f, 0, 16, lt);
    for (long i = 16; i < n; ++i) unguarded_linear_insert(f, i, lt);
  } else {
    insertion_sort(f, 0, n, lt);
  }
}

// Reorders f[0..n) exactly as std::nth_element(f, f + nth, f + n, lt) does (GCC 13 bits/stl_algo.h __introselect:
// median-of-3 partitioning around the nth position until at most 3 elements remain, heap-select after 2*floor(log2 n)
// rounds, insertion sort of the last range).  cv::KeyPointsFilter::retainBest, which ComputeKeyPointsOld calls per cell
// and per level (Features/ORBextractor.cpp:1452,1482), keeps the FIRST n elements of that arrangement, so which of the
// keypoints with tied responses survive, and in which order, is decided here.
template <class T, class Less>
ORBX_HD void stdnth_emul_t(T* f, long n, long nth, Less lt) {
  using namespace ss;
  if (n <= 0 || nth >= n) return;
  int lg = 0;
  for (long t = n; t > 1; t >>= 1) ++lg;
  long first = 0, last = n;
  int depth = 2 * lg;
  while (last - first > 3) {
    if (depth == 0) {
      // __heap_select(first, nth + 1, last), then iter_swap(first, nth)
      T* b = f + first;
      const long mid = nth + 1 - first, len = last - first;
      if (mid >= 2) {
        long parent = (mid - 2) / 2;
        for (;;) {
          T v = b[parent];
          adjust_heap(b, parent, mid, v, lt);
          if (parent == 0) break;
          parent--;
        }
      }
      for (long i = mid; i < len; ++i)
        if (lt(b[i], b[0])) {  // __pop_heap(first, middle, i)
          T v = b[i];
          b[i] = b[0];
          adjust_heap(b, 0, mid, v, lt);
        }
      swp(f[first], f[nth]);
      return;
    }
    --depth;
    const long a = first + 1, b = first + (last - first) / 2, c = last - 1;  // __move_median_to_first(first, first+1, mid, last-1)
    if (lt(f[a], f[b])) {
      if (lt(f[b], f[c])) swp(f[first], f[b]);
      else if (lt(f[a], f[c])) swp(f[first], f[c]);
      else swp(f[first], f[a]);
    } else if (lt(f[a], f[c])) swp(f[first], f[a]);
    else if (lt(f[b], f[c])) swp(f[first], f[c]);
    else swp(f[first], f[b]);
    long lo = first + 1, hi = last;  // __unguarded_partition(first+1, last, first)
    for (;;) {
      while (lt(f[lo], f[first])) ++lo;
      --hi;
      while (lt(f[first], f[hi])) --hi;
      if (!(lo < hi)) break;
      swp(f[lo], f[hi]);
      ++lo;
    }
    if (lo <= nth) first = lo; else last = lo;
  }
  insertion_sort(f, first, last, lt);
}

// Sorts f[0..n) exactly as std::sort(f, f + n, [](a, b){ return a.key < b.key; }) does.
ORBX_HD void stdsort_emul(SortElem* f, long n) { stdsort_emul_t(f, n, ss::KeyLess()); }


}  // namespace orbx
