// gosort.h -- sort.Sort of Go 1.21 (the reference's toolchain, go.mod:3) over an
// index interface, because makePodOrderdByAge (podmanager.go:326-347) sorts with a
// NON-strict Less (`<=`), so the order of pods with equal assume times is whatever
// this exact algorithm leaves behind: for n <= 12 an insertion sort (ties come out
// reversed), above that pattern-defeating quicksort. Restated from the published
// algorithm (sort/zsortinterface.go); tie order for n > 12 is unpinned.
#pragma once
#include <stdint.h>

namespace gs_gosort {

template <class Less, class Swap>
struct Sorter {
  Less less;
  Swap swap;

  void insertion(int a, int b) {
    for (int i = a + 1; i < b; ++i)
      for (int j = i; j > a && less(j, j - 1); --j) swap(j, j - 1);
  }

  void sift_down(int lo, int hi, int first) {
    int root = lo;
    for (;;) {
      int child = 2 * root + 1;
      if (child >= hi) return;
      if (child + 1 < hi && less(first + child, first + child + 1)) ++child;
      if (!less(first + root, first + child)) return;
      swap(first + root, first + child);
      root = child;
    }
  }

  void heap_sort(int a, int b) {
    const int first = a, lo = 0, hi = b - a;
    for (int i = (hi - 1) / 2; i >= 0; --i) sift_down(i, hi, first);
    for (int i = hi - 1; i >= 0; --i) {
      swap(first, first + i);
      sift_down(lo, i, first);
    }
  }

  // returns mid; *already = the range was already partitioned
  int partition(int a, int b, int pivot, bool *already) {
    swap(a, pivot);
    int i = a + 1, j = b - 1;
    while (i <= j && less(i, a)) ++i;
    while (i <= j && !less(j, a)) --j;
    if (i > j) {
      swap(j, a);
      *already = true;
      return j;
    }
    swap(i, j);
    ++i;
    --j;
    for (;;) {
      while (i <= j && less(i, a)) ++i;
      while (i <= j && !less(j, a)) --j;
      if (i > j) break;
      swap(i, j);
      ++i;
      --j;
    }
    swap(j, a);
    *already = false;
    return j;
  }

  int partition_equal(int a, int b, int pivot) {
    swap(a, pivot);
    int i = a + 1, j = b - 1;
    for (;;) {
      while (i <= j && !less(a, i)) ++i;
      while (i <= j && less(a, j)) --j;
      if (i > j) break;
      swap(i, j);
      ++i;
      --j;
    }
    return i;
  }

  bool partial_insertion(int a, int b) {
    const int max_steps = 5, shortest_shifting = 50;
    int i = a + 1;
    for (int step = 0; step < max_steps; ++step) {
      while (i < b && !less(i, i - 1)) ++i;
      if (i == b) return true;
      if (b - a < shortest_shifting) return false;
      swap(i, i - 1);
      if (i - a >= 2) {
        for (int j = i - 1; j >= 1; --j) {
          if (!less(j, j - 1)) break;
          swap(j, j - 1);
        }
      }
      if (b - i >= 2) {
        for (int j = i + 1; j < b; ++j) {
          if (!less(j, j - 1)) break;
          swap(j, j - 1);
        }
      }
    }
    return false;
  }

  static int bits_len(unsigned v) {
    int n = 0;
    while (v) {
      ++n;
      v >>= 1;
    }
    return n;
  }

  void break_patterns(int a, int b) {
    const int length = b - a;
    if (length >= 8) {
      uint64_t random = (uint64_t)length;
      const unsigned modulus = 1u << bits_len((unsigned)length);
      const int idx = a + (length / 4) * 2 - 1;
      for (int i = 0; i < 3; ++i) {
        random ^= random << 13;
        random ^= random >> 7;
        random ^= random << 17;
        int other = (int)((unsigned)random & (modulus - 1));
        if (other >= length) other -= length;
        swap(idx - 1 + i, a + other);
      }
    }
  }

  void order2(int &a, int &b, int *swaps) {
    if (less(b, a)) {
      ++*swaps;
      const int t = a;
      a = b;
      b = t;
    }
  }
  int median(int a, int b, int c, int *swaps) {
    order2(a, b, swaps);
    order2(b, c, swaps);
    order2(a, b, swaps);
    return b;
  }
  int median_adjacent(int a, int *swaps) { return median(a - 1, a, a + 1, swaps); }

  enum Hint { kUnknown, kIncreasing, kDecreasing };

  int choose_pivot(int a, int b, Hint *hint) {
    const int shortest_ninther = 50, max_swaps = 4 * 3;
    const int l = b - a;
    int swaps = 0;
    int i = a + l / 4 * 1, j = a + l / 4 * 2, k = a + l / 4 * 3;
    if (l >= 8) {
      if (l >= shortest_ninther) {
        i = median_adjacent(i, &swaps);
        j = median_adjacent(j, &swaps);
        k = median_adjacent(k, &swaps);
      }
      j = median(i, j, k, &swaps);
    }
    *hint = swaps == 0 ? kIncreasing : (swaps == max_swaps ? kDecreasing : kUnknown);
    return j;
  }

  void reverse_range(int a, int b) {
    for (int i = a, j = b - 1; i < j; ++i, --j) swap(i, j);
  }

  void pdqsort(int a, int b, int limit) {
    const int max_insertion = 12;
    bool was_balanced = true, was_partitioned = true;
    for (;;) {
      const int length = b - a;
      if (length <= max_insertion) {
        insertion(a, b);
        return;
      }
      if (limit == 0) {
        heap_sort(a, b);
        return;
      }
      if (!was_balanced) {
        break_patterns(a, b);
        --limit;
      }
      Hint hint;
      int pivot = choose_pivot(a, b, &hint);
      if (hint == kDecreasing) {
        reverse_range(a, b);
        pivot = (b - 1) - (pivot - a);
        hint = kIncreasing;
      }
      if (was_balanced && was_partitioned && hint == kIncreasing) {
        if (partial_insertion(a, b)) return;
      }
      if (a > 0 && !less(a - 1, pivot)) {
        a = partition_equal(a, b, pivot);
        continue;
      }
      bool already = false;
      const int mid = partition(a, b, pivot, &already);
      was_partitioned = already;
      const int left_len = mid - a, right_len = b - mid;
      const int balance_threshold = length / 8;
      if (left_len < right_len) {
        was_balanced = left_len >= balance_threshold;
        pdqsort(a, mid, limit);
        a = mid + 1;
      } else {
        was_balanced = right_len >= balance_threshold;
        pdqsort(mid + 1, b, limit);
        b = mid;
      }
    }
  }
};

// sort.Sort(data): n = data.Len(), limit = bits.Len(uint(n))
template <class Less, class Swap>
inline void sort(int n, Less less, Swap swap) {
  if (n <= 1) return;
  Sorter<Less, Swap> s{less, swap};
  s.pdqsort(0, n, Sorter<Less, Swap>::bits_len((unsigned)n));
}

}  // namespace gs_gosort
