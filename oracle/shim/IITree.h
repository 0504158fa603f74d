// TEST INFRASTRUCTURE (oracle shim) -- not product code.
// Restatement of cgranges cpp/IITree.h (lh3/cgranges @b3d5e2c; subprojects/cgranges.wrap in
// the reference; un-vendored, no network).  Implicit augmented interval tree: index() =
// std::sort by start only + bottom-up max; overlap() returns every i with
// a[i].st < en && st < a[i].en in ascending i.  Call sites: reference src/g2t.cpp:57,61,76,110.
#pragma once
#include <vector>
#include <algorithm>
#include <cstddef>

template<typename S, typename T>
class IITree {
  struct StackCell {
    size_t x; int k, w;
    StackCell() {}
    StackCell(int k_, size_t x_, int w_) : x(x_), k(k_), w(w_) {}
  };
  struct Interval {
    S st, en, max; T data;
    Interval(const S &s, const S &e, const T &d) : st(s), en(e), max(e), data(d) {}
  };
  struct IntervalLess {
    bool operator()(const Interval &a, const Interval &b) const { return a.st < b.st; }
  };
  std::vector<Interval> a;
  int max_level;
  int index_core(std::vector<Interval> &a) {
    size_t i, last_i = 0; S last = S(); int k;
    if (a.size() == 0) return -1;
    for (i = 0; i < a.size(); i += 2) last_i = i, last = a[i].max = a[i].en;
    for (k = 1; (size_t)(1LL << k) <= a.size(); ++k) {
      size_t x = 1LL << (k - 1), i0 = (x << 1) - 1, step = x << 2;
      for (i = i0; i < a.size(); i += step) {
        S el = a[i - x].max;
        S er = i + x < a.size() ? a[i + x].max : last;
        S e = a[i].en;
        e = e > el ? e : el;
        e = e > er ? e : er;
        a[i].max = e;
      }
      last_i = (last_i >> k & 1) ? last_i - x : last_i + x;
      if (last_i < a.size() && a[last_i].max > last) last = a[last_i].max;
    }
    return k - 1;
  }
public:
  IITree() : max_level(-1) {}
  void add(const S &s, const S &e, const T &d) { a.push_back(Interval(s, e, d)); }
  void index(void) {
    std::sort(a.begin(), a.end(), IntervalLess());
    max_level = index_core(a);
  }
  bool overlap(const S &st, const S &en, std::vector<size_t> &out) const {
    int t = 0; StackCell stack[64];
    out.clear();
    if (max_level < 0) return false;
    stack[t++] = StackCell(max_level, (1LL << max_level) - 1, 0);
    while (t) {
      StackCell z = stack[--t];
      if (z.k <= 3) {
        size_t i, i0 = z.x >> z.k << z.k, i1 = i0 + (1LL << (z.k + 1)) - 1;
        if (i1 >= a.size()) i1 = a.size();
        for (i = i0; i < i1 && a[i].st < en; ++i)
          if (st < a[i].en) out.push_back(i);
      } else if (z.w == 0) {
        size_t y = z.x - (1LL << (z.k - 1));
        stack[t++] = StackCell(z.k, z.x, 1);
        if (y >= a.size() || a[y].max > st) stack[t++] = StackCell(z.k - 1, y, 0);
      } else if (z.x < a.size() && a[z.x].st < en) {
        if (st < a[z.x].en) out.push_back(z.x);
        stack[t++] = StackCell(z.k - 1, z.x + (1LL << (z.k - 1)), 0);
      }
    }
    return out.size() > 0;
  }
  size_t size(void) const { return a.size(); }
  const S &start(size_t i) const { return a[i].st; }
  const S &end(size_t i) const { return a[i].en; }
  const T &data(size_t i) const { return a[i].data; }
};
