// Host-side helpers of the engine library (no device work).
#include "../../include/bramble_b200.h"
#include <algorithm>
#include <cstdint>
#include <vector>

extern "C" {

// Order in which cgranges' IITree::index() leaves the intervals of one tree: libstdc++ std::sort
// (introsort, unstable) with a comparator on `st` only (reference: cgranges cpp/IITree.h index(),
// called from src/g2t.cpp:61).  Tie order among equal starts is whatever introsort yields for the
// insertion sequence, so the same algorithm from the same libstdc++ is run on (st, original index).
int brb_sort_by_start(uint64_t n, const int32_t* st, uint32_t* perm) {
  if (n && (!st || !perm)) return BRB_E_ARG;
  struct Key { int32_t st; uint32_t idx; };
  std::vector<Key> a(n);
  for (uint64_t i = 0; i < n; i++) { a[i].st = st[i]; a[i].idx = (uint32_t)i; }
  std::sort(a.begin(), a.end(), [](const Key& x, const Key& y) { return x.st < y.st; });
  for (uint64_t i = 0; i < n; i++) perm[i] = a[i].idx;
  return BRB_OK;
}

}  // extern "C"
