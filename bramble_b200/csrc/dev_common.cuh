// Device-side data layout and the small arithmetic helpers shared by all kernels.
// Everything here is integer work; citations are to the reference (zrudnick/bramble, C++).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/bramble_b200.h"

namespace brb {

enum { FIRST_EXON = 0, MIDDLE_EXON = 1, LAST_EXON = 2, ONLY_EXON = 3 };
enum { CM = 0, CI = 1, CD = 2, CN = 3, CS = 4, CH = 5, CP = 6, CEQ = 7, CX = 8,
       OV_M = 10, OV_D = 11, OV_I = 12, OV_S = 13 };   // override codes: include/evaluate.h:10-13
#define BRB_KSW_NEG_INF (-0x40000000)

// One guide exon = one IITree node (include/g2t.h:12-43), 48 B, three 16-byte sectors:
//   .k  search/filter key   .nb neighbours in transcript order   .x  transcript length + exon id
struct __align__(16) Rec {
  int32_t st, en; uint32_t tid, pos_start;
  uint32_t prev_st, prev_en, next_st, next_en;
  uint32_t tlen, meta /* exon_id | has_prev<<8 | has_next<<9 */, pad0, pad1;
};

struct DevIndex {
  uint32_t n_tables; uint32_t n_chrom; uint64_t n_rec;
  const uint32_t* table_off;     // n_tables+1
  const int32_t* st;             // sorted starts (search key, SoA)
  const int32_t* pmax;           // running max of `en` inside each table (monotone -> binary search)
  const Rec* rec;
  const uint64_t* chrom_off; const uint8_t* genome;   // codes 0..4, or null
};

struct DevBatch {
  uint32_t n_aln, n_groups, n_exon, n_cigar;
  const uint32_t* group_off; const int32_t* refid; const int8_t* strand;
  const uint32_t* exon_off; const uint32_t* exon_st; const uint32_t* exon_en;
  const uint32_t* cigar_off; const uint32_t* cigar; const int32_t* mate_idx;
  const uint64_t* group_seq_off; const uint8_t* seq; const uint32_t* aln_group;
};

__device__ __forceinline__ uint4 ldg4(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

struct RecKey { int32_t st, en; uint32_t tid, pos_start; };
__device__ __forceinline__ RecKey rec_key(const Rec* r, uint32_t i) {
  uint4 v = ldg4(&r[i]); RecKey k; k.st = (int32_t)v.x; k.en = (int32_t)v.y; k.tid = v.z; k.pos_start = v.w; return k;
}

__device__ __forceinline__ int exon_status(uint32_t n, uint32_t j) {   // src/evaluate.cpp:44-56
  return n == 1 ? ONLY_EXON : (j == 0 ? FIRST_EXON : (j < n - 1 ? MIDDLE_EXON : LAST_EXON));
}

struct Hit { uint32_t pos, left_gap, left_ins, right_gap, right_ins; };

// IntervalTree::findOverlapping per-hit arithmetic and tolerance filters, src/g2t.cpp:156-227.
// Returns false when the hit is dropped.  minus == query strand '-'.
__device__ __forceinline__ bool filter_hit(const brb_params& c, bool minus, int status, uint32_t qs, uint32_t qe,
                                           uint32_t s, uint32_t e, uint32_t pos_start, Hit& h) {
  h.pos = h.left_gap = h.left_ins = h.right_gap = h.right_ins = 0;
  const bool ml = (status == MIDDLE_EXON || status == LAST_EXON), fm = (status == FIRST_EXON || status == MIDDLE_EXON);
  if (!minus) {
    if (s <= qs) { h.pos = (qs - s) + pos_start; h.left_gap = qs - s; if (ml && h.left_gap > c.max_junc_gap) return false; }
    else { h.pos = pos_start; h.left_ins = s - qs; if (h.left_ins > (ml ? c.max_junc_ins : c.max_clip)) return false; }
    if (e < qe) { h.right_ins = qe - e; if (h.right_ins > (fm ? c.max_junc_ins : c.max_clip)) return false; }
    else if (qe < e) { h.right_gap = e - qe; if (fm && h.right_gap > c.max_junc_gap) return false; }
  } else {
    if (qe <= e) { h.pos = (e - qe) + pos_start; h.right_gap = e - qe; if (fm && h.right_gap > c.max_junc_gap) return false; }
    else { h.pos = pos_start; h.right_ins = qe - e;
           if (h.right_ins > c.max_junc_ins) return false; }   // src/g2t.cpp:204: `status == FIRST_EXON || MIDDLE_EXON` is always true
    if (qs < s) { h.left_ins = s - qs; if (h.left_ins > (ml ? c.max_junc_ins : c.max_clip)) return false; }
    else if (s < qs) { h.left_gap = qs - s; if (ml && h.left_gap > c.max_junc_gap) return false; }
  }
  return true;
}

// cgranges IITree::overlap(qs,qe) == { i : st[i] < qe && qs < en[i] } in ascending i.
// On the flattened table: hi = first i with st[i] >= qe; lo = first i with pmax[i] > qs; members of
// [lo,hi) still need the `qs < en[i]` test.  Two branch-free binary searches.
__device__ __forceinline__ void overlap_range(const DevIndex& ix, int32_t refid, bool minus, uint32_t qs, uint32_t qe,
                                              uint32_t& lo, uint32_t& hi) {
  lo = hi = 0;
  if (refid < 0 || (uint32_t)refid * 2u + 1u >= ix.n_tables) return;            // src/g2t.cpp:275-276
  uint32_t t = (uint32_t)refid * 2u + (minus ? 1u : 0u);
  uint32_t b = __ldg(&ix.table_off[t]), e = __ldg(&ix.table_off[t + 1]);
  uint32_t l = b, n = e - b;
  while (n > 0) { uint32_t h = n >> 1; if (__ldg(&ix.st[l + h]) < (int32_t)qe) { l += h + 1; n -= h + 1; } else n = h; }
  hi = l;
  l = b; n = hi - b;
  while (n > 0) { uint32_t h = n >> 1; if (__ldg(&ix.pmax[l + h]) <= (int32_t)qs) { l += h + 1; n -= h + 1; } else n = h; }
  lo = l;
}

// IntervalTree::findOverlappingForTid, src/g2t.cpp:70-101: first overlapping record (ascending index) of `tid`.
__device__ __forceinline__ bool find_for_tid(const DevIndex& ix, int32_t refid, bool minus, uint32_t tid, uint32_t qs, uint32_t qe, uint32_t& out) {
  if (qs == 0 && qe == 0) return false;
  uint32_t lo, hi; overlap_range(ix, refid, minus, qs, qe, lo, hi);
  for (uint32_t i = lo; i < hi; i++) {
    RecKey k = rec_key(ix.rec, i);
    if ((int32_t)qs < k.en && k.tid == tid) { out = i; return true; }
  }
  return false;
}

}  // namespace brb
