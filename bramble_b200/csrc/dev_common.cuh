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
constexpr uint32_t NONE = 0xffffffffu;

// One guide exon = one IITree node (include/g2t.h:12-43) = one 32-byte, 32-byte-aligned record: exactly one
// L2/HBM sector.  First half = search/filter key (the only part the overlap scan reads), second half = the
// links of the exon chain in GENOMIC order.  gclib merges overlapping and adjacent exons of a transcript
// (gclib/gff.cpp:988-1075), so the exons of one tid inside a table are disjoint and ascending; that is
// validated at brb_index_create and lets a candidate follow its transcript with one gather per read exon
// instead of re-scanning the hit range of every exon.
struct __align__(32) Rec {
  int32_t st, en; uint32_t tid, pos_start;
  int32_t gprev_en;        // end of the genomically previous exon of this tid (0 if none)
  int32_t gnext_st;        // start of the genomically next exon of this tid (INT32_MAX if none)
  uint32_t gnext;          // its record index (NONE if none)
  uint32_t meta;           // exon_id | has_prev<<8 | has_next<<9  (prev/next in TRANSCRIPT order, include/g2t.h)
};

struct DevIndex {
  uint32_t n_tables; uint32_t n_chrom; uint64_t n_rec; uint32_t bin_shift;
  uint32_t max_chain;            // most exons of one transcript (exon_id is a uint8_t: differences wrap beyond 256)
  const uint32_t* table_off;     // n_tables+1
  const int32_t* st;             // sorted starts (search key, SoA)
  const int32_t* pmax;           // running max of `en` inside each table (monotone -> binary search)
  const uint4* rkey; const uint4* rlink;   // the records' two halves (see rec_key / rec_link)
  const uint32_t* gprev;         // record index of the genomically previous exon of the same tid (NONE if none)
  const uint32_t* bin_off;       // n_tables+1 offsets into bins
  const uint2* bins;             // per table, per bin b: .x = first i with pmax[i] > b<<shift, .y = first i with st[i] >= b<<shift
  const uint64_t* chrom_off; const uint8_t* genome;   // base codes 0..4, TWO per byte (low nibble = even offset), or null
};
// base code at offset o (in bases from the start of the first chromosome)
__device__ __forceinline__ uint32_t genome_code(const DevIndex& ix, uint64_t o) { return (uint32_t)(__ldg(&ix.genome[o >> 1]) >> ((o & 1) << 2)) & 15u; }

// Batches of 32 alignments a persistent warp takes per ticket (one same-address atomic each).  Measured on the C2
// workload: two per ticket take 5 us off k_front_fast and add 5 us to k_back (uneven batches), four cost both.
#ifndef VAR_TICKET_FRONT
#define VAR_TICKET_FRONT 2
#endif
#ifndef VAR_TICKET_BACK
#define VAR_TICKET_BACK 1
#endif
constexpr uint32_t kTicketFront = VAR_TICKET_FRONT, kTicketBack = VAR_TICKET_BACK;

// Stores of data that streams through once per bundle (match records, rows): VAR_STREAM_HINTS=1 marks them
// evict-first (st.global.cs) so that they displace less of the index in L2 (measured: 5 us off k_front_fast).
// An L2 persisting window on the record array (cudaStreamAttributeAccessPolicyWindow) was measured too and made
// the front kernel slower (0.58 -> 0.61 ms with half of the set-aside limit, 1.02 ms with all of it).
#ifndef VAR_STREAM_HINTS
#define VAR_STREAM_HINTS 1
#endif
template <typename T> __device__ __forceinline__ void st_stream(T* p, T v) {
#if VAR_STREAM_HINTS
  __stcs(p, v);
#else
  *p = v;
#endif
}

// Loads of batch columns and match records that are read once per bundle: VAR_STREAM_LOADS >= 1 marks the front kernel's,
// >= 2 the back kernel's too, evict-first (ld.global.cs), so that the index (plain loads) outlives them in L2.
#ifndef VAR_STREAM_LOADS
#define VAR_STREAM_LOADS 0
#endif
template <int LEVEL, typename T> __device__ __forceinline__ T ld_stream(const T* p) {
#if VAR_STREAM_LOADS
  if (VAR_STREAM_LOADS >= LEVEL) return __ldcs(p);
#endif
  return __ldg(p);
}

struct DevBatch {
  uint32_t n_aln, n_groups, n_exon, n_cigar;
  const uint32_t* group_off; const int32_t* refid; const int8_t* strand;
  const uint32_t* exon_off; const uint32_t* exon_st; const uint32_t* exon_en;
  const uint32_t* cigar_off; const uint32_t* cigar; const int32_t* mate_idx;
  const uint64_t* group_seq_off; const uint8_t* seq; const uint32_t* aln_group;
  uint32_t back_shift;           // k_back / k_place: log2 of the alignments per batch (5 for short reads, 3 for long reads)
  uint32_t front_shift;          // k_front: the same for the front kernel (long reads: 3, or 0 when the engine has measured that to be faster on this data)
  uint32_t compact_rows;         // k_back<.., PACKED, ..>: 1 = 12-byte rows without the source alignment, 16-bit unit sizes and one row offset per batch (BRB_OPT_PACKED_ROWS = 2); 2 = the same with 8-byte rows (= 3)
  uint32_t count_hits;           // k_front: count EVERY index hit of every read exon (the reference-defined figure); else later exons stop at the first survivor
};

// Device-side control block of one brb_project_resident call: capacities of the pre-allocated pools and the
// running totals.  No kernel of the sequence needs a host round trip; a pool that turns out too small makes
// the producers stop writing (they keep counting), every later kernel returns at once, and the host grows
// the pool and replays the sequence.
struct Ctl {
  unsigned long long n_hits, out_words, rescue_bytes, n_matches, n_rescue, ksw_max_mat, ksw_max_tlen, n_mates;   // work counters
  unsigned long long claimed;         // k_back: rows | CIGAR words << 32 handed out so far
  unsigned long long ksw_max_qlen, ksw_cells;   // clip rescue: longest query of a big problem; DP cells (query x target) of all problems
  unsigned long long ksw_cells_run;             // ... of the problems that ran (identical problems of a read share one DP)
  uint32_t n_cand, cap_cand;          // candidate slots allocated / capacity
  uint32_t n_ideal, cap_ideal;        // ideal-CIGAR pool words
  uint32_t n_rows, cap_rows;          // output rows
  uint32_t n_tmpw, cap_tmpw;          // merge scratch pool words
  uint32_t n_outw, cap_outw;          // dense output CIGAR words
  uint32_t n_tb, cap_tb;              // clip rescue: traceback ops / override-coded clip CIGAR words
  uint32_t bad_input;                 // 1: mate link not symmetric, 2: read exons not ascending/disjoint, 3: >= 2^28 CIGAR words in one batch of 32 alignments, 4: packed rows, 5: CSR offsets, 6: compact rows (a unit of 65536 rows or more), 7: 8-byte rows (tid >= 2^22, pos >= 2^24 or a CIGAR field >= 2^13)
  uint32_t ticket;                    // next batch of 32 alignments for k_front (dynamic scheduling)
  uint32_t ticket_a, ticket_b;        // k_back: next batch; CTAs that have left (the last one publishes the totals)
  uint32_t kq_n[6], kq_t[6];          // clip rescue: problems per size class, next problem of each class
  uint32_t ksw_big_skipped;           // big problems that did not fit the per-warp scratch (the host grows it and replays)
  uint32_t ticket_m;                  // k_merge_long: next 32 candidate slots
};
constexpr int N_CAPS = 6;
// End of a kernel sequence, ONE thread, every other writer of `ctl` done and fenced: hand the counters to the host
// through mapped pinned memory (no separate copy operation on the stream) and leave `ctl` ready for the next
// sequence of the same pool capacities (counters and tickets zero, caps kept), so that a steady-state step needs no
// host->device initialisation either.
__device__ __forceinline__ void ctl_publish_and_reset(Ctl* ctl, Ctl* host_result) {
  static_assert(sizeof(Ctl) % 8 == 0, "Ctl is copied in 8-byte words");
  volatile unsigned long long* s = reinterpret_cast<volatile unsigned long long*>(ctl);
  volatile unsigned long long* d = reinterpret_cast<volatile unsigned long long*>(host_result);
  const uint32_t cc = ctl->cap_cand, ci = ctl->cap_ideal, cr = ctl->cap_rows, ct = ctl->cap_tmpw, co = ctl->cap_outw, cb = ctl->cap_tb;
  for (uint32_t i = 0; i < sizeof(Ctl) / 8; i++) { d[i] = s[i]; s[i] = 0ull; }
  ctl->cap_cand = cc; ctl->cap_ideal = ci; ctl->cap_rows = cr; ctl->cap_tmpw = ct; ctl->cap_outw = co; ctl->cap_tb = cb;
  __threadfence_system();
}
__device__ __forceinline__ bool ctl_overflow(const Ctl* c) {
  return c->n_cand > c->cap_cand || c->n_ideal > c->cap_ideal || c->n_rows > c->cap_rows || c->n_tmpw > c->cap_tmpw ||
         c->n_outw > c->cap_outw || c->n_tb > c->cap_tb || c->bad_input != 0;
}

__device__ __forceinline__ uint4 ldg4(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// Record halves.  VAR_SPLIT_REC=0 (default): the 32-byte records as one interleaved array (key, link, key, link ...) --
// a candidate that passes the filter finds its link half in the sector its key came in.  VAR_SPLIT_REC=1: two dense
// arrays (a warp walking consecutive keys touches half the lines); measured slower, 0.627 vs 0.590 ms for k_front_fast
// on the C2 workload, because 56 % of the candidates go on to read the link half from a second line.
#ifndef VAR_SPLIT_REC
#define VAR_SPLIT_REC 0
#endif
struct RecKey { int32_t st, en; uint32_t tid, pos_start; };
__device__ __forceinline__ const uint4* rec_key_ptr(const DevIndex& ix, uint32_t i) { return VAR_SPLIT_REC ? ix.rkey + i : ix.rkey + 2 * (size_t)i; }
__device__ __forceinline__ RecKey rec_key(const DevIndex& ix, uint32_t i) {
  uint4 v = ldg4(rec_key_ptr(ix, i)); RecKey k; k.st = (int32_t)v.x; k.en = (int32_t)v.y; k.tid = v.z; k.pos_start = v.w; return k;
}
struct RecLink { int32_t gprev_en, gnext_st; uint32_t gnext, meta; };
__device__ __forceinline__ RecLink rec_link(const DevIndex& ix, uint32_t i) {
  uint4 v = ldg4(VAR_SPLIT_REC ? ix.rlink + i : ix.rkey + 2 * (size_t)i + 1);
  RecLink k; k.gprev_en = (int32_t)v.x; k.gnext_st = (int32_t)v.y; k.gnext = v.z; k.meta = v.w; return k;
}
// the second half of the link half only: .x = gnext, .y = meta
__device__ __forceinline__ uint2 rec_link_tail(const DevIndex& ix, uint32_t i) {
  const uint4* q = VAR_SPLIT_REC ? ix.rlink + i : ix.rkey + 2 * (size_t)i + 1;
  return __ldg(reinterpret_cast<const uint2*>(q) + 1);
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
// Same filters, result only (the overlap scan does not need the Hit fields).
__device__ __forceinline__ bool passes_filter(const brb_params& c, bool minus, int status, uint32_t qs, uint32_t qe, uint32_t s, uint32_t e) {
  const bool ml = (status == MIDDLE_EXON || status == LAST_EXON), fm = (status == FIRST_EXON || status == MIDDLE_EXON);
  const uint32_t lim_l = ml ? c.max_junc_ins : c.max_clip;
  const uint32_t lim_r = minus ? c.max_junc_ins : (fm ? c.max_junc_ins : c.max_clip);
  bool ok = true;
  if (s <= qs) { if (ml && qs - s > c.max_junc_gap) ok = false; } else { if (s - qs > lim_l) ok = false; }
  if (e < qe) { if (qe - e > lim_r) ok = false; } else { if (fm && e - qe > c.max_junc_gap) ok = false; }
  return ok;
}

// cgranges IITree::overlap(qs,qe) == { i : st[i] < qe && qs < en[i] } in ascending i.
// On the flattened table: hi = first i with st[i] >= qe; lo = first i with pmax[i] > qs; members of
// [lo,hi) still need the `qs < en[i]` test.  A coarse bin table (one uint2 per 2^shift bases) brackets both
// bounds, then a branch-free binary search over the few records of the bracket finishes them exactly.
__device__ __forceinline__ void overlap_range(const DevIndex& ix, int32_t refid, bool minus, uint32_t qs, uint32_t qe,
                                              uint32_t& lo, uint32_t& hi) {
  lo = hi = 0;
  if (refid < 0 || (uint32_t)refid * 2u + 1u >= ix.n_tables) return;            // src/g2t.cpp:275-276
  const uint32_t t = (uint32_t)refid * 2u + (minus ? 1u : 0u);
  const uint32_t bo = __ldg(&ix.bin_off[t]), nb = __ldg(&ix.bin_off[t + 1]) - bo;
  if (nb < 2) return;                                                           // empty table
  uint32_t b0 = qs >> ix.bin_shift, b1 = qe >> ix.bin_shift;
  if (b0 > nb - 2) b0 = nb - 2;
  if (b1 > nb - 2) b1 = nb - 2;
  const uint2 A0 = __ldg(&ix.bins[bo + b0]), A1 = __ldg(&ix.bins[bo + b0 + 1]);
  const uint2 B0 = __ldg(&ix.bins[bo + b1]), B1 = __ldg(&ix.bins[bo + b1 + 1]);
  // both searches step together: two independent loads in flight per round instead of one
  uint32_t l = B0.y, n = B1.y - B0.y, l2 = A0.x, n2 = A1.x - A0.x;
  while (n > 0 || n2 > 0) {
    const uint32_t h = n >> 1, h2 = n2 >> 1;
    const int32_t v = n > 0 ? __ldg(&ix.st[l + h]) : 0, v2 = n2 > 0 ? __ldg(&ix.pmax[l2 + h2]) : 0;
    if (n > 0) { if (v < (int32_t)qe) { l += h + 1; n -= h + 1; } else n = h; }
    if (n2 > 0) { if (v2 <= (int32_t)qs) { l2 += h2 + 1; n2 -= h2 + 1; } else n2 = h2; }
  }
  hi = l; lo = l2 < hi ? l2 : hi;
}

// First record of table t (refid, strand) whose start is >= x, through the bin directory (the `hi` half of overlap_range).
__device__ __forceinline__ uint32_t st_lower_bound(const DevIndex& ix, int32_t refid, bool minus, uint32_t x) {
  if (refid < 0 || (uint32_t)refid * 2u + 1u >= ix.n_tables) return 0;
  const uint32_t t = (uint32_t)refid * 2u + (minus ? 1u : 0u);
  const uint32_t bo = __ldg(&ix.bin_off[t]), nb = __ldg(&ix.bin_off[t + 1]) - bo;
  if (nb < 2) return 0;
  uint32_t b1 = x >> ix.bin_shift;
  if (b1 > nb - 2) b1 = nb - 2;
  const uint2 B0 = __ldg(&ix.bins[bo + b1]), B1 = __ldg(&ix.bins[bo + b1 + 1]);
  uint32_t l = B0.y, n = B1.y - B0.y;
  while (n > 0) { const uint32_t h = n >> 1; if (__ldg(&ix.st[l + h]) < (int32_t)x) { l += h + 1; n -= h + 1; } else n = h; }
  return l;
}

// ---- warp helpers
__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, unsigned lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, v, d); if (lane >= (unsigned)d) v += y; }
  return v;
}
// Owner lane of flat item `item` given the warp's inclusive prefix sums (item < total): the first lane L with incl[L] > item.
__device__ __forceinline__ unsigned warp_owner(uint32_t incl, uint32_t item) {
  unsigned pos = 0;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) { const uint32_t v = __shfl_sync(0xffffffffu, incl, pos + d - 1); if (v <= item) pos += d; }
  return pos;
}
// The same search with the prefix sums in shared memory (any lane may ask for any item, no convergence needed).
__device__ __forceinline__ unsigned warp_owner_smem(const uint32_t* incl, uint32_t item) {
  unsigned pos = 0;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) if (incl[pos + d - 1] <= item) pos += d;
  return pos;
}
// bits [a,b) of a 32-bit word, 0 <= a <= b <= 32
__device__ __forceinline__ uint32_t bit_range(uint32_t a, uint32_t b) {
  const uint32_t hi = b >= 32 ? 0xffffffffu : ((1u << b) - 1u);
  const uint32_t lo = a >= 32 ? 0xffffffffu : ((1u << a) - 1u);
  return hi & ~lo;
}

// Warp-aggregated allocation of `want` words per lane from a bump pool; returns the lane's offset.
// `ok` is false for every lane of the warp when the pool is exhausted (the counter still advances).
__device__ __forceinline__ uint32_t warp_alloc(uint32_t* __restrict__ counter, uint32_t cap, uint32_t want, unsigned lane, bool& ok) {
  const uint32_t incl = warp_incl_scan(want, lane), tot = __shfl_sync(0xffffffffu, incl, 31);
  uint32_t base = 0;
  if (lane == 31 && tot) base = atomicAdd(counter, tot);
  base = __shfl_sync(0xffffffffu, base, 31);
  ok = (unsigned long long)base + tot <= cap;
  return base + incl - want;
}


// Bump allocation for PERSISTENT warps: the warp owns a private chunk of the pool and serves its lanes from
// it; the global counter is touched once per chunk (a same-address atomic per 32 items was the top stall of
// the first fused kernels).  Chunk tails left at the end are simply unused pool words.
struct WarpPool { uint32_t cur, end; };
constexpr uint32_t POOL_CHUNK = 2048;
__device__ __forceinline__ uint32_t warp_pool_alloc(WarpPool& wp, uint32_t* __restrict__ counter, uint32_t cap, uint32_t want,
                                                    unsigned lane, bool& ok) {
  const uint32_t incl = warp_incl_scan(want, lane), tot = __shfl_sync(0xffffffffu, incl, 31);
  if (tot > wp.end - wp.cur) {
    const uint32_t grab = tot > POOL_CHUNK ? tot : POOL_CHUNK;
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(counter, grab);
    base = __shfl_sync(0xffffffffu, base, 0);
    wp.cur = base; wp.end = base + grab;
  }
  ok = wp.end <= cap && wp.end >= wp.cur;          // the whole chunk lies inside the pool (and did not wrap)
  const uint32_t off = wp.cur + incl - want;
  wp.cur += tot;
  return off;
}

}  // namespace brb
