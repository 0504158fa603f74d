// Kernel group 1: exon-overlap lookup on the flattened sorted-interval table.
// Replaces g2tTree::getGuideExons / IntervalTree::findOverlapping (reference src/g2t.cpp:103-257,334-344)
// and the per-exon bookkeeping of get_intervals that does not depend on a candidate
// (src/evaluate.cpp:184-282: "no hit" -> small-middle-exon INS path or failure of the strand).
//
// Work mapping: a warp owns 32 consecutive alignments (lane = alignment).  For every (strand, exon index)
// round each lane brackets its query's index range [lo,hi) through the bin table, then the warp walks the
// CONCATENATION of the 32 ranges 32 records at a time (lane = record), so a read that hits 200 isoforms of a
// hot gene costs 7 fully occupied rounds instead of 200 single-lane ones.  Results come back to the owning
// lane through ballots: survivors per query, first/last surviving record.
#pragma once
#include "dev_common.cuh"

namespace brb {

// A task is (alignment a, strand s): task id = 2a+s.  Strands to check: src/evaluate.cpp:58-67.
__device__ __forceinline__ bool task_active(const brb_params& p, int8_t read_strand, int s) {
  if (p.long_reads) return true;
  if (read_strand == '+') return s == 0;
  if (read_strand == '-') return s == 1;
  return true;
}

// The three CSR offset arrays of a bundle (brb_batch_in): non-decreasing and inside their columns.  Runs once per upload,
// before any kernel follows an offset; a violation sets bad_input = 5 and every kernel of the sequence returns at once.
__global__ void __launch_bounds__(256) k_validate_batch(DevBatch b, Ctl* __restrict__ ctl) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  bool bad = false;
  if (i < b.n_aln) {
    const uint32_t e0 = b.exon_off[i], e1 = b.exon_off[i + 1], c0 = b.cigar_off[i], c1 = b.cigar_off[i + 1];
    bad = e1 < e0 || e1 > b.n_exon || c1 < c0 || c1 > b.n_cigar;
  }
  if (i < b.n_groups) { const uint32_t g0 = b.group_off[i], g1 = b.group_off[i + 1]; bad = bad || g1 < g0 || g1 > b.n_aln; }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicMax(&ctl->bad_input, 5u);
}

struct FlatOut { uint32_t nf, first, last, nhit; };   // survivors, first survivor, last survivor + 1, overlapping records

// Warp-cooperative filtered scan of 32 index ranges, counting pass.  Lane inputs: lo, n (0 = nothing), qs, qe,
// status; lane outputs: FlatOut of its own range.
// `minus` is per lane (the strand of the lane's query).
__device__ __forceinline__ void warp_flat_count(const DevIndex& ix, const brb_params& p, bool minus, unsigned lane,
                                                uint32_t lo, uint32_t n, uint32_t qs, uint32_t qe, int status, FlatOut& o) {
  status |= minus ? 4 : 0;
  const uint32_t incl = warp_incl_scan(n, lane), excl = incl - n;
  const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
  o.nf = 0; o.first = 0; o.last = 0; o.nhit = 0;
  for (uint32_t base = 0; base < total; base += 32) {
    const uint32_t item = base + lane;
    const bool act = item < total;
    const unsigned L = warp_owner(incl, act ? item : total - 1);
    const uint32_t o_lo = __shfl_sync(0xffffffffu, lo, L), o_ex = __shfl_sync(0xffffffffu, excl, L);
    const uint32_t o_qs = __shfl_sync(0xffffffffu, qs, L), o_qe = __shfl_sync(0xffffffffu, qe, L);
    const int o_sm = __shfl_sync(0xffffffffu, status, L), o_st = o_sm & 3;
    bool ov = false, sv = false;
    if (act) {
      const RecKey k = rec_key(ix, o_lo + (item - o_ex));
      ov = (int32_t)o_qs < k.en;                           // st < qe holds for every record below hi
      sv = ov && passes_filter(p, (o_sm & 4) != 0, o_st, o_qs, o_qe, (uint32_t)k.st, (uint32_t)k.en);
    }
    const uint32_t B = __ballot_sync(0xffffffffu, sv), H = __ballot_sync(0xffffffffu, ov);
    if (n && incl > base && excl < base + 32) {            // bits of this chunk that belong to MY range
      const uint32_t mine = bit_range(excl > base ? excl - base : 0, incl - base < 32 ? incl - base : 32);
      const uint32_t m = B & mine;
      if (m) {
        if (o.nf == 0) o.first = lo + (base + (uint32_t)__ffs(m) - 1 - excl);
        o.last = lo + (base + 31u - (uint32_t)__clz(m) - excl) + 1;
        o.nf += __popc(m);
      }
      o.nhit += __popc(H & mine);
    }
  }
}

// Standalone lookup for the parity tests (brb_lookup): one thread per query.
__global__ void k_lookup_queries(DevIndex ix, brb_params p, uint32_t nq, const int32_t* refid, const int8_t* strand,
                                 const uint32_t* qs_, const uint32_t* qe_, const uint8_t* status_, uint32_t cap,
                                 uint32_t* hits, uint32_t* n_hits) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  uint32_t n = 0;
  const int8_t sd = strand[q];
  if (sd == '+' || sd == '-') {
    const bool minus = sd == '-';
    const uint32_t qs = qs_[q], qe = qe_[q];
    uint32_t lo, hi; overlap_range(ix, refid[q], minus, qs, qe, lo, hi);
    for (uint32_t i = lo; i < hi; i++) {
      RecKey k = rec_key(ix, i);
      if ((int32_t)qs >= k.en) continue;
      Hit h;
      if (!filter_hit(p, minus, status_[q], qs, qe, (uint32_t)k.st, (uint32_t)k.en, k.pos_start, h)) continue;
      if (n < cap) hits[(size_t)q * cap + n] = i;
      n++;
    }
  }
  n_hits[q] = n;
}

}  // namespace brb
