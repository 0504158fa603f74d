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

struct FlatOut { uint32_t nf, first, last, nhit; };   // survivors, first survivor, last survivor + 1, overlapping records

// Warp-cooperative filtered scan of 32 index ranges.  Lane inputs: lo, n (0 = nothing), qs, qe, status.
// WRITE=false: per-lane FlatOut.  WRITE=true: survivors of lane L's range are appended, in index order, at
// c_rec[dst + k] (k = 0,1,..) together with c_task = task.
template <bool WRITE>
__device__ __forceinline__ void warp_flat_scan(const DevIndex& ix, const brb_params& p, bool minus, unsigned lane,
                                               uint32_t lo, uint32_t n, uint32_t qs, uint32_t qe, int status, FlatOut& o,
                                               uint32_t dst, uint32_t task, uint32_t* __restrict__ c_rec, uint32_t* __restrict__ c_task) {
  const uint32_t incl = warp_incl_scan(n, lane), excl = incl - n;
  const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
  o.nf = 0; o.first = 0; o.last = 0; o.nhit = 0;
  uint32_t written = 0;
  for (uint32_t base = 0; base < total; base += 32) {
    const uint32_t item = base + lane;
    const bool act = item < total;
    const unsigned L = warp_owner(incl, act ? item : total - 1);
    const uint32_t o_lo = __shfl_sync(0xffffffffu, lo, L), o_ex = __shfl_sync(0xffffffffu, excl, L);
    const uint32_t o_qs = __shfl_sync(0xffffffffu, qs, L), o_qe = __shfl_sync(0xffffffffu, qe, L);
    const int o_st = __shfl_sync(0xffffffffu, status, L);
    const uint32_t i = o_lo + (item - o_ex);
    bool ov = false, sv = false;
    if (act) {
      const RecKey k = rec_key(ix.rec, i);
      ov = (int32_t)o_qs < k.en;                           // st < qe holds for every record below hi
      sv = ov && passes_filter(p, minus, o_st, o_qs, o_qe, (uint32_t)k.st, (uint32_t)k.en);
    }
    const uint32_t B = __ballot_sync(0xffffffffu, sv);
    // bits of this chunk that belong to MY range
    uint32_t mine = 0;
    if (n && incl > base && excl < base + 32) mine = bit_range(excl > base ? excl - base : 0, incl - base < 32 ? incl - base : 32);
    if (!WRITE) {
      const uint32_t H = __ballot_sync(0xffffffffu, ov);
      const uint32_t m = B & mine;
      if (m) {
        if (o.nf == 0) o.first = lo + (base + (uint32_t)__ffs(m) - 1 - excl);
        o.last = lo + (base + 31u - (uint32_t)__clz(m) - excl) + 1;
        o.nf += __popc(m);
      }
      o.nhit += __popc(H & mine);
    } else {
      const uint32_t o_mine = __shfl_sync(0xffffffffu, mine, L), o_wr = __shfl_sync(0xffffffffu, written, L);
      const uint32_t o_dst = __shfl_sync(0xffffffffu, dst, L), o_task = __shfl_sync(0xffffffffu, task, L);
      if (sv) {
        const uint32_t pos = o_dst + o_wr + __popc(B & o_mine & ((1u << lane) - 1u));
        c_rec[pos] = i; c_task[pos] = o_task;
      }
      written += __popc(B & mine);
    }
  }
}

// Per alignment: where its candidates live (both strands, '+' first) -- written by k_lookup, read by everyone after.
//   .x = first candidate slot, .y = candidates of task 2a, .z = candidates of task 2a+1
__global__ void __launch_bounds__(256) k_lookup(DevIndex ix, DevBatch b, brb_params p, Ctl* __restrict__ ctl,
                                                uint8_t* __restrict__ q_any, uint4* __restrict__ a_cand,
                                                uint32_t* __restrict__ c_rec, uint32_t* __restrict__ c_task) {
  const unsigned lane = threadIdx.x & 31;
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = a < b.n_aln;
  uint32_t e0 = 0, n_ex = 0; int32_t refid = -1; bool act[2] = {false, false};
  uint32_t bad = 0; bool has_mate = false;
  if (valid) {
    e0 = __ldg(&b.exon_off[a]); n_ex = __ldg(&b.exon_off[a + 1]) - e0; refid = __ldg(&b.refid[a]);
    const int8_t sd = b.strand[a];
    act[0] = task_active(p, sd, 0); act[1] = task_active(p, sd, 1);
    const int32_t m = __ldg(&b.mate_idx[a]);               // process_pairs links are single and symmetric (src/bramble.cpp:272-311)
    if (m >= 0 && ((uint32_t)m >= b.n_aln || (uint32_t)m == a || __ldg(&b.mate_idx[m]) != (int32_t)a)) bad = 1;
    has_mate = m >= 0;
  }
  bool fail[2]; uint32_t ncand[2] = {0, 0}, clo[2] = {0, 0}, chi[2] = {0, 0};
  uint32_t st0 = 0, en0 = 0;                                // exon 0 (kept for the fill pass)
  unsigned long long my_hits = 0;
#pragma unroll
  for (int s = 0; s < 2; s++) {
    fail[s] = (n_ex == 0);
    const uint32_t maxj = __reduce_max_sync(0xffffffffu, (valid && (act[s] || s == 0)) ? n_ex : 0u);
    uint32_t prev_en = 0;
    for (uint32_t j = 0; j < maxj; j++) {
      const bool in = valid && j < n_ex;
      const bool have = in && act[s] && !fail[s];
      uint32_t qs = 0, qe = 0, lo = 0, n = 0; int status = 0;
      if (in && (have || s == 0)) {
        qs = __ldg(&b.exon_st[e0 + j]); qe = __ldg(&b.exon_en[e0 + j]);
        if (s == 0) { if (qe < qs || (j > 0 && qs < prev_en)) bad = 2; prev_en = qe; }   // read exons ascending and disjoint (CIGAR order)
        if (j == 0) { st0 = qs; en0 = qe; }
      }
      if (have) {
        status = exon_status(n_ex, j);
        uint32_t hi; overlap_range(ix, refid, s, qs, qe, lo, hi); n = hi - lo;
      }
      FlatOut f;
      warp_flat_scan<false>(ix, p, s, lane, lo, n, qs, qe, status, f, 0, 0, nullptr, nullptr);
      if (have) {
        my_hits += f.nhit;
        q_any[2 * (e0 + j) + s] = f.nf ? 1 : 0;
        if (j == 0) { ncand[s] = f.nf; clo[s] = f.first; chi[s] = f.last; }
        if (f.nf == 0) {
          // src/evaluate.cpp:249-281: only a small MIDDLE exon may be skipped (data is non-empty for j>0)
          const bool small = (qe - qs <= p.max_error_exon);
          if (!(status == MIDDLE_EXON && p.ignore_small_exons && small)) fail[s] = true;
        }
      }
    }
    if (fail[s] || !act[s]) ncand[s] = 0;
  }
  // one slot allocation per warp: candidates of a warp's alignments are contiguous, '+' task before '-' task
  const uint32_t tot = ncand[0] + ncand[1];
  const uint32_t incl = warp_incl_scan(tot, lane), wtot = __shfl_sync(0xffffffffu, incl, 31);
  uint32_t base = 0;
  if (lane == 31 && wtot) base = atomicAdd(&ctl->n_cand, wtot);
  base = __shfl_sync(0xffffffffu, base, 31);
  const uint32_t off0 = base + incl - tot, off1 = off0 + ncand[0];
  if (valid) a_cand[a] = make_uint4(off0, ncand[0], ncand[1], 0);
  const bool fits = (unsigned long long)base + wtot <= ctl->cap_cand;     // cap_cand is constant during the launch
  if (wtot && fits) {
    const int st_code = exon_status(n_ex ? n_ex : 1, 0);
#pragma unroll
    for (int s = 0; s < 2; s++) {
      FlatOut f;
      warp_flat_scan<true>(ix, p, s, lane, clo[s], ncand[s] ? chi[s] - clo[s] : 0, st0, en0, st_code, f, s ? off1 : off0, 2 * a + s, c_rec, c_task);
    }
  }
  // work counter (feeds the algorithmic-bytes formula) + input validation, one atomic per warp
  for (int d = 16; d; d >>= 1) my_hits += __shfl_down_sync(0xffffffffu, my_hits, d);
  if (lane == 0 && my_hits) atomicAdd(&ctl->n_hits, my_hits);
  const uint32_t nm = __popc(__ballot_sync(0xffffffffu, has_mate));
  if (lane == 0 && nm) atomicAdd(&ctl->n_mates, (unsigned long long)nm);
  const uint32_t anybad = __reduce_max_sync(0xffffffffu, bad);
  if (lane == 0 && anybad) atomicMax(&ctl->bad_input, anybad);
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
      RecKey k = rec_key(ix.rec, i);
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
