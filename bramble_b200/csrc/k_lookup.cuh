// Kernel group 1: exon-overlap lookup on the flattened sorted-interval table.
// Replaces g2tTree::getGuideExons / IntervalTree::findOverlapping (reference src/g2t.cpp:103-257,334-344)
// and the per-exon bookkeeping of get_intervals that does not depend on a candidate
// (src/evaluate.cpp:184-282: "no hit" -> small-middle-exon INS path or failure of the strand).
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

// One thread per task.  For every read exon j: the index range [q_lo,q_hi) tightened to the first /
// last hit that passes the tolerance filters (q_lo == q_hi  <=>  no surviving hit), and for the task
// the number of exon-0 hits (= candidate transcripts before de-duplication) or 0 if the strand fails.
__global__ void __launch_bounds__(256) k_lookup(DevIndex ix, DevBatch b, brb_params p,
                                                uint32_t* __restrict__ q_lo, uint32_t* __restrict__ q_hi,
                                                uint32_t* __restrict__ t_ncand, unsigned long long* __restrict__ n_hits) {
  const uint32_t task = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long my_hits = 0;
  if (task < 2u * b.n_aln) {
    const uint32_t a = task >> 1; const int s = task & 1;
    uint32_t ncand = 0;
    if (task_active(p, b.strand[a], s)) {
      const uint32_t e0 = __ldg(&b.exon_off[a]), n_ex = __ldg(&b.exon_off[a + 1]) - e0;
      const int32_t refid = __ldg(&b.refid[a]);
      bool fail = (n_ex == 0);
      for (uint32_t j = 0; j < n_ex && !fail; j++) {
        const uint32_t qs = __ldg(&b.exon_st[e0 + j]), qe = __ldg(&b.exon_en[e0 + j]);
        const int status = exon_status(n_ex, j);
        uint32_t lo, hi; overlap_range(ix, refid, s, qs, qe, lo, hi);
        uint32_t nf = 0, first = 0, last = 0;
        for (uint32_t i = lo; i < hi; i++) {
          RecKey k = rec_key(ix.rec, i);
          if ((int32_t)qs >= k.en) continue;
          my_hits++;
          Hit h;
          if (!filter_hit(p, s, status, qs, qe, (uint32_t)k.st, (uint32_t)k.en, k.pos_start, h)) continue;
          if (nf == 0) first = i;
          last = i + 1; nf++;
        }
        q_lo[2 * (e0 + j) + s] = nf ? first : 0;
        q_hi[2 * (e0 + j) + s] = nf ? last : 0;
        if (j == 0) ncand = nf;
        if (nf == 0) {
          // src/evaluate.cpp:249-281: only a small MIDDLE exon may be skipped (data is non-empty for j>0)
          const bool small = (qe - qs <= p.max_error_exon);
          if (!(status == MIDDLE_EXON && p.ignore_small_exons && small)) fail = true;
        }
      }
      if (fail) ncand = 0;
    }
    t_ncand[task] = ncand;
  }
  // one atomic per warp for the work counter (feeds the algorithmic-bytes formula)
  for (int d = 16; d; d >>= 1) my_hits += __shfl_down_sync(0xffffffffu, my_hits, d);
  if ((threadIdx.x & 31) == 0 && my_hits) atomicAdd(n_hits, my_hits);
}

// One thread per task: materialise the candidate list (index records of the surviving exon-0 hits, in
// hit order == insertion order of the reference's `data` map, src/evaluate.cpp:218-224).
__global__ void __launch_bounds__(256) k_fill_cand(DevIndex ix, DevBatch b, brb_params p,
                                                   const uint32_t* __restrict__ q_lo, const uint32_t* __restrict__ q_hi,
                                                   const uint32_t* __restrict__ t_coff,
                                                   uint32_t* __restrict__ c_rec, uint32_t* __restrict__ c_task) {
  const uint32_t task = blockIdx.x * blockDim.x + threadIdx.x;
  if (task >= 2u * b.n_aln) return;
  uint32_t off = t_coff[task]; const uint32_t n = t_coff[task + 1] - off;
  if (n == 0) return;
  const uint32_t a = task >> 1; const int s = task & 1;
  const uint32_t e0 = __ldg(&b.exon_off[a]), n_ex = __ldg(&b.exon_off[a + 1]) - e0;
  const uint32_t qs = __ldg(&b.exon_st[e0]), qe = __ldg(&b.exon_en[e0]);
  const int status = exon_status(n_ex, 0);
  const uint32_t lo = q_lo[2 * e0 + s], hi = q_hi[2 * e0 + s];
  for (uint32_t i = lo; i < hi; i++) {
    RecKey k = rec_key(ix.rec, i);
    if ((int32_t)qs >= k.en) continue;
    Hit h;
    if (!filter_hit(p, s, status, qs, qe, (uint32_t)k.st, (uint32_t)k.en, k.pos_start, h)) continue;
    c_rec[off] = i; c_task[off] = task; off++;
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
