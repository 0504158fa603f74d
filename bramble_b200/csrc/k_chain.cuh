// Kernel group 2: per-candidate exon-chain evaluation (one thread per candidate transcript).
// The reference evaluates exon-major over an insertion-ordered map (src/evaluate.cpp:184-282,
// 888-1134); here every candidate (alignment, strand, exon-0 hit) walks the read's exons on its own,
// re-reading the already located index ranges, which is equivalent because candidates only interact
// through (a) the strand-level failure decided in k_lookup and (b) duplicate tids, handled below.
#pragma once
#include "dev_common.cuh"

namespace brb {

// Ideal-CIGAR / score accumulator: ExonChainMatch + AlignInfo, include/evaluate.h:152-181.
struct Builder {
  uint32_t* cig; uint32_t n; uint32_t prev_op;
  double cov, ops; int32_t ref, junc, clip;
  __device__ __forceinline__ void init(uint32_t* pool) { cig = pool; n = 0; prev_op = CM; cov = 0.0; ops = 0.0; ref = 0; junc = 0; clip = 0; }
  __device__ __forceinline__ void add(uint32_t len, uint32_t op) {          // Cigar::add_operation, include/evaluate.h:108-126
    if (n > 0 && (cig[n - 1] & 0xfu) == op) cig[n - 1] = (((cig[n - 1] >> 4) + len) << 4) | op;
    else cig[n++] = (len << 4) | op;
  }
};

// build_cigar_match, src/evaluate.cpp:675-786
__device__ __forceinline__ void build_match(Builder& m, int st, uint32_t qs, uint32_t qe, uint32_t gs, uint32_t ge, const Hit& h,
                                            bool has_lc, bool has_rc, bool first_match, bool last_match) {
  if (h.left_ins > 0) {
    if (st == FIRST_EXON || st == ONLY_EXON) {
      if (!has_lc) { m.add(h.left_ins, CS); m.ops += h.left_ins; m.prev_op = CS; }
    } else {   // MIDDLE / LAST
      m.add(h.left_ins, CI); m.ops += h.left_ins;
      if (m.prev_op == CD) m.cov += h.left_ins;
      else if (m.prev_op == CI) m.ops += (m.ops * 0.2);
      m.prev_op = CI;
    }
  } else if (h.left_gap > 0) {
    if (!first_match && (st == MIDDLE_EXON || st == LAST_EXON || has_lc)) {
      m.add(h.left_gap, CD); m.ops += h.left_gap; m.ref += (int32_t)h.left_gap;
      if (m.prev_op == CI) m.cov += h.left_gap;
      else if (m.prev_op == CD) m.ops += (m.ops * 0.2);
      m.prev_op = CD;
    }
  } else m.junc++;
  const uint32_t os = qs > gs ? qs : gs, oe = qe < ge ? qe : ge;
  if (oe >= os) { const uint32_t len = oe - os; m.add(len, CM); m.ops += len; m.cov += len; m.ref += (int32_t)len; m.prev_op = CM; }
  if (h.right_ins > 0) {
    if (st == LAST_EXON || st == ONLY_EXON) {
      if (!has_rc) { m.add(h.right_ins, CS); m.ops += h.right_ins; m.prev_op = CS; }
    } else {   // FIRST / MIDDLE
      m.add(h.right_ins, CI); m.ops += h.right_ins;
      if (m.prev_op == CD) m.cov += h.right_ins;
      m.prev_op = CI;
    }
  } else if (h.right_gap > 0) {
    if (!last_match && (st == FIRST_EXON || st == MIDDLE_EXON || has_rc)) {
      m.add(h.right_gap, CD); m.ops += h.right_gap; m.ref += (int32_t)h.right_gap;
      if (m.prev_op == CI) m.cov += h.right_gap;
      m.prev_op = CD;
    }
  } else m.junc++;
}

struct RescueSeg {       // outcome of one clip rescue (k_ksw.cuh); ok==0 -> no clip segment
  uint32_t ok; int32_t score; int32_t rc; uint32_t cig_off, cig_len;
};

struct WalkOut {         // pass-1 facts about a surviving candidate
  uint32_t first_rec, last_rec, n_seg, n_guide, cap;
};

// Scan exon j's hit range for records of transcript `tid` that pass the filters.
__device__ __forceinline__ uint32_t find_tid_hits(const DevIndex& ix, const brb_params& p, bool minus, int status, uint32_t qs, uint32_t qe,
                                                  uint32_t lo, uint32_t hi, uint32_t tid, uint32_t& rec, RecKey& key, Hit& hit) {
  uint32_t cnt = 0;
  for (uint32_t i = lo; i < hi; i++) {
    RecKey k = rec_key(ix.rec, i);
    if (k.tid != tid || (int32_t)qs >= k.en) continue;
    Hit h;
    if (!filter_hit(p, minus, status, qs, qe, (uint32_t)k.st, (uint32_t)k.en, k.pos_start, h)) continue;
    if (cnt == 0) { rec = i; key = k; hit = h; }
    cnt++;
  }
  return cnt;
}

// The chain walk.  BUILD=false: pass 1 (continuity, injectivity, counts; src/evaluate.cpp:111-182, 1004-1065).
// BUILD=true: pass 2 (ideal CIGAR and score accumulators; :1067-1109).  Returns false if eliminated.
template <bool BUILD>
__device__ bool walk(const DevIndex& ix, const DevBatch& b, const brb_params& p, const uint32_t* __restrict__ q_lo,
                     const uint32_t* __restrict__ q_hi, uint32_t a, int s, uint32_t first_rec, WalkOut& w,
                     Builder* m, const RescueSeg* lres, const RescueSeg* rres, const uint32_t* __restrict__ rpool,
                     uint32_t& first_pos, uint32_t& last_pos) {
  const bool minus = s;
  const uint32_t e0 = __ldg(&b.exon_off[a]), n_ex = __ldg(&b.exon_off[a + 1]) - e0;
  const int32_t refid = __ldg(&b.refid[a]);
  const bool has_lc = BUILD && lres && lres->ok, has_rc = BUILD && rres && rres->ok;
  uint32_t k = 0;                                   // segment index
  uint32_t n_seg = 0, n_guide = 0, cap = 0;
  uint32_t last_idx = BUILD ? w.n_guide - 1 : 0;    // last_match_idx: a COUNTER compared with the segment index (:1010-1011,1075-1076)
  uint32_t n_seg_total = BUILD ? w.n_seg : 0;
  // segment 0
  uint32_t qs = __ldg(&b.exon_st[e0]), qe = __ldg(&b.exon_en[e0]);
  RecKey key = rec_key(ix.rec, first_rec);
  const uint32_t tid = key.tid;
  Hit h; int status = exon_status(n_ex, 0);
  filter_hit(p, minus, status, qs, qe, (uint32_t)key.st, (uint32_t)key.en, key.pos_start, h);
  uint32_t prev_gs = (uint32_t)key.st, prev_ge = (uint32_t)key.en, prev_qs = qs, prev_qe = qe;
  uint32_t prev_eid = ix.rec[first_rec].meta & 0xffu;
  uint32_t last_rec = first_rec;
  if (BUILD) {
    if (has_lc) {                                   // LEFTC segment, build_left_clip_segment :397-448 + build_cigar_clip :824-841
      first_pos = key.pos_start - (uint32_t)lres->rc;
      for (uint32_t x = 0; x < lres->cig_len; x++) {
        const uint32_t wd = rpool[lres->cig_off + x], op = wd & 0xfu, len = wd >> 4;
        m->add(len, op); if (op == OV_M || op == OV_D) m->ref += (int32_t)len;
      }
      m->clip += lres->score;
      last_pos = first_pos; k++;
      h.left_ins = 0;                               // :482 gexon.left_ins = 0
    } else first_pos = h.pos;
    if (has_rc && n_ex == 1) h.right_ins = 0;
    last_pos = h.pos;                               // rcpos follows every later guide-bearing segment (:1057-1062)
    build_match(*m, status, qs, qe, (uint32_t)key.st, (uint32_t)key.en, h, has_lc, has_rc, k == 0, k == last_idx);
    k++;
  } else { n_seg = 1; n_guide = 1; cap = 3; }
  for (uint32_t j = 1; j < n_ex; j++) {
    qs = __ldg(&b.exon_st[e0 + j]); qe = __ldg(&b.exon_en[e0 + j]);
    status = exon_status(n_ex, j);
    const uint32_t lo = q_lo[2 * (e0 + j) + s], hi = q_hi[2 * (e0 + j) + s];
    if (lo == hi) {                                 // INS_EXON segment (:256-268); build_cigar_ins :788-806
      if (!BUILD) {
        if (qs == prev_qs && qe == prev_qe) return false;
        n_seg++; cap++;
      } else {
        const uint32_t len = qe - qs; const bool edge = (k == 0 || k == n_seg_total - 1);
        if (edge) { m->add(len, CS); m->prev_op = CS; } else { m->add(len, CI); m->prev_op = CI; }
        m->ops += len; m->cov += len; m->junc -= edge ? 1 : 2;
        k++;
      }
      prev_qs = qs; prev_qe = qe;
      continue;
    }
    uint32_t rec; const uint32_t cnt = find_tid_hits(ix, p, minus, status, qs, qe, lo, hi, tid, rec, key, h);
    if (!BUILD) { if (cnt != 1) return false; }     // 0: tid not hit by this exon; >=2: non-injective (same q exon twice)
    const Rec* R = &ix.rec[rec];
    const uint32_t eid = R->meta & 0xffu;
    const uint32_t gap = (eid - prev_eid) & 0xffu;  // uint8_t wrap, src/evaluate.cpp:131
    if (!p.long_reads) { if (!BUILD && gap != 1) return false; }
    else {
      if (!BUILD && gap > 2) return false;
      if (gap == 2) {                               // skipped guide exon -> GAP_EXON segment (:148-176), build_cigar_gap :808-822
        uint4 nb = ldg4(&R->prev_st);
        const uint32_t gs = minus ? nb.z : nb.x, ge = minus ? nb.w : nb.y;
        if (!BUILD) {
          if ((gs == 0 && ge == 0) || (ge - gs > p.max_error_exon)) return false;
        }
        uint32_t grec;
        if (!find_for_tid(ix, refid, minus, tid, gs, ge, grec)) return false;
        RecKey gk = rec_key(ix.rec, grec);
        if (!BUILD) {
          if ((uint32_t)gk.st == prev_gs && (uint32_t)gk.en == prev_ge) return false;
          n_seg++; n_guide++; cap++;
        } else {
          const uint32_t len = (uint32_t)(gk.en - gk.st);
          m->add(len, CD); m->prev_op = CD; m->ops += len; m->cov += len; m->ref += (int32_t)len; m->junc -= 2;
          k++;
        }
        prev_gs = (uint32_t)gk.st; prev_ge = (uint32_t)gk.en;
      }
    }
    if (!BUILD) {
      if ((uint32_t)key.st == prev_gs && (uint32_t)key.en == prev_ge) return false;
      if (qs == prev_qs && qe == prev_qe) return false;
      n_seg++; n_guide++; cap += 3;
    } else {
      if (has_rc && j == n_ex - 1) h.right_ins = 0;  // :631 gexon.right_ins = 0
      last_pos = h.pos;
      build_match(*m, status, qs, qe, (uint32_t)key.st, (uint32_t)key.en, h, has_lc, has_rc, k == 0, k == last_idx);
      k++;
    }
    prev_gs = (uint32_t)key.st; prev_ge = (uint32_t)key.en; prev_qs = qs; prev_qe = qe; prev_eid = eid; last_rec = rec;
  }
  if (BUILD) {
    if (has_rc) {                                   // RIGHTC segment, build_right_clip_segment :548-598
      RecKey lk = rec_key(ix.rec, w.last_rec);
      last_pos = lk.pos_start - (uint32_t)rres->rc;
      for (uint32_t x = 0; x < rres->cig_len; x++) {
        const uint32_t wd = rpool[rres->cig_off + x], op = wd & 0xfu, len = wd >> 4;
        m->add(len, op); if (op == OV_M || op == OV_D) m->ref += (int32_t)len;
      }
      m->clip += rres->score;
    }
    if (m->junc < 0) m->junc = 0;
  } else { w.first_rec = first_rec; w.last_rec = last_rec; w.n_seg = n_seg; w.n_guide = n_guide; w.cap = cap; }
  return true;
}

// Exon-0 duplicate handling: `data[tid] = td` (src/evaluate.cpp:218-224) keeps the FIRST hit's slot
// and the LAST hit's content.  Returns false if this candidate slot is not the first hit of its tid.
__device__ __forceinline__ bool dedupe_first(const DevIndex& ix, const DevBatch& b, const brb_params& p, const uint32_t* q_lo,
                                             const uint32_t* q_hi, uint32_t a, int s, uint32_t& rec0) {
  const uint32_t e0 = __ldg(&b.exon_off[a]), n_ex = __ldg(&b.exon_off[a + 1]) - e0;
  const uint32_t qs = __ldg(&b.exon_st[e0]), qe = __ldg(&b.exon_en[e0]);
  const int status = exon_status(n_ex, 0);
  const uint32_t lo = q_lo[2 * e0 + s], hi = q_hi[2 * e0 + s];
  const uint32_t tid = rec_key(ix.rec, rec0).tid;
  uint32_t last = rec0;
  for (uint32_t i = lo; i < hi; i++) {
    if (i == rec0) continue;
    RecKey k = rec_key(ix.rec, i);
    if (k.tid != tid || (int32_t)qs >= k.en) continue;
    Hit h;
    if (!filter_hit(p, s, status, qs, qe, (uint32_t)k.st, (uint32_t)k.en, k.pos_start, h)) continue;
    if (i < rec0) return false;
    last = i;
  }
  rec0 = last;
  return true;
}

// Pass 1, one thread per candidate.  c_cap[c] = ideal-CIGAR capacity (0 = eliminated).
__global__ void __launch_bounds__(256) k_chain(DevIndex ix, DevBatch b, brb_params p, uint32_t n_cand,
                                               const uint32_t* __restrict__ q_lo, const uint32_t* __restrict__ q_hi,
                                               const uint32_t* __restrict__ c_rec, const uint32_t* __restrict__ c_task,
                                               uint32_t* __restrict__ c_first, uint32_t* __restrict__ c_last,
                                               uint32_t* __restrict__ c_nseg, uint32_t* __restrict__ c_nguide,
                                               uint32_t* __restrict__ c_cap) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cand) return;
  const uint32_t task = c_task[c], a = task >> 1; const int s = task & 1;
  uint32_t rec0 = c_rec[c];
  WalkOut w; w.cap = 0; w.first_rec = rec0; w.last_rec = rec0; w.n_seg = 0; w.n_guide = 0;
  uint32_t fp, lp;
  bool ok = dedupe_first(ix, b, p, q_lo, q_hi, a, s, rec0);
  if (ok) ok = walk<false>(ix, b, p, q_lo, q_hi, a, s, rec0, w, nullptr, nullptr, nullptr, nullptr, fp, lp);
  c_first[c] = w.first_rec; c_last[c] = w.last_rec; c_nseg[c] = w.n_seg; c_nguide[c] = w.n_guide;
  c_cap[c] = ok ? w.cap : 0;
}

// Pass 2 + filter_by_similarity (src/evaluate.cpp:843-865), one thread per surviving candidate.
// Writes the match row: valid flag, tid, position (fwpos or rcpos by strand), scores, ideal CIGAR length.
__device__ __forceinline__ uint32_t build_one(const DevIndex& ix, const DevBatch& b, const brb_params& p, uint32_t c,
                                              const uint32_t* __restrict__ q_lo, const uint32_t* __restrict__ q_hi,
                                              const uint32_t* __restrict__ c_task,
                                              const uint32_t* __restrict__ c_first, const uint32_t* __restrict__ c_last,
                                              const uint32_t* __restrict__ c_nseg, const uint32_t* __restrict__ c_nguide,
                                              const uint32_t* __restrict__ c_cap, const uint32_t* __restrict__ c_ioff,
                                              const RescueSeg* __restrict__ r_left, const RescueSeg* __restrict__ r_right,
                                              const uint32_t* __restrict__ rpool, uint32_t* __restrict__ ipool,
                                              uint8_t* __restrict__ m_valid, uint32_t* __restrict__ m_tid, uint32_t* __restrict__ m_pos,
                                              double* __restrict__ m_sim, int32_t* __restrict__ m_clip, uint32_t* __restrict__ m_ilen) {
  if (c_cap[c] == 0) { m_valid[c] = 0; return 0; }
  const uint32_t task = c_task[c], a = task >> 1; const int s = task & 1;
  WalkOut w; w.first_rec = c_first[c]; w.last_rec = c_last[c]; w.n_seg = c_nseg[c]; w.n_guide = c_nguide[c]; w.cap = c_cap[c];
  RescueSeg L, R; L.ok = 0; R.ok = 0;
  if (r_left) { L = r_left[c]; R = r_right[c]; }
  if (L.ok) { w.n_seg++; w.n_guide++; }
  if (R.ok) { w.n_seg++; w.n_guide++; }
  Builder m; m.init(ipool + c_ioff[c]);
  uint32_t fp = 0, lp = 0;
  walk<true>(ix, b, p, q_lo, q_hi, a, s, w.first_rec, w, &m, r_left ? &L : nullptr, r_left ? &R : nullptr, rpool, fp, lp);
  double sim_score = 0.0; bool keep = true;
  if (p.filter_by_similarity) {
    const double thr = (double)p.similarity_threshold;
    const double sim = (m.ops > 0) ? (m.cov / m.ops) : 0.0;
    if (sim > thr) { const double x = (sim - thr) / (1.0 - thr); sim_score = x * x * (double)(m.junc + 1); }
    else keep = false;
  }
  m_valid[c] = keep ? 1 : 0;
  m_tid[c] = rec_key(ix.rec, w.first_rec).tid;
  m_pos[c] = s ? lp : fp;
  m_sim[c] = sim_score; m_clip[c] = m.clip; m_ilen[c] = m.n;
  return keep ? 1u : 0u;
}

__global__ void __launch_bounds__(256) k_build(DevIndex ix, DevBatch b, brb_params p, uint32_t n_cand,
                                               const uint32_t* __restrict__ q_lo, const uint32_t* __restrict__ q_hi,
                                               const uint32_t* __restrict__ c_task,
                                               const uint32_t* __restrict__ c_first, const uint32_t* __restrict__ c_last,
                                               const uint32_t* __restrict__ c_nseg, const uint32_t* __restrict__ c_nguide,
                                               const uint32_t* __restrict__ c_cap, const uint32_t* __restrict__ c_ioff,
                                               const RescueSeg* __restrict__ r_left, const RescueSeg* __restrict__ r_right,
                                               const uint32_t* __restrict__ rpool, uint32_t* __restrict__ ipool,
                                               uint8_t* __restrict__ m_valid, uint32_t* __restrict__ m_tid, uint32_t* __restrict__ m_pos,
                                               double* __restrict__ m_sim, int32_t* __restrict__ m_clip, uint32_t* __restrict__ m_ilen,
                                               unsigned long long* __restrict__ n_matches) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t kept = 0;
  if (c < n_cand) kept = build_one(ix, b, p, c, q_lo, q_hi, c_task, c_first, c_last, c_nseg, c_nguide, c_cap, c_ioff, r_left, r_right, rpool,
                                   ipool, m_valid, m_tid, m_pos, m_sim, m_clip, m_ilen);
  for (int d = 16; d; d >>= 1) kept += __shfl_down_sync(0xffffffffu, kept, d);
  if ((threadIdx.x & 31) == 0 && kept) atomicAdd(n_matches, (unsigned long long)kept);
}

}  // namespace brb
