// Kernel group 2: per-candidate exon-chain evaluation (one thread per candidate transcript).
// The reference evaluates exon-major over an insertion-ordered map (src/evaluate.cpp:184-282,
// 888-1134); here every candidate (alignment, strand, exon-0 hit) walks the read's exons on its own.
// That is equivalent because candidates only interact through (a) the strand-level failure decided in
// k_lookup and (b) duplicate tids, handled below.  The walk follows the transcript's own exon chain
// (Rec::gnext) instead of re-scanning each exon's hit range: the exons of a tid are disjoint and ascending
// (validated at brb_index_create), and read exons are ascending, so every record of the tid that overlaps
// read exon j is reachable from the record matched at exon j-1 by gnext links.
#pragma once
#include "dev_common.cuh"

namespace brb {

// Ideal-CIGAR / score accumulator: ExonChainMatch + AlignInfo, include/evaluate.h:152-181.
// The open (last) op lives in a register and is stored when the next op starts.
struct Builder {
  uint32_t* cig; uint32_t n; uint32_t last; uint32_t prev_op;
  double cov, ops; int32_t ref, junc, clip;
  __device__ __forceinline__ void init(uint32_t* pool) { cig = pool; n = 0; last = 0; prev_op = CM; cov = 0.0; ops = 0.0; ref = 0; junc = 0; clip = 0; }
  __device__ __forceinline__ void add(uint32_t len, uint32_t op) {          // Cigar::add_operation, include/evaluate.h:108-126
    if (n > 0 && (last & 0xfu) == op) last = (((last >> 4) + len) << 4) | op;
    else { if (n > 0) cig[n - 1] = last; last = (len << 4) | op; n++; }
  }
  __device__ __forceinline__ void finish() { if (n > 0) cig[n - 1] = last; }
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

struct ReadExons {       // the read side of a candidate: exon list of the alignment and its first exon
  uint32_t e0, n_ex, qs0, qe0;
};
__device__ __forceinline__ ReadExons read_exons(const DevBatch& b, uint32_t a) {
  ReadExons r; r.e0 = __ldg(&b.exon_off[a]); r.n_ex = __ldg(&b.exon_off[a + 1]) - r.e0;
  r.qs0 = __ldg(&b.exon_st[r.e0]); r.qe0 = __ldg(&b.exon_en[r.e0]); return r;
}

// The chain walk.  BUILD=false: pass 1 (continuity, injectivity, counts; src/evaluate.cpp:111-182, 1004-1065).
// BUILD=true: pass 2 (ideal CIGAR and score accumulators; :1067-1109).  Returns false if eliminated.
template <bool BUILD>
__device__ bool walk(const DevIndex& ix, const DevBatch& b, const brb_params& p, const uint8_t* q_any,
                     const ReadExons& rx, int s, uint32_t first_rec, WalkOut& w,
                     Builder* m, const RescueSeg* lres, const RescueSeg* rres, const uint32_t* __restrict__ rpool,
                     uint32_t& first_pos, uint32_t& last_pos) {
  const bool minus = s;
  const uint32_t e0 = rx.e0, n_ex = rx.n_ex;
  const bool has_lc = BUILD && lres && lres->ok, has_rc = BUILD && rres && rres->ok;
  uint32_t k = 0;                                   // segment index
  uint32_t n_seg = 0, n_guide = 0, cap = 0;
  uint32_t last_idx = BUILD ? w.n_guide - 1 : 0;    // last_match_idx: a COUNTER compared with the segment index (:1010-1011,1075-1076)
  uint32_t n_seg_total = BUILD ? w.n_seg : 0;
  // segment 0
  uint32_t qs = rx.qs0, qe = rx.qe0;
  RecKey key = rec_key(ix, first_rec);
  RecLink lk = rec_link(ix, first_rec);
  Hit h; int status = exon_status(n_ex, 0);
  filter_hit(p, minus, status, qs, qe, (uint32_t)key.st, (uint32_t)key.en, key.pos_start, h);
  uint32_t prev_gs = (uint32_t)key.st, prev_ge = (uint32_t)key.en, prev_qs = qs, prev_qe = qe;
  uint32_t prev_eid = lk.meta & 0xffu;
  uint32_t cur = first_rec;                         // record of the last guide-bearing segment (key/lk describe it)
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
    // k_front records "no surviving hit" only for MIDDLE exons (any other hit-less exon failed the strand); the
    // flag may have been written by another lane of this very kernel, hence the volatile load
    if (p.ignore_small_exons && status == MIDDLE_EXON && !*(volatile const uint8_t*)(q_any + 2 * (e0 + j) + s)) {   // INS_EXON segment (:256-268); build_cigar_ins :788-806
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
    // surviving hits of this transcript on read exon j: `cur` itself (a guide exon spanning the read's
    // intron) and its genomic successors that start before qe.  Exactly one must survive
    // (0: tid not hit by this exon; >=2: non-injective, src/evaluate.cpp:1031-1052).
    uint32_t cnt = 0, rec = cur; RecKey nk = key; RecLink nl = lk;
    if ((int32_t)qs < key.en) {
      Hit hh;
      if (filter_hit(p, minus, status, qs, qe, (uint32_t)key.st, (uint32_t)key.en, key.pos_start, hh)) { cnt = 1; h = hh; }
    }
    {
      uint32_t r = lk.gnext; int32_t r_st = lk.gnext_st;
      while (r != NONE && r_st < (int32_t)qe) {
        const RecKey K = rec_key(ix, r); const RecLink Lr = rec_link(ix, r);
        if ((int32_t)qs < K.en) {
          Hit hh;
          if (filter_hit(p, minus, status, qs, qe, (uint32_t)K.st, (uint32_t)K.en, K.pos_start, hh)) {
            if (cnt == 0) { rec = r; nk = K; nl = Lr; h = hh; }
            cnt++;
          }
        }
        r = Lr.gnext; r_st = Lr.gnext_st;
      }
    }
    if (cnt != 1) return false;                     // (cannot happen in pass 2)
    const uint32_t eid = nl.meta & 0xffu;
    const uint32_t gap = (eid - prev_eid) & 0xffu;  // uint8_t wrap, src/evaluate.cpp:131
    if (!p.long_reads) { if (!BUILD && gap != 1) return false; }
    else {
      if (!BUILD && gap > 2) return false;
      if (gap == 2) {                               // skipped guide exon -> GAP_EXON segment (:148-176), build_cigar_gap :808-822
        // prev_* ('+') / next_* ('-') of seg2 = the genomically previous exon of the same transcript;
        // getGuideExonForTid on its own coordinates returns that very record
        const uint32_t grec = __ldg(&ix.gprev[rec]);
        if (grec == NONE) return false;             // coordinates (0,0)
        const RecKey gk = rec_key(ix, grec);
        if (!BUILD) {
          if ((uint32_t)(gk.en - gk.st) > p.max_error_exon) return false;
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
    key = nk; lk = nl; cur = rec;
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
    prev_gs = (uint32_t)key.st; prev_ge = (uint32_t)key.en; prev_qs = qs; prev_qe = qe; prev_eid = eid;
  }
  if (BUILD) {
    if (has_rc) {                                   // RIGHTC segment, build_right_clip_segment :548-598
      RecKey lastk = rec_key(ix, w.last_rec);
      last_pos = lastk.pos_start - (uint32_t)rres->rc;
      for (uint32_t x = 0; x < rres->cig_len; x++) {
        const uint32_t wd = rpool[rres->cig_off + x], op = wd & 0xfu, len = wd >> 4;
        m->add(len, op); if (op == OV_M || op == OV_D) m->ref += (int32_t)len;
      }
      m->clip += rres->score;
    }
    if (m->junc < 0) m->junc = 0;
    m->finish();
  } else { w.first_rec = first_rec; w.last_rec = cur; w.n_seg = n_seg; w.n_guide = n_guide; w.cap = cap; }
  return true;
}

// Exon-0 duplicate handling: `data[tid] = td` (src/evaluate.cpp:218-224) keeps the FIRST hit's slot
// and the LAST hit's content.  Returns false if this candidate slot is not the first surviving hit of its
// tid.  Duplicates can only be genomic neighbours of rec0 in the same transcript that also overlap read
// exon 0, so the common case is decided by two compares on rec0's own link words.
__device__ __forceinline__ bool dedupe_first(const DevIndex& ix, const brb_params& p, const ReadExons& rx, int s, uint32_t& rec0) {
  const uint32_t n_ex = rx.n_ex, qs = rx.qs0, qe = rx.qe0;
  const RecLink lk = rec_link(ix, rec0);
  if (lk.gprev_en <= (int32_t)qs && lk.gnext_st >= (int32_t)qe) return true;
  const int status = exon_status(n_ex, 0);
  for (uint32_t r = __ldg(&ix.gprev[rec0]); r != NONE; r = __ldg(&ix.gprev[r])) {      // an earlier surviving hit owns the slot
    const RecKey k = rec_key(ix, r);
    if (k.en <= (int32_t)qs) break;
    if (passes_filter(p, s, status, qs, qe, (uint32_t)k.st, (uint32_t)k.en)) return false;
  }
  uint32_t r = lk.gnext; int32_t r_st = lk.gnext_st;
  while (r != NONE && r_st < (int32_t)qe) {                                            // the last surviving hit provides the content
    const RecKey k = rec_key(ix, r); const RecLink l2 = rec_link(ix, r);
    if ((int32_t)qs < k.en && passes_filter(p, s, status, qs, qe, (uint32_t)k.st, (uint32_t)k.en)) rec0 = r;
    r = l2.gnext; r_st = l2.gnext_st;
  }
  return true;
}

struct Matches {          // per candidate slot
  uint4* rec;             // .x pos  .y ideal-CIGAR length << 1 | strand  .z its offset in the ideal pool, or the CIGAR word itself when the
                          // length is 1  .w tid (NONE marks "no match")
  double* sim; int32_t* clip;     // null when every score is 0 by construction
};

// Pass 2 for one candidate whose pass-1 facts are in `w`: ideal CIGAR into pool, scores, similarity filter
// (src/evaluate.cpp:843-865, FP64, operation order kept), match row.
__device__ __forceinline__ bool build_one(const DevIndex& ix, const DevBatch& b, const brb_params& p, uint32_t c, uint32_t task, const ReadExons& rx, int s,
                                          const uint8_t* q_any, WalkOut w, const RescueSeg* L, const RescueSeg* R,
                                          const uint32_t* __restrict__ rpool, uint32_t* __restrict__ ipool, uint32_t ioff, const Matches& M) {
  if (L && L->ok) { w.n_seg++; w.n_guide++; }
  if (R && R->ok) { w.n_seg++; w.n_guide++; }
  Builder m; m.init(ipool + ioff);
  uint32_t fp = 0, lp = 0;
  walk<true>(ix, b, p, q_any, rx, s, w.first_rec, w, &m, L, R, rpool, fp, lp);
  double sim_score = 0.0; bool keep = true;
  if (p.filter_by_similarity) {
    const double thr = (double)p.similarity_threshold;
    const double sim = (m.ops > 0) ? (m.cov / m.ops) : 0.0;
    if (sim > thr) { const double x = (sim - thr) / (1.0 - thr); sim_score = x * x * (double)(m.junc + 1); }
    else keep = false;
  }
  if (keep) {
    M.rec[c] = make_uint4(s ? lp : fp, (m.n << 1) | (uint32_t)s, m.n == 1 ? m.last : ioff, rec_key(ix, w.first_rec).tid);
    if (M.sim) { M.sim[c] = sim_score; M.clip[c] = m.clip; }
  } else M.rec[c] = make_uint4(0, 0, 0, NONE);
  return keep;
}

// Split form for the clip-rescue path (long reads with a genome): pass 1 here, then k_rescue_* / k_ksw
// (which add their CIGAR lengths to c_cap), then k_build.
struct CandFacts { uint32_t* first; uint32_t* last; uint32_t* nseg; uint32_t* nguide; uint32_t* cap; };

__global__ void __launch_bounds__(256) k_chain(DevIndex ix, DevBatch b, brb_params p, const Ctl* __restrict__ ctl,
                                               const uint8_t* q_any, const uint32_t* __restrict__ c_rec,
                                               const uint32_t* __restrict__ c_task, CandFacts F) {
  if (ctl_overflow(ctl)) return;
  const uint32_t n_cand = ctl->n_cand;
  for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < n_cand; c += gridDim.x * blockDim.x) {
    const uint32_t task = c_task[c], a = task >> 1; const int s = task & 1;
    uint32_t rec0 = c_rec[c];
    WalkOut w; w.cap = 0; w.first_rec = rec0; w.last_rec = rec0; w.n_seg = 0; w.n_guide = 0;
    uint32_t fp, lp;
    const ReadExons rx = read_exons(b, a);
    bool ok = dedupe_first(ix, p, rx, s, rec0);
    if (ok) ok = walk<false>(ix, b, p, q_any, rx, s, rec0, w, nullptr, nullptr, nullptr, nullptr, fp, lp);
    F.first[c] = w.first_rec; F.last[c] = w.last_rec; F.nseg[c] = w.n_seg; F.nguide[c] = w.n_guide;
    F.cap[c] = ok ? w.cap : 0;
  }
}

__global__ void __launch_bounds__(256) k_build(DevIndex ix, DevBatch b, brb_params p, Ctl* __restrict__ ctl,
                                               const uint8_t* q_any, const uint32_t* __restrict__ c_task, CandFacts F,
                                               const RescueSeg* __restrict__ r_left, const RescueSeg* __restrict__ r_right,
                                               const uint32_t* __restrict__ rpool, uint32_t* __restrict__ ipool, Matches M) {
  if (ctl_overflow(ctl)) return;
  const unsigned lane = threadIdx.x & 31;
  const uint32_t n_cand = ctl->n_cand;
  const uint32_t stride = gridDim.x * blockDim.x;
  uint32_t kept = 0;
  WarpPool wp; wp.cur = 0; wp.end = 0;
  const uint32_t cap_ideal = ctl->cap_ideal;
  for (uint32_t base = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); base < n_cand; base += stride) {
    const uint32_t c = base + lane;
    const bool act = c < n_cand;
    const uint32_t cap = act ? F.cap[c] : 0;
    bool fits;
    const uint32_t ioff = warp_pool_alloc(wp, &ctl->n_ideal, cap_ideal, cap, lane, fits);
    if (!act) continue;
    if (cap == 0 || !fits) { M.rec[c] = make_uint4(0, 0, 0, NONE); continue; }
    const uint32_t task = c_task[c], a = task >> 1; const int s = task & 1;
    WalkOut w; w.first_rec = F.first[c]; w.last_rec = F.last[c]; w.n_seg = F.nseg[c]; w.n_guide = F.nguide[c]; w.cap = cap;
    RescueSeg L, R; L.ok = 0; R.ok = 0;
    if (r_left) { L = r_left[c]; R = r_right[c]; }
    kept += build_one(ix, b, p, c, task, read_exons(b, a), s, q_any, w, r_left ? &L : nullptr, r_left ? &R : nullptr, rpool, ipool, ioff, M) ? 1u : 0u;
  }
  for (int d = 16; d; d >>= 1) kept += __shfl_down_sync(0xffffffffu, kept, d);
  if (lane == 0 && kept) atomicAdd(&ctl->n_matches, (unsigned long long)kept);
}

}  // namespace brb

namespace brb {

// ---- short-read specialisation -----------------------------------------------------------------------------
// With the short-read presets (ShortReadEvaluator, src/evaluate.cpp:1136-1175: !LONG_READS, max_error_exon == 0)
// there are no INS_EXON / GAP_EXON / clip segments, so a surviving candidate has exactly one guide-bearing
// segment per read exon: n_seg = n_guide = n_ex is known up front, and pass 1 (elimination) and pass 2
// (ideal CIGAR + scores) collapse into one walk.  SIM=false removes the FP64 accumulators, which only feed
// filter_by_similarity.
template <bool SIM>
struct ShortBuilder {
  uint32_t* cig; uint32_t n, last, prev_op; int32_t junc; double cov, ops;
  __device__ __forceinline__ void init(uint32_t* pool) { cig = pool; n = 0; last = 0; prev_op = CM; junc = 0; cov = 0.0; ops = 0.0; }
  __device__ __forceinline__ void add(uint32_t len, uint32_t op) {
    if (n > 0 && (last & 0xfu) == op) last = (((last >> 4) + len) << 4) | op;
    else { if (n > 0) cig[n - 1] = last; last = (len << 4) | op; n++; }
  }
  __device__ __forceinline__ void finish() { if (n > 1) cig[n - 1] = last; }      // a one-word CIGAR stays in the match record
  __device__ __forceinline__ void add_ops(uint32_t v) { if (SIM) ops += v; }
  __device__ __forceinline__ void add_cov(uint32_t v) { if (SIM) cov += v; }
  __device__ __forceinline__ void penal() { if (SIM) ops += (ops * 0.2); }
};

// build_cigar_match (src/evaluate.cpp:675-786) without clip segments (has_left_clip = has_right_clip = false)
template <bool SIM>
__device__ __forceinline__ void build_match_short(ShortBuilder<SIM>& m, int st, uint32_t qs, uint32_t qe, uint32_t gs, uint32_t ge, const Hit& h,
                                                  bool first_match, bool last_match) {
  if (h.left_ins > 0) {
    if (st == FIRST_EXON || st == ONLY_EXON) { m.add(h.left_ins, CS); m.add_ops(h.left_ins); m.prev_op = CS; }
    else {
      m.add(h.left_ins, CI); m.add_ops(h.left_ins);
      if (m.prev_op == CD) m.add_cov(h.left_ins); else if (m.prev_op == CI) m.penal();
      m.prev_op = CI;
    }
  } else if (h.left_gap > 0) {
    if (!first_match && (st == MIDDLE_EXON || st == LAST_EXON)) {
      m.add(h.left_gap, CD); m.add_ops(h.left_gap);
      if (m.prev_op == CI) m.add_cov(h.left_gap); else if (m.prev_op == CD) m.penal();
      m.prev_op = CD;
    }
  } else m.junc++;
  const uint32_t os = qs > gs ? qs : gs, oe = qe < ge ? qe : ge;
  if (oe >= os) { const uint32_t len = oe - os; m.add(len, CM); m.add_ops(len); m.add_cov(len); m.prev_op = CM; }
  if (h.right_ins > 0) {
    if (st == LAST_EXON || st == ONLY_EXON) { m.add(h.right_ins, CS); m.add_ops(h.right_ins); m.prev_op = CS; }
    else { m.add(h.right_ins, CI); m.add_ops(h.right_ins); if (m.prev_op == CD) m.add_cov(h.right_ins); m.prev_op = CI; }
  } else if (h.right_gap > 0) {
    if (!last_match && (st == FIRST_EXON || st == MIDDLE_EXON)) {
      m.add(h.right_gap, CD); m.add_ops(h.right_gap); if (m.prev_op == CI) m.add_cov(h.right_gap); m.prev_op = CD;
    }
  } else m.junc++;
}

// One candidate, one walk.  `key`/`lk` describe rec0 (already the duplicate-rule winner).  Returns false when the
// candidate is eliminated; otherwise the ideal CIGAR is in cig[0..ilen) and pos / sim_score are set.
template <bool SIM>
__device__ __forceinline__ bool eval_short(const DevIndex& ix, const DevBatch& b, const brb_params& p, const ReadExons& rx, int s,
                                           RecKey key, RecLink lk, uint32_t* __restrict__ cig, uint32_t& pos, uint32_t& ilen, uint32_t& word,
                                           double& sim_score) {
  const bool minus = s;
  const uint32_t n_ex = rx.n_ex;
  ShortBuilder<SIM> m; m.init(cig);
  uint32_t qs = rx.qs0, qe = rx.qe0;
  int status = exon_status(n_ex, 0);
  Hit h;
  filter_hit(p, minus, status, qs, qe, (uint32_t)key.st, (uint32_t)key.en, key.pos_start, h);
  const uint32_t first_pos = h.pos; uint32_t last_pos = h.pos;
  if (n_ex > 1 && lk.gnext != NONE) asm volatile("prefetch.global.L1 [%0];" ::"l"(rec_key_ptr(ix, lk.gnext)));   // the walk continues there
  build_match_short<SIM>(m, status, qs, qe, (uint32_t)key.st, (uint32_t)key.en, h, true, n_ex == 1);
  uint32_t prev_eid = lk.meta & 0xffu;
  for (uint32_t j = 1; j < n_ex; j++) {
    const uint32_t prev_qs = qs, prev_qe = qe;
    qs = __ldg(&b.exon_st[rx.e0 + j]); qe = __ldg(&b.exon_en[rx.e0 + j]);
    status = exon_status(n_ex, j);
    // surviving hits of this transcript on read exon j (see walk<>): exactly one, and not the previous record again
    uint32_t cnt = 0; RecKey nk = key; RecLink nl = lk;
    if ((int32_t)qs < key.en && passes_filter(p, minus, status, qs, qe, (uint32_t)key.st, (uint32_t)key.en)) return false;   // same guide exon twice: gap 0
    uint32_t r = lk.gnext; int32_t r_st = lk.gnext_st;
    while (r != NONE && r_st < (int32_t)qe) {
      const RecKey K = rec_key(ix, r); const RecLink Lr = rec_link(ix, r);
      if ((int32_t)qs < K.en) {
        Hit hh;
        if (filter_hit(p, minus, status, qs, qe, (uint32_t)K.st, (uint32_t)K.en, K.pos_start, hh)) { if (cnt == 0) { nk = K; nl = Lr; h = hh; } cnt++; }
      }
      r = Lr.gnext; r_st = Lr.gnext_st;
    }
    if (cnt != 1) return false;
    const uint32_t eid = nl.meta & 0xffu;
    if (((eid - prev_eid) & 0xffu) != 1) return false;            // correct_for_gaps, short reads: uint8_t gap != 1 (src/evaluate.cpp:131-140)
    if (qs == prev_qs && qe == prev_qe) return false;             // injectivity on the read side (:1031-1052); guide side differs since gap == 1
    key = nk; lk = nl; prev_eid = eid;
    last_pos = h.pos;
    build_match_short<SIM>(m, status, qs, qe, (uint32_t)key.st, (uint32_t)key.en, h, false, j == n_ex - 1);
  }
  m.finish();
  if (m.junc < 0) m.junc = 0;
  sim_score = 0.0;
  if (SIM) {                                                      // filter_by_similarity, src/evaluate.cpp:843-865 (FP64, operation order kept)
    const double thr = (double)p.similarity_threshold;
    const double sim = (m.ops > 0) ? (m.cov / m.ops) : 0.0;
    if (!(sim > thr)) return false;
    const double x = (sim - thr) / (1.0 - thr); sim_score = x * x * (double)(m.junc + 1);
  }
  pos = minus ? last_pos : first_pos; ilen = m.n; word = m.last;
  return true;
}

}  // namespace brb
