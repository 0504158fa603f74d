// TEST INFRASTRUCTURE -- CPU restatement ("port") of the reference's projection hot path.
// NOT product code: bramble_b200/ never links, loads or calls anything in this file.
//
// Parity status: PINNED.  The port is checked (tests/test_oracle_*.py) against
//   * the reference's own golden vectors (bramble-rs/tests/update_cigar.rs,
//     evaluate.rs presets, short_read_projection.rs fixture, ksw2 smoke vector), and
//   * the UNMODIFIED reference compiled into oracle/_ref/bramble_ref (oracle/Makefile), record for
//     record, on seeded synthetic inputs for every mode (default, --fr/--rf, --lr -S, --lr-hq --strict).
//
// Structure follows the reference function by function (exon-major evaluation over an insertion-
// ordered tid map), which is deliberately NOT how the CUDA kernels are organised (candidate-major),
// so the two disagree unless both implement the reference's semantics.
#include "brb_port.h"
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <string>
#include <random>
#include <limits>
#include <vector>

namespace {

enum { FIRST_EXON = 0, MIDDLE_EXON = 1, LAST_EXON = 2, ONLY_EXON = 3, INS_EXON = 4, GAP_EXON = 5, LEFTC_EXON = 6, RIGHTC_EXON = 7 };
enum { CM = 0, CI = 1, CD = 2, CN = 3, CS = 4, CH = 5, CP = 6, CEQ = 7, CX = 8, OV_M = 10, OV_D = 11, OV_I = 12, OV_S = 13 };
const int KSW_NEG_INF = -0x40000000;

// ---- include/evaluate.h:17-126 Cigar::add_operation
struct Cigar {
  std::vector<uint32_t> v;
  void add(uint32_t len, uint8_t op) {
    if (!v.empty() && (v.back() & 0xf) == op) v.back() = ((v.back() >> 4) + len) << 4 | op;
    else v.push_back(len << 4 | op);
  }
};

// ---- include/evaluate.h:130-150 GuideExon
struct GExon {
  uint32_t tid = 0, start = 0, end = 0, pos = 0, pos_start = 0;
  uint8_t exon_id = 0;
  int32_t left_ins = 0, right_ins = 0, left_gap = 0, right_gap = 0;
  bool has_prev = false, has_next = false;
  uint32_t prev_start = 0, prev_end = 0, next_start = 0, next_end = 0, transcript_len = 0;
  int64_t rec = -1;   // index record (gives the exon sequence)
};
struct Segment {  // include/evaluate.h:194-207
  bool has_gexon = false, has_qexon = false;
  GExon gexon; uint32_t qs = 0, qe = 0;
  int status = 0; bool is_small_exon = false;
  Cigar cigar; int score = 0;
};
struct Match {    // include/evaluate.h:152-181 ExonChainMatch + AlignInfo
  uint32_t tid = 0, fwpos = 0, rcpos = 0; char strand = '+';
  Cigar cigar; int clip_score = 0; double similarity_score = 0;
  double total_coverage = 0, total_operations = 0; int32_t ref_consumed = 0, junc_hits = 0, transcript_len = 0;
  uint8_t prev_op = CM;
};
struct TidData {  // include/evaluate.h:209-219
  uint32_t tid; bool elim = false, has_left_clip = false, has_right_clip = false;
  Match match; std::vector<Segment> segments;
};

struct Ctx {
  const brb_index_desc* ix; const brb_params* p; const brb_batch_in* in;
  std::vector<int32_t> pmax;   // prefix max of en within each table
  uint64_t n_hits = 0, n_rescue = 0, n_matches = 0;
};

// ---- cgranges IITree::overlap: all i with st[i] < qe && qs < en[i], ascending.
void overlap(const Ctx& c, int refid, char strand, int32_t qs, int32_t qe, std::vector<uint64_t>& out) {
  out.clear();
  if (refid < 0 || (uint32_t)refid * 2 + 1 >= c.ix->n_tables) return;      // src/g2t.cpp:275-276
  if (strand != '+' && strand != '-') return;
  uint32_t t = (uint32_t)refid * 2 + (strand == '-');
  uint64_t lo = c.ix->table_off[t], hi = c.ix->table_off[t + 1];
  const int32_t* st = c.ix->st; const int32_t* en = c.ix->en;
  uint64_t end = std::lower_bound(st + lo, st + hi, qe) - st;               // first st >= qe
  uint64_t beg = std::upper_bound(c.pmax.begin() + lo, c.pmax.begin() + end, qs) - c.pmax.begin();  // first pmax > qs
  for (uint64_t i = beg; i < end; i++) if (qs < en[i]) out.push_back(i);
}

void fill_gexon(const Ctx& c, uint64_t i, GExon& g) {
  const brb_index_desc* ix = c.ix;
  g.start = ix->st[i]; g.end = ix->en[i]; g.transcript_len = ix->transcript_len[i]; g.pos_start = ix->pos_start[i];
  g.exon_id = ix->exon_id[i]; g.has_prev = ix->has_prev_next[i] & 1; g.has_next = (ix->has_prev_next[i] >> 1) & 1;
  g.prev_start = ix->prev_st[i]; g.prev_end = ix->prev_en[i]; g.next_start = ix->next_st[i]; g.next_end = ix->next_en[i];
  g.rec = (int64_t)i;
}

// ---- src/g2t.cpp:70-101 IntervalTree::findOverlappingForTid
bool find_for_tid(const Ctx& c, int refid, char strand, uint32_t tid, uint32_t qs, uint32_t qe, GExon& g) {
  if (qs == 0 && qe == 0) return false;
  std::vector<uint64_t> hits;
  overlap(c, refid, strand, (int32_t)qs, (int32_t)qe, hits);
  for (uint64_t i : hits)
    if (c.ix->tid[i] == tid) { fill_gexon(c, i, g); return true; }   // tid/pos/ins/gap left as they were
  return false;
}

// ---- src/g2t.cpp:103-257 IntervalTree::findOverlapping.  Returns false when the hit passes no filter.
bool filter_hit(const brb_params& cfg, char strand, int status, uint32_t qstart, uint32_t qend, uint32_t s, uint32_t e,
                uint32_t pos_start, uint32_t& pos, uint32_t& left_gap, uint32_t& left_ins, uint32_t& right_gap, uint32_t& right_ins) {
  pos = left_gap = left_ins = right_gap = right_ins = 0;
  if (strand == '+') {
    if (s <= qstart) {
      pos = (qstart - s) + pos_start; left_gap = qstart - s;
      if (status == MIDDLE_EXON || status == LAST_EXON) { if (left_gap > cfg.max_junc_gap) return false; }
    } else {
      pos = pos_start; left_ins = s - qstart;
      if (status == MIDDLE_EXON || status == LAST_EXON) { if (left_ins > cfg.max_junc_ins) return false; }
      else { if (left_ins > cfg.max_clip) return false; }
    }
    if (e < qend) {
      right_ins = qend - e;
      if (status == FIRST_EXON || status == MIDDLE_EXON) { if (right_ins > cfg.max_junc_ins) return false; }
      else { if (right_ins > cfg.max_clip) return false; }
    } else if (qend < e) {
      right_gap = e - qend;
      if (status == FIRST_EXON || status == MIDDLE_EXON) { if (right_gap > cfg.max_junc_gap) return false; }
    }
  } else {
    if (qend <= e) {
      pos = (e - qend) + pos_start; right_gap = e - qend;
      if (status == FIRST_EXON || status == MIDDLE_EXON) { if (right_gap > cfg.max_junc_gap) return false; }
    } else {
      pos = pos_start; right_ins = qend - e;
      // src/g2t.cpp:204 `status == FIRST_EXON || MIDDLE_EXON` is always true: never max_clip here
      if (right_ins > cfg.max_junc_ins) return false;
    }
    if (qstart < s) {
      left_ins = s - qstart;
      if (status == MIDDLE_EXON || status == LAST_EXON) { if (left_ins > cfg.max_junc_ins) return false; }
      else { if (left_ins > cfg.max_clip) return false; }
    } else if (s < qstart) {
      left_gap = qstart - s;
      if (status == MIDDLE_EXON || status == LAST_EXON) { if (left_gap > cfg.max_junc_gap) return false; }
    }
  }
  return true;
}

bool find_overlapping(Ctx& c, int refid, char strand, uint32_t qs, uint32_t qe, int status, std::vector<GExon>& out) {
  std::vector<uint64_t> hits;
  overlap(c, refid, strand, (int32_t)qs, (int32_t)qe, hits);
  c.n_hits += hits.size();
  if (hits.empty()) return false;
  for (uint64_t i : hits) {
    uint32_t pos, lg, li, rg, ri;
    if (!filter_hit(*c.p, strand, status, qs, qe, (uint32_t)c.ix->st[i], (uint32_t)c.ix->en[i], c.ix->pos_start[i], pos, lg, li, rg, ri)) continue;
    GExon g; fill_gexon(c, i, g);
    g.tid = c.ix->tid[i]; g.pos = pos; g.left_gap = (int32_t)lg; g.left_ins = (int32_t)li; g.right_gap = (int32_t)rg; g.right_ins = (int32_t)ri;
    out.push_back(g);
  }
  return !out.empty();
}

// ---- ksw2: subprojects/packagefiles/ksw2/ksw2_extz2_sse.cpp:37-318 with bramble's arguments
// (src/evaluate.cpp:285-317: m=5, match 1, mismatch -4, N -> -e, q=4, e=1, w=-1 (full band), zdrop=40,
// end_bonus=0, flag = EXTZ_ONLY|APPROX_MAX|APPROX_DROP), plus ksw2.h helpers.
struct KswRes { int max = 0, score = KSW_NEG_INF, max_q = -1, max_t = -1; bool zdropped = false; std::vector<uint32_t> cigar; };

void ksw_push(std::vector<uint32_t>& cg, uint32_t op, int len) {   // ksw_push_cigar
  if (cg.empty() || op != (cg.back() & 0xf)) cg.push_back((uint32_t)len << 4 | op);
  else cg.back() += (uint32_t)len << 4;
}

KswRes ksw_extend(const uint8_t* query, int qlen, const uint8_t* target, int tlen) {
  KswRes ez;
  const int q = 4, e = 1, qe = q + e, zdrop = 40, max_sc = 1 + qe * 2;
  if (qlen <= 0 || tlen <= 0) return ez;
  std::vector<int8_t> u(tlen + 1, 0), v(tlen + 1, 0), x(tlen + 1, 0), y(tlen + 1, 0);
  int n_rows = qlen + tlen - 1, ncol = std::min(qlen, tlen);
  std::vector<uint8_t> pmat((size_t)n_rows * ncol);
  std::vector<int> off(n_rows);
  int32_t H0 = 0; int last_H0_t = 0;
  for (int r = 0; r < n_rows; ++r) {
    int st = 0, en = tlen - 1;
    if (st < r - qlen + 1) st = r - qlen + 1;
    if (en > r) en = r;
    if (st > en) { ez.zdropped = true; break; }
    int8_t x1, v1;
    if (st > 0) { x1 = x[st - 1]; v1 = v[st - 1]; }     // (r-1, st-1) was computed in the previous round
    else { x1 = 0; v1 = r ? q : 0; }
    if (en >= r) { y[r] = 0; u[r] = r ? q : 0; }
    off[r] = st;
    uint8_t* pr = &pmat[(size_t)r * ncol];
    for (int t = st; t <= en; ++t) {
      uint8_t sq = target[t], sr = query[r - t];
      int s = (sq == 4 || sr == 4) ? -e : (sq == sr ? 1 : -4);
      int z = s + qe * 2;
      int xt1 = x1; x1 = x[t];
      int vt1 = v1; v1 = v[t];
      int a = xt1 + vt1, ut = u[t], b = y[t] + ut;
      int d = a > z ? 1 : 0;
      z = z > a ? z : a;
      d = b > z ? 2 : d;
      z = z > b ? z : b;
      z = z < max_sc ? z : max_sc;
      u[t] = (int8_t)(z - vt1); v[t] = (int8_t)(z - ut);
      z -= q; a -= z; b -= z;
      x[t] = (int8_t)(a > 0 ? a : 0); if (a > 0) d |= 0x08;
      y[t] = (int8_t)(b > 0 ? b : 0); if (b > 0) d |= 0x10;
      pr[t - st] = (uint8_t)d;
    }
    if (r > 0) {   // approximate max / z-drop along a greedy diagonal cursor, lines 284-297
      if (last_H0_t >= st && last_H0_t <= en && last_H0_t + 1 >= st && last_H0_t + 1 <= en) {
        int32_t d0 = (uint8_t)v[last_H0_t] - qe, d1 = (uint8_t)u[last_H0_t + 1] - qe;
        if (d0 > d1) H0 += d0; else H0 += d1, ++last_H0_t;
      } else if (last_H0_t >= st && last_H0_t <= en) {
        H0 += (uint8_t)v[last_H0_t] - qe;
      } else {
        ++last_H0_t; H0 += (uint8_t)u[last_H0_t] - qe;
      }
      // ksw_apply_zdrop(ez, 1, H0, r, last_H0_t, zdrop, e)
      int t = last_H0_t;
      if (H0 > ez.max) { ez.max = H0; ez.max_t = t; ez.max_q = r - t; }
      else if (t >= ez.max_t && r - t >= ez.max_q) {
        int tl = t - ez.max_t, ql = (r - t) - ez.max_q, l = tl > ql ? tl - ql : ql - tl;
        if (ez.max - H0 > zdrop + l * e) { ez.zdropped = true; break; }
      }
    } else { H0 = (uint8_t)v[0] - qe - qe; last_H0_t = 0; }
    if (r == qlen + tlen - 2 && en == tlen - 1) ez.score = H0;
  }
  if (ez.max_t >= 0 && ez.max_q >= 0) {   // ksw_backtrack(is_rot=1) from (max_t, max_q)
    int i = ez.max_t, j = ez.max_q, state = 0;
    std::vector<uint32_t>& cg = ez.cigar;
    while (i >= 0 && j >= 0) {
      int r = i + j;
      uint32_t tmp = pmat[(size_t)r * ncol + i - off[r]];
      if (state == 0) state = tmp & 7;
      else if (!(tmp >> (state + 2) & 1)) state = 0;
      if (state == 0) state = tmp & 7;
      if (state == 0) { ksw_push(cg, 0, 1); --i; --j; }
      else if (state == 1 || state == 3) { ksw_push(cg, 2, 1); --i; }
      else { ksw_push(cg, 1, 1); --j; }
    }
    if (i >= 0) ksw_push(cg, 2, i + 1);
    if (j >= 0) ksw_push(cg, 1, j + 1);
    std::reverse(cg.begin(), cg.end());
  }
  return ez;
}

// ---- exon sequence of an index record: GFaSeqGet::copyRange(start, end-1, false, true) encoded like
// src/evaluate.cpp:299-305 (A,C,G,T -> 0..3, anything else -> 4); src/g2t.cpp:50-55.
void exon_codes(const Ctx& c, int refid, const GExon& g, std::vector<uint8_t>& out) {
  out.clear();
  if (!c.ix->genome_codes || refid < 0 || (uint32_t)refid >= c.ix->n_chrom) return;
  uint64_t base = c.ix->chrom_off[refid], lim = c.ix->chrom_off[refid + 1];
  for (uint64_t pos = g.start; pos < g.end; pos++) {        // 1-based [start, end)
    uint64_t o = base + pos - 1;
    out.push_back(o < lim ? (c.ix->genome_format == 1 ? (uint8_t)((c.ix->genome_codes[o >> 1] >> ((o & 1) * 4)) & 15) : c.ix->genome_codes[o]) : 4);
  }
}

struct ReadView {
  int refid; char strand; uint32_t n_exons; const uint32_t* es; const uint32_t* ee;
  uint32_t n_cigar; const uint32_t* cigar; const uint8_t* seq; int seq_len;
};

int get_exon_status(uint32_t exon_count, uint32_t j) {   // src/evaluate.cpp:44-56
  if (exon_count == 1) return ONLY_EXON;
  if (j == 0) return FIRST_EXON;
  if (j < exon_count - 1) return MIDDLE_EXON;
  return LAST_EXON;
}

typedef std::vector<TidData> TidMap;   // ankerl::unordered_dense::map: iteration == insertion order
TidData* tid_find(TidMap& m, uint32_t tid) { for (auto& t : m) if (t.tid == tid) return &t; return nullptr; }

// ---- src/evaluate.cpp:111-182
bool correct_for_gaps(Ctx& c, TidData& td, uint32_t tid, Segment& seg2, char strand, int refid) {
  const Segment* prev_guide = nullptr;
  for (int k = (int)td.segments.size() - 1; k >= 0; k--) if (td.segments[k].has_gexon) { prev_guide = &td.segments[k]; break; }
  if (!prev_guide) return true;
  uint8_t gap = (uint8_t)(seg2.gexon.exon_id - prev_guide->gexon.exon_id);
  if (!c.p->long_reads) { if (gap != 1) { td.elim = true; return false; } return true; }
  if (gap > 2) { td.elim = true; return false; }
  if (gap == 2) {
    uint32_t gs = strand == '+' ? seg2.gexon.prev_start : seg2.gexon.next_start;
    uint32_t ge = strand == '+' ? seg2.gexon.prev_end : seg2.gexon.next_end;
    if ((gs == 0 && ge == 0) || (ge - gs > c.p->max_error_exon)) { td.elim = true; return false; }
    GExon gap_exon;
    if (!find_for_tid(c, refid, strand, tid, gs, ge, gap_exon)) { td.elim = true; return false; }
    Segment gseg; gseg.gexon = gap_exon; gseg.has_gexon = true; gseg.status = GAP_EXON;
    gseg.is_small_exon = (gap_exon.end - gap_exon.start <= c.p->max_error_exon);
    td.segments.push_back(gseg);
  }
  return true;
}

// ---- src/evaluate.cpp:184-282
void get_intervals(Ctx& c, TidMap& data, const ReadView& rd, uint32_t j, char strand, bool has_left_clip, bool has_right_clip, bool& failure) {
  uint32_t qs = rd.es[j], qe = rd.ee[j];
  int status = get_exon_status(rd.n_exons, j);
  bool is_small = (qe - qs <= c.p->max_error_exon);
  bool data_empty = data.empty();
  std::vector<GExon> gexons; std::vector<uint32_t> cand;
  if (find_overlapping(c, rd.refid, strand, qs, qe, status, gexons)) {
    for (auto& g : gexons) {
      cand.push_back(g.tid);
      Segment seg; seg.has_gexon = seg.has_qexon = true; seg.gexon = g; seg.qs = qs; seg.qe = qe; seg.status = status; seg.is_small_exon = is_small;
      if (data_empty) {
        TidData td; td.tid = g.tid; td.has_left_clip = has_left_clip; td.has_right_clip = has_right_clip; td.segments.push_back(seg);
        TidData* ex = tid_find(data, g.tid);
        if (ex) *ex = td; else data.push_back(td);      // data[tid] = td
      } else {
        TidData* td = tid_find(data, g.tid);
        if (!td || td->elim) continue;
        correct_for_gaps(c, *td, g.tid, seg, strand, rd.refid);
        td->segments.push_back(seg);
      }
    }
    for (auto& td : data) if (std::find(cand.begin(), cand.end(), td.tid) == cand.end()) td.elim = true;
    return;
  }
  if (status != ONLY_EXON && c.p->ignore_small_exons && is_small) {
    if (status == MIDDLE_EXON) {
      if (data.empty()) { failure = true; return; }
      for (auto& td : data) {
        Segment ig; ig.qs = qs; ig.qe = qe; ig.has_qexon = true; ig.has_gexon = false; ig.status = INS_EXON; ig.is_small_exon = true;
        td.segments.push_back(ig);
      }
      return;
    }
    failure = true; return;
  }
  failure = true;
}

// ---- src/evaluate.cpp:69-109
void get_clips(const brb_params& p, const ReadView& rd, bool& hl, bool& hr, uint32_t& nl, uint32_t& nr) {
  const uint32_t* cg = rd.cigar; uint32_t n = rd.n_cigar;
  if (n == 0) return;
  if ((cg[0] & 0xf) == CH) { if (1 < n && (cg[1] & 0xf) == CS) { hl = p.use_fasta; nl = cg[1] >> 4; } }
  else if ((cg[0] & 0xf) == CS) { hl = p.use_fasta; nl = cg[0] >> 4; }
  if ((cg[n - 1] & 0xf) == CH) { if (n >= 2 && (cg[n - 2] & 0xf) == CS) { hr = p.use_fasta; nr = cg[n - 2] >> 4; } }
  else if ((cg[n - 1] & 0xf) == CS) { hr = p.use_fasta; nr = cg[n - 1] >> 4; }
}

void count_consumed(const std::vector<uint32_t>& cg, int& qc, int& rc) {
  qc = rc = 0;
  for (uint32_t w : cg) { int op = w & 0xf, len = w >> 4; if (op == 0 || op == 1) qc += len; if (op == 0 || op == 2) rc += len; }
}

// ---- src/evaluate.cpp:451-498 (+320-448)
void left_clip_rescue(Ctx& c, TidData& td, char strand, const ReadView& rd, uint32_t tid, uint32_t n_left_clip) {
  td.has_left_clip = false;
  Segment& seg = td.segments[0];
  if (!seg.has_gexon || seg.gexon.left_gap > 0) return;
  GExon& gexon = seg.gexon;
  uint32_t total = n_left_clip + (uint32_t)gexon.left_ins;
  if ((int)total > rd.seq_len) total = (uint32_t)rd.seq_len;
  std::vector<uint8_t> qseq(rd.seq, rd.seq + total), gseq, tmp;
  GExon curr = gexon; int i = 0, n_ex = 0;
  while (qseq.size() > gseq.size()) {                       // collect_left_exons
    i++;
    bool has_nb = strand == '+' ? curr.has_prev : curr.has_next;
    if (!has_nb) { if (i == 1) return; break; }
    GExon next;
    bool ok = strand == '+' ? find_for_tid(c, rd.refid, strand, tid, curr.prev_start, curr.prev_end, next)
                            : find_for_tid(c, rd.refid, strand, tid, curr.next_start, curr.next_end, next);
    if (!ok) return;                                         // undefined behaviour in the reference
    exon_codes(c, rd.refid, next, tmp);
    gseq.insert(gseq.begin(), tmp.begin(), tmp.end());
    n_ex++; curr = next;
  }
  if (n_ex == 0) return;
  c.n_rescue++;
  int len1 = (int)qseq.size(), len2 = (int)gseq.size();
  int start_pos = std::max(0, len2 - (len1 + 40));          // align_reversed
  std::vector<uint8_t> g2(gseq.begin() + start_pos, gseq.end()), q2(qseq);
  std::reverse(g2.begin(), g2.end()); std::reverse(q2.begin(), q2.end());
  KswRes r = ksw_extend(q2.data(), (int)q2.size(), g2.data(), (int)g2.size());
  if (r.max < 10 || r.score == KSW_NEG_INF) return;
  if (gexon.left_ins > 0) gexon.left_ins = 0;
  int qc, rc; count_consumed(r.cigar, qc, rc);               // build_left_clip_segment
  int left_clip = len1 - qc;
  Segment s; s.has_qexon = false; s.has_gexon = true; s.status = LEFTC_EXON; s.is_small_exon = (len1 <= (int)c.p->max_error_exon); s.score = r.max;
  s.gexon.start = gexon.start - rc; s.gexon.end = gexon.start; s.gexon.pos = gexon.pos_start - rc;
  if (left_clip > 0) s.cigar.add(left_clip, OV_S);
  int n = (int)r.cigar.size();
  for (int k = n - 1; k >= 0; --k) {
    int len = r.cigar[k] >> 4, op = r.cigar[k] & 0xf;       // 0 M, 1 I, 2 D
    if (k == n - 1 && op == 2) {}
    else if (k == n - 1 && op == 1) s.cigar.add(len, OV_S);
    else if (op == 2) s.cigar.add(len, OV_D);
    else if (op == 1) s.cigar.add(len, OV_I);
    else s.cigar.add(len, OV_M);
  }
  td.segments.insert(td.segments.begin(), s);
  td.has_left_clip = true;
}

// ---- src/evaluate.cpp:600-656 (+500-598)
void right_clip_rescue(Ctx& c, TidData& td, char strand, const ReadView& rd, uint32_t tid, uint32_t n_right_clip) {
  td.has_right_clip = false;
  Segment& seg = td.segments.back();
  if (!seg.has_gexon || seg.gexon.right_gap > 0) return;
  GExon& gexon = seg.gexon;
  uint32_t total = (uint32_t)gexon.right_ins + n_right_clip;
  if ((int)total > rd.seq_len) total = (uint32_t)rd.seq_len;
  int start = rd.seq_len - (int)total;
  std::vector<uint8_t> qseq(rd.seq + start, rd.seq + start + total), gseq, tmp;
  GExon curr = gexon; int i = 0, n_ex = 0;
  while (qseq.size() > gseq.size()) {                       // collect_right_exons
    i++;
    bool has_nb = strand == '+' ? curr.has_next : curr.has_prev;
    if (!has_nb) { if (i == 1) return; break; }
    GExon next;
    bool ok = strand == '+' ? find_for_tid(c, rd.refid, strand, tid, curr.next_start, curr.next_end, next)
                            : find_for_tid(c, rd.refid, strand, tid, curr.prev_start, curr.prev_end, next);
    if (!ok) return;
    exon_codes(c, rd.refid, next, tmp);
    gseq.insert(gseq.end(), tmp.begin(), tmp.end());
    n_ex++; curr = next;
  }
  if (n_ex == 0) return;
  c.n_rescue++;
  int len1 = (int)qseq.size();
  if ((int)gseq.size() > len1 + 40) gseq.resize(len1 + 40);
  KswRes r = ksw_extend(qseq.data(), len1, gseq.data(), (int)gseq.size());
  if (r.max < 10 || r.score == KSW_NEG_INF) return;
  if (gexon.right_ins > 0) gexon.right_ins = 0;
  int qc, rc; count_consumed(r.cigar, qc, rc);
  int right_clip = len1 - qc;
  Segment s; s.has_qexon = false; s.has_gexon = true; s.status = RIGHTC_EXON; s.is_small_exon = (len1 <= (int)c.p->max_error_exon); s.score = r.max;
  s.gexon.start = gexon.end; s.gexon.end = gexon.end + rc; s.gexon.pos = gexon.pos_start - rc;
  int n = (int)r.cigar.size();
  for (int k = 0; k < n; ++k) {
    int len = r.cigar[k] >> 4, op = r.cigar[k] & 0xf;
    if (k == n - 1 && op == 2) {}
    else if (k == n - 1 && op == 1) s.cigar.add(len, OV_S);
    else if (op == 2) s.cigar.add(len, OV_D);
    else if (op == 1) s.cigar.add(len, OV_I);
    else s.cigar.add(len, OV_M);
  }
  if (right_clip > 0) s.cigar.add(right_clip, OV_S);
  td.segments.push_back(s);
  td.has_right_clip = true;
}

// ---- src/evaluate.cpp:675-786
void build_cigar_match(const Segment& seg, const TidData& td, Match& m, bool first_match, bool last_match) {
  uint32_t qstart = seg.qs, qend = seg.qe, gstart = seg.gexon.start, gend = seg.gexon.end;
  uint32_t left_ins = seg.gexon.left_ins, left_gap = seg.gexon.left_gap, right_ins = seg.gexon.right_ins, right_gap = seg.gexon.right_gap;
  Cigar& cg = m.cigar; int st = seg.status;
  if (left_ins > 0) {
    if (st == FIRST_EXON || st == ONLY_EXON) {
      if (!td.has_left_clip) { cg.add(left_ins, CS); m.total_operations += left_ins; m.prev_op = CS; }
    } else if (st == MIDDLE_EXON || st == LAST_EXON || td.has_left_clip) {
      cg.add(left_ins, CI); m.total_operations += left_ins;
      if (m.prev_op == CD) m.total_coverage += left_ins;
      else if (m.prev_op == CI) m.total_operations += (m.total_operations * 0.2);
      m.prev_op = CI;
    }
  } else if (left_gap > 0) {
    if (!first_match && (st == MIDDLE_EXON || st == LAST_EXON || td.has_left_clip)) {
      cg.add(left_gap, CD); m.total_operations += left_gap; m.ref_consumed += left_gap;
      if (m.prev_op == CI) m.total_coverage += left_gap;
      else if (m.prev_op == CD) m.total_operations += (m.total_operations * 0.2);
      m.prev_op = CD;
    }
  } else m.junc_hits++;
  uint32_t os = std::max(qstart, gstart), oe = std::min(qend, gend);
  if (oe >= os) { uint32_t len = oe - os; cg.add(len, CM); m.total_operations += len; m.total_coverage += len; m.ref_consumed += len; m.prev_op = CM; }
  if (right_ins > 0) {
    if (st == LAST_EXON || st == ONLY_EXON) {
      if (!td.has_right_clip) { cg.add(right_ins, CS); m.total_operations += right_ins; m.prev_op = CS; }
    } else if (st == FIRST_EXON || st == MIDDLE_EXON || td.has_right_clip) {
      cg.add(right_ins, CI); m.total_operations += right_ins;
      if (m.prev_op == CD) m.total_coverage += right_ins;
      m.prev_op = CI;
    }
  } else if (right_gap > 0) {
    if (!last_match && (st == FIRST_EXON || st == MIDDLE_EXON || td.has_right_clip)) {
      cg.add(right_gap, CD); m.total_operations += right_gap; m.ref_consumed += right_gap;
      if (m.prev_op == CI) m.total_coverage += right_gap;
      m.prev_op = CD;
    }
  } else m.junc_hits++;
}

// ---- src/evaluate.cpp:888-1134 (+843-886 filter_by_similarity)
void evaluate_exon_chains(Ctx& c, const ReadView& rd, std::vector<Match>& out) {
  out.clear();
  bool has_left_clip = false, has_right_clip = false; uint32_t n_left_clip = 0, n_right_clip = 0;
  if (c.p->long_reads) get_clips(*c.p, rd, has_left_clip, has_right_clip, n_left_clip, n_right_clip);
  char strands[2]; int ns = 0;                                   // get_strands_to_check :58-67
  if (c.p->long_reads) { strands[0] = '+'; strands[1] = '-'; ns = 2; }
  else if (rd.strand == '+' || rd.strand == '-') { strands[0] = rd.strand; ns = 1; }
  else { strands[0] = '+'; strands[1] = '-'; ns = 2; }
  TidMap data;
  for (int si = 0; si < ns; si++) {
    char strand = strands[si];
    data.clear();
    bool failure = false;
    for (uint32_t j = 0; j < rd.n_exons; j++) {
      get_intervals(c, data, rd, j, strand, has_left_clip, has_right_clip, failure);
      if (failure) break;
    }
    if (failure) continue;
    if (c.p->long_reads && c.p->use_fasta) {
      for (auto& td : data) {
        if (td.elim) continue;
        if (td.has_left_clip) { if (n_left_clip >= 5) left_clip_rescue(c, td, strand, rd, td.tid, n_left_clip); else td.has_left_clip = false; }
        if (td.has_right_clip) { if (n_right_clip >= 5) right_clip_rescue(c, td, strand, rd, td.tid, n_right_clip); else td.has_right_clip = false; }
      }
    }
    for (auto& td : data) {
      if (td.elim) continue;
      uint32_t n_segments = (uint32_t)td.segments.size();
      bool match_created = false;
      uint32_t first_match_idx = (uint32_t)-1, last_match_idx = (uint32_t)-1;
      uint32_t prev_gs = UINT32_MAX, prev_ge = UINT32_MAX, prev_qs = UINT32_MAX, prev_qe = UINT32_MAX;
      bool qset = false, gset = false;
      for (uint32_t k = 0; k < n_segments; k++) {
        Segment& seg = td.segments[k];
        if (seg.has_gexon) {
          if (gset && seg.gexon.start == prev_gs && seg.gexon.end == prev_ge) { td.elim = true; break; }
          prev_gs = seg.gexon.start; prev_ge = seg.gexon.end; gset = true;
        }
        if (seg.has_qexon) {
          if (qset && seg.qs == prev_qs && seg.qe == prev_qe) { td.elim = true; break; }
          prev_qs = seg.qs; prev_qe = seg.qe; qset = true;
        }
        if (!match_created && seg.has_gexon) {               // create_match :658-673
          Match& m = td.match; m = Match();
          m.tid = td.tid; m.fwpos = m.rcpos = seg.gexon.pos; m.transcript_len = seg.gexon.transcript_len; m.strand = strand;
          match_created = true; first_match_idx++; last_match_idx++;
        } else if (match_created && seg.has_gexon && seg.status != INS_EXON) {
          last_match_idx++;
          if (strand == '-') td.match.rcpos = seg.gexon.pos;
        }
      }
      for (uint32_t k = 0; k < n_segments; k++) {
        Segment& seg = td.segments[k];
        if (td.elim) break;
        bool first_match = (k == first_match_idx), last_match = (k == last_match_idx);
        Match& m = td.match;
        if (seg.status <= ONLY_EXON) build_cigar_match(seg, td, m, first_match, last_match);
        else if (seg.status == INS_EXON) {                   // build_cigar_ins :788-806
          uint32_t len = seg.qe - seg.qs;
          if (k == 0 || k == n_segments - 1) { m.cigar.add(len, CS); m.prev_op = CS; } else { m.cigar.add(len, CI); m.prev_op = CI; }
          m.total_operations += len; m.total_coverage += len;
          m.junc_hits -= (k == 0 || k == n_segments - 1) ? 1 : 2;
        } else if (seg.status == GAP_EXON) {                 // build_cigar_gap :808-822
          uint32_t len = seg.gexon.end - seg.gexon.start;
          m.cigar.add(len, CD); m.prev_op = CD; m.total_operations += len; m.total_coverage += len; m.ref_consumed += len;
          m.junc_hits -= 2;
        } else {                                             // build_cigar_clip :824-841
          for (uint32_t w : seg.cigar.v) {
            uint8_t op = w & 0xf; uint32_t len = w >> 4;
            m.cigar.add(len, op);
            if (op == OV_M || op == OV_D) m.ref_consumed += len;
          }
          m.clip_score += seg.score;
        }
      }
      if (td.match.junc_hits < 0) td.match.junc_hits = 0;
      if (!td.elim) out.push_back(td.match);
    }
  }
  if (!out.empty() && c.p->filter_by_similarity) {           // filter_by_similarity :843-865
    std::vector<Match> kept;
    for (auto& m : out) {
      double sim = (m.total_operations > 0) ? (m.total_coverage / m.total_operations) : 0.0;
      if (sim > c.p->similarity_threshold) {
        double x = ((sim - c.p->similarity_threshold) / (1.0 - c.p->similarity_threshold));
        m.similarity_score = (x * x * static_cast<double>(m.junc_hits + 1));
        kept.push_back(m);
      }
    }
    out.swap(kept);
  }
}

// ---- src/bam.cpp:22-111
int merge_ops(int real_op, int ideal_op) {
  bool rMS = (real_op == CM || real_op == CS);
  if (rMS && ideal_op == OV_S) return CS;
  if (rMS && ideal_op == OV_M) return CM;
  if (rMS && ideal_op == OV_I) return CI;
  if (rMS && ideal_op == OV_D) return CD;
  if (real_op == CD && (ideal_op == CS || ideal_op == OV_S)) return '_';
  if (real_op == CD && ideal_op == OV_M) return CD;
  if (real_op == CI && ideal_op == OV_S) return CS;
  if (real_op == CI && ideal_op == OV_M) return CI;
  if (ideal_op == OV_S) return CS;
  if (ideal_op == OV_M) return CM;
  if (ideal_op == OV_I) return CI;
  if (ideal_op == OV_D) return CD;
  if (real_op == CP) return ideal_op;
  if (real_op == CH) return CH;
  if (real_op == CI && ideal_op == CS) return CS;
  if (ideal_op == CS || ideal_op == CD || ideal_op == CI) return ideal_op;
  if (real_op == CS || real_op == CD || real_op == CI) return real_op;
  if (ideal_op == CM || ideal_op == CEQ) return CM;
  if (ideal_op == CX) return CX;
  if (real_op == CM || real_op == CEQ) return CM;
  if (real_op == CX) return CX;
  return ideal_op;
}

// ---- src/bam.cpp:113-315 merge_cigars + :443-472 get_new_cigar
void merge_cigars(const uint32_t* real, uint32_t n_real, const uint32_t* ideal, uint32_t n_ideal, std::vector<uint32_t>& res) {
  res.clear();
  uint32_t front_hard = 0, front_soft = 0, ci = 0;
  if (n_real > 0 && (real[0] & 0xf) == CH) { front_hard = real[0] >> 4; ci++; }
  if (ci < n_real && (real[ci] & 0xf) == CS) front_soft = real[ci] >> 4;
  uint32_t ri = 0, ii = 0, real_pos = 0, ideal_pos = 0;
  auto add_op = [&](uint8_t op, uint32_t len) {
    if (len == 0) return;
    if (op == '_') return;
    if (!res.empty() && (res.back() & 0xf) == op) res.back() += (len << 4);
    else res.push_back(len << 4 | op);
  };
  auto rem = [](uint32_t w, uint32_t pos) { return (w >> 4) - pos; };
  uint32_t clips = front_hard;
  while (clips > 0 && ri < n_real) {
    uint32_t avail = rem(real[ri], real_pos), chunk = clips < avail ? clips : avail;
    add_op(real[ri] & 0xf, chunk);
    clips -= chunk; real_pos += chunk;
    if (real_pos >= (real[ri] >> 4)) { ri++; real_pos = 0; }
  }
  clips = front_soft;
  while (clips > 0 && ri < n_real) {
    uint8_t real_op = real[ri] & 0xf, ideal_op = ii < n_ideal ? (ideal[ii] & 0xf) : 0xff;
    uint32_t real_rem = rem(real[ri], real_pos), ideal_rem = ii < n_ideal ? rem(ideal[ii], ideal_pos) : UINT32_MAX;
    bool ov = ii < n_ideal && (ideal_op == OV_M || ideal_op == OV_D || ideal_op == OV_I || ideal_op == OV_S);
    if (ov) {
      if (ideal_op == OV_D) {
        uint32_t chunk = ideal_rem;
        add_op((uint8_t)merge_ops(real_op, ideal_op), chunk);
        ideal_pos += chunk;
        if (ideal_pos >= (ideal[ii] >> 4)) { ii++; ideal_pos = 0; }
      } else {
        uint32_t chunk = clips; if (chunk > real_rem) chunk = real_rem; if (chunk > ideal_rem) chunk = ideal_rem;
        add_op((uint8_t)merge_ops(real_op, ideal_op), chunk);
        clips -= chunk; real_pos += chunk; ideal_pos += chunk;
        if (real_pos >= (real[ri] >> 4)) { ri++; real_pos = 0; }
        if (ideal_pos >= (ideal[ii] >> 4)) { ii++; ideal_pos = 0; }
      }
    } else {
      uint32_t chunk = clips; if (chunk > real_rem) chunk = real_rem;
      add_op((uint8_t)merge_ops(real_op, ideal_op), chunk);
      clips -= chunk; real_pos += chunk;
      if (real_pos >= (real[ri] >> 4)) { ri++; real_pos = 0; }
    }
  }
  while (ri < n_real || ii < n_ideal) {
    if (ri >= n_real) { add_op(ideal[ii] & 0xf, rem(ideal[ii], ideal_pos)); ii++; ideal_pos = 0; continue; }
    if (ii >= n_ideal) { add_op(real[ri] & 0xf, rem(real[ri], real_pos)); ri++; real_pos = 0; continue; }
    uint8_t real_op = real[ri] & 0xf, ideal_op = ideal[ii] & 0xf;
    uint32_t real_rem = rem(real[ri], real_pos), ideal_rem = rem(ideal[ii], ideal_pos);
    if (real_op == CN) { ri++; real_pos = 0; }
    else if (real_op == CD && (ideal_op == CS || ideal_op == OV_S || ideal_op == CI || ideal_op == OV_I)) {
      uint32_t chunk = real_rem < ideal_rem ? real_rem : ideal_rem;
      real_pos += chunk; ideal_pos += chunk;
      if (real_pos >= (real[ri] >> 4)) { ri++; real_pos = 0; }
      if (ideal_pos >= (ideal[ii] >> 4)) { ii++; ideal_pos = 0; }
    } else if (real_op == CI) { add_op(CI, real_rem); ri++; real_pos = 0; }
    else if (ideal_op == CD || ideal_op == OV_D) { add_op(CD, ideal_rem); ii++; ideal_pos = 0; }
    else {
      uint32_t chunk = real_rem < ideal_rem ? real_rem : ideal_rem;
      add_op((uint8_t)merge_ops(real_op, ideal_op), chunk);
      real_pos += chunk; ideal_pos += chunk;
      if (real_pos >= (real[ri] >> 4)) { ri++; real_pos = 0; }
      if (ideal_pos >= (ideal[ii] >> 4)) { ii++; ideal_pos = 0; }
    }
  }
  for (size_t i = 1; i + 1 < res.size(); i++) {
    if ((res[i] & 0xf) != CI) continue;
    uint8_t prev = res[i - 1] & 0xf, next = res[i + 1] & 0xf;
    if ((prev == CS || prev == CH) && (next == CS || next == CH)) res[i] = (res[i] >> 4) << 4 | prev;
  }
  size_t n = 0;
  for (size_t i = 0; i < res.size(); i++) {
    uint8_t op = res[i] & 0xf; uint32_t len = res[i] >> 4;
    if (n > 0 && (res[n - 1] & 0xf) == op) res[n - 1] += (len << 4); else res[n++] = len << 4 | op;
  }
  res.resize(n);
}

struct OutRec { uint32_t src_aln, tid, pos; int8_t strand; uint8_t flags; double sim; int32_t clip; Cigar ideal; };

// ---- src/mates.cpp:150-261 process_mate_pair + :28-119 add_mate_info.  `mate` null == no mate / mate had no match.
void process_mate_pair(std::vector<Match>* rd, uint32_t r_aln, std::vector<Match>* mate, uint32_t m_aln, std::vector<OutRec>& out) {
  if (!rd) return;
  auto emit = [&](const Match& m, uint32_t aln, uint8_t fl) {
    OutRec o; o.src_aln = aln; o.tid = m.tid; o.pos = m.strand == '+' ? m.fwpos : m.rcpos; o.strand = m.strand; o.flags = fl;
    o.sim = m.similarity_score; o.clip = m.clip_score; o.ideal = m.cigar; out.push_back(o);
  };
  auto by_tid = [](std::vector<Match>& v, uint32_t tid) -> const Match& {   // read_alignments[tid] = align: last one wins
    const Match* f = nullptr; for (auto& m : v) if (m.tid == tid) f = &m; return *f;
  };
  std::vector<uint32_t> rt, mt;
  for (auto& m : *rd) rt.push_back(m.tid);
  std::sort(rt.begin(), rt.end());
  if (!mate) { for (uint32_t tid : rt) emit(by_tid(*rd, tid), r_aln, BRB_REC_FIRST); return; }
  for (auto& m : *mate) mt.push_back(m.tid);
  std::sort(mt.begin(), mt.end());
  std::vector<uint32_t> common, fin;
  std::set_intersection(rt.begin(), rt.end(), mt.begin(), mt.end(), std::back_inserter(common));
  int mate_case;
  if (!common.empty()) { fin = common; mate_case = 1; }
  else if (rt.size() == 1 && mt.size() == 1) { fin.push_back(rt[0]); fin.push_back(mt[0]); mate_case = 2; }
  else return;
  std::vector<Match> r_before = *rd, m_before = *mate;     // alignments captured before update_read_matches
  auto upd = [&](std::vector<Match>& v) { std::vector<Match> nv; for (auto& m : v) if (std::find(fin.begin(), fin.end(), m.tid) != fin.end()) nv.push_back(m); v.swap(nv); };
  upd(*rd); upd(*mate);
  if (mate_case == 1) {
    for (uint32_t tid : fin) {
      emit(by_tid(r_before, tid), r_aln, BRB_REC_FIRST | BRB_REC_PAIRED | BRB_REC_SAMETX);
      emit(by_tid(m_before, tid), m_aln, BRB_REC_PAIRED | BRB_REC_SAMETX);
    }
  } else {
    emit(by_tid(r_before, rt[0]), r_aln, BRB_REC_FIRST | BRB_REC_PAIRED);
    emit(by_tid(m_before, mt[0]), m_aln, BRB_REC_PAIRED);
  }
}

template <typename T> T* dup_vec(const std::vector<T>& v) {
  T* p = (T*)malloc((v.size() ? v.size() : 1) * sizeof(T));
  if (!v.empty()) memcpy(p, v.data(), v.size() * sizeof(T));
  return p;
}

void init_ctx(Ctx& c, const brb_index_desc* ix, const brb_params* p, const brb_batch_in* in) {
  c.ix = ix; c.p = p; c.in = in;
  uint64_t n = ix->n_tables ? ix->table_off[ix->n_tables] : 0;
  c.pmax.resize(n);
  for (uint32_t t = 0; t < ix->n_tables; t++) {
    int32_t m = INT32_MIN;
    for (uint64_t i = ix->table_off[t]; i < ix->table_off[t + 1]; i++) { m = std::max(m, ix->en[i]); c.pmax[i] = m; }
  }
}
}  // namespace

extern "C" {

// ---- src/core.cpp:220-427 convert_reads (group loop, evaluation, pairing); flush()/write_to_bam stay with the host.
int brb_port_project(const brb_index_desc* ix, const brb_params* p, const brb_batch_in* in, brb_port_out* out) {
  if (!ix || !p || !in || !out) return BRB_E_ARG;
  Ctx c; init_ctx(c, ix, p, in);
  const auto t_start = std::chrono::steady_clock::now();
  std::vector<OutRec> recs; std::vector<uint32_t> group_rec_off(in->n_groups + 1, 0);
  std::vector<uint8_t> seen(in->n_aln, 0);
  for (uint32_t g = 0; g < in->n_groups; g++) {
    group_rec_off[g] = (uint32_t)recs.size();
    const uint8_t* seq = nullptr; int seq_len = 0;
    if (p->use_fasta && in->group_seq_off) { seq = in->seq_codes + in->group_seq_off[g]; seq_len = (int)(in->group_seq_off[g + 1] - in->group_seq_off[g]); }
    auto view = [&](uint32_t a) {
      ReadView rd; rd.refid = in->refid[a]; rd.strand = (char)in->strand[a];
      rd.n_exons = in->exon_off[a + 1] - in->exon_off[a]; rd.es = in->exon_st + in->exon_off[a]; rd.ee = in->exon_en + in->exon_off[a];
      rd.n_cigar = in->cigar_off[a + 1] - in->cigar_off[a]; rd.cigar = in->cigar + in->cigar_off[a]; rd.seq = seq; rd.seq_len = seq_len;
      return rd;
    };
    for (uint32_t i = in->group_off[g]; i < in->group_off[g + 1]; i++) {
      if (seen[i]) continue;
      std::vector<Match> mi; evaluate_exon_chains(c, view(i), mi); c.n_matches += mi.size();
      std::vector<Match>* this_read = mi.empty() ? nullptr : &mi;
      int32_t mate = in->mate_idx[i];
      if (mate < 0) { process_mate_pair(this_read, i, nullptr, 0, recs); seen[i] = 1; continue; }
      if (!((uint32_t)mate >= in->n_aln) && !seen[mate]) {
        std::vector<Match> mm; evaluate_exon_chains(c, view((uint32_t)mate), mm); c.n_matches += mm.size();
        process_mate_pair(this_read, i, mm.empty() ? nullptr : &mm, (uint32_t)mate, recs);
        seen[mate] = 1;
      }
      seen[i] = 1;
    }
  }
  group_rec_off[in->n_groups] = (uint32_t)recs.size();
  size_t n = recs.size();
  std::vector<uint32_t> src(n), tid(n), pos(n), coff(n + 1), clen(n), pool; std::vector<int8_t> strand(n); std::vector<uint8_t> flags(n);
  std::vector<double> sim(n); std::vector<int32_t> clip(n); std::vector<uint32_t> merged;
  for (size_t r = 0; r < n; r++) {
    const OutRec& o = recs[r]; uint32_t a = o.src_aln;
    merge_cigars(in->cigar + in->cigar_off[a], in->cigar_off[a + 1] - in->cigar_off[a], o.ideal.v.data(), (uint32_t)o.ideal.v.size(), merged);
    if (o.strand == '-') std::reverse(merged.begin(), merged.end());     // reverse_complement_bam, src/bam.cpp:636-702
    src[r] = a; tid[r] = o.tid; pos[r] = o.pos; strand[r] = o.strand; flags[r] = o.flags; sim[r] = o.sim; clip[r] = o.clip;
    coff[r] = (uint32_t)pool.size(); clen[r] = (uint32_t)merged.size(); pool.insert(pool.end(), merged.begin(), merged.end());
  }
  coff[n] = (uint32_t)pool.size();
  out->n_rec = (uint32_t)n; out->n_cigar_words = pool.size();
  out->group_rec_off = dup_vec(group_rec_off); out->src_aln = dup_vec(src); out->tid = dup_vec(tid); out->pos = dup_vec(pos);
  out->tx_strand = dup_vec(strand); out->flags = dup_vec(flags); out->sim_score = dup_vec(sim); out->clip_score = dup_vec(clip);
  out->cigar_off = dup_vec(coff); out->cigar_len = dup_vec(clen); out->cigar = dup_vec(pool);
  out->n_matches = c.n_matches; out->n_rescue = c.n_rescue; out->n_hits = c.n_hits;
  out->seconds_project = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
  return BRB_OK;
}

void brb_port_free(brb_port_out* o) {
  if (!o) return;
  free(o->group_rec_off); free(o->src_aln); free(o->tid); free(o->pos); free(o->tx_strand); free(o->flags);
  free(o->sim_score); free(o->clip_score); free(o->cigar_off); free(o->cigar_len); free(o->cigar);
  memset(o, 0, sizeof *o);
}

int brb_port_lookup(const brb_index_desc* ix, const brb_params* p, uint32_t nq, const int32_t* refid, const int8_t* strand,
                    const uint32_t* qs, const uint32_t* qe, const uint8_t* status, uint32_t cap, uint32_t* hits, uint32_t* n_hits) {
  Ctx c; init_ctx(c, ix, p, nullptr);
  std::vector<uint64_t> ov;
  for (uint32_t q = 0; q < nq; q++) {
    overlap(c, refid[q], (char)strand[q], (int32_t)qs[q], (int32_t)qe[q], ov);
    uint32_t n = 0;
    for (uint64_t i : ov) {
      uint32_t pos, lg, li, rg, ri;
      if (!filter_hit(*p, (char)strand[q], status[q], qs[q], qe[q], (uint32_t)ix->st[i], (uint32_t)ix->en[i], ix->pos_start[i], pos, lg, li, rg, ri)) continue;
      if (n < cap) hits[(size_t)q * cap + n] = (uint32_t)i;
      n++;
    }
    n_hits[q] = n;
  }
  return BRB_OK;
}

int brb_port_merge_cigars(uint32_t n, const uint32_t* real_off, const uint32_t* real, const uint32_t* ideal_off, const uint32_t* ideal,
                          uint32_t* out_off, uint32_t* out_len, uint32_t* out) {
  std::vector<uint32_t> m;
  for (uint32_t i = 0; i < n; i++) {
    merge_cigars(real + real_off[i], real_off[i + 1] - real_off[i], ideal + ideal_off[i], ideal_off[i + 1] - ideal_off[i], m);
    out_len[i] = (uint32_t)m.size();
    memcpy(out + out_off[i], m.data(), m.size() * 4);
  }
  return BRB_OK;
}

// ---- src/core.cpp:237-332 flush(): per name, in emit order; :46-58 get_mapq; :214-218 get_rand (the real library)
int brb_port_flush_tags(uint32_t n_groups, const uint32_t* group_rec_off, const uint8_t* flags, const double* sim,
                        const uint64_t* name_off, const uint8_t* names, int long_reads,
                        int32_t* hi, int32_t* nh, uint8_t* mapq, uint8_t* primary) {
  for (uint32_t g = 0; g < n_groups; g++) {
    const uint32_t r0 = group_rec_off[g], r1 = group_rec_off[g + 1];
    if (r0 == r1) continue;
    struct Unit { uint32_t row; bool paired; };
    std::vector<Unit> pairs;
    for (uint32_t r = r0; r < r1;) { const bool pr = flags[r] & BRB_REC_PAIRED; pairs.push_back({r, pr}); r += pr ? 2 : 1; }
    auto score_of = [&](const Unit& u) { double s = sim ? sim[u.row] : 0.0; if (u.paired && sim) s = std::max(s, sim[u.row + 1]); return s; };
    double best_score = -std::numeric_limits<double>::infinity(); int count_at_best = 0; size_t best = pairs.size(); int32_t hit_index = 1; int total = 0;
    for (size_t i = 0; i < pairs.size(); i++) {
      hi[pairs[i].row] = hit_index++; total++;
      if (pairs[i].paired) { hi[pairs[i].row + 1] = hit_index++; total++; }
      const double ps = score_of(pairs[i]);
      if (ps > best_score) { best_score = ps; best = i; count_at_best = 1; } else if (ps == best_score) count_at_best++;
    }
    if (count_at_best > 1) {
      std::vector<size_t> tied;
      for (size_t i = 0; i < pairs.size(); i++) if (score_of(pairs[i]) == best_score) tied.push_back(i);
      const std::string name((const char*)names + name_off[g], (size_t)(name_off[g + 1] - name_off[g]));
      std::mt19937_64 gen(std::hash<std::string>{}(name));
      std::uniform_int_distribution<uint32_t> dis(0, (uint32_t)tied.size() - 1);
      best = tied[dis(gen)];
    }
    const uint32_t mq = !long_reads ? (total == 1 ? 255 : total == 2 ? 3 : (total == 3 || total == 4) ? 1 : 0) : (total > 1 ? 0 : 3);
    for (size_t i = 0; i < pairs.size(); i++)
      for (uint32_t e = 0; e < (pairs[i].paired ? 2u : 1u); e++) { const uint32_t r = pairs[i].row + e; nh[r] = total; mapq[r] = (uint8_t)mq; primary[r] = i == best; }
  }
  return BRB_OK;
}

int brb_port_ksw_extend(uint32_t n, const uint64_t* q_off, const uint8_t* q, const uint64_t* t_off, const uint8_t* t,
                        int32_t* max, int32_t* score, int32_t* max_q, int32_t* max_t, uint8_t* zdropped,
                        uint32_t cigar_cap, uint32_t* n_cigar, uint32_t* cigar_out) {
  for (uint32_t i = 0; i < n; i++) {
    KswRes r = ksw_extend(q + q_off[i], (int)(q_off[i + 1] - q_off[i]), t + t_off[i], (int)(t_off[i + 1] - t_off[i]));
    max[i] = r.max; score[i] = r.score; max_q[i] = r.max_q; max_t[i] = r.max_t; zdropped[i] = r.zdropped;
    n_cigar[i] = (uint32_t)r.cigar.size();
    for (uint32_t k = 0; k < r.cigar.size() && k < cigar_cap; k++) cigar_out[(size_t)i * cigar_cap + k] = r.cigar[k];
  }
  return BRB_OK;
}
}  // extern "C"
