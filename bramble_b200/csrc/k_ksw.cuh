// Kernel group 5: long-read soft-clip rescue.
//   * k_rescue_plan / k_rescue_gather : left_clip_rescue / right_clip_rescue drivers of the reference
//     (src/evaluate.cpp:320-365, 451-546, 600-656): which candidates get a rescue, the query slice of the
//     group-shared read sequence and the neighbour-exon target sequence.
//   * k_ksw : ksw_extz2_sse (subprojects/packagefiles/ksw2/ksw2_extz2_sse.cpp:37-318) with bramble's fixed
//     arguments (src/evaluate.cpp:285-317: match 1, mismatch -4, N -1, q=4, e=1, full band, zdrop 40,
//     EXTZ_ONLY|APPROX_MAX|APPROX_DROP) + ksw_backtrack (ksw2.h).  One warp per problem, lanes over the
//     cells of an anti-diagonal (the SSE code does 16 int8 lanes; here 32 threads), int8 difference state
//     u,v,x,y in shared memory, one direction byte per cell to a per-warp scratch, greedy H0 cursor and
//     z-drop evaluated warp-uniformly after every diagonal, traceback by lane 0.
//   * k_rescue_finish : build_left/right_clip_segment (:397-448, 548-598).
#pragma once
#include "dev_common.cuh"
#include "k_chain.cuh"

namespace brb {

struct KswOut { int32_t max, score, max_q, max_t; uint32_t zdropped, n_cig; };

constexpr int KSW_SH_T = 768;          // cells of difference state kept in shared memory per warp
constexpr int KSW_WARPS = 4;           // warps per CTA

// DP + traceback for one problem by one warp.  `tb` receives the run-length ops in TRACEBACK order
// (last alignment op first; op 0=M 1=I 2=D), i.e. ksw2's cigar before its final reversal.
__device__ void ksw_extend_warp(const uint8_t* __restrict__ query, int qlen, const uint8_t* __restrict__ target, int tlen,
                                int8_t* su, int8_t* sv, int8_t* sx, int8_t* sy, uint8_t* __restrict__ pmat,
                                uint32_t* __restrict__ tb, KswOut& out) {
  const int lane = threadIdx.x & 31;
  const int q = 4, e = 1, qe = q + e, zdrop = 40, max_sc = 1 + qe * 2;
  out.max = 0; out.score = BRB_KSW_NEG_INF; out.max_q = -1; out.max_t = -1; out.zdropped = 0; out.n_cig = 0;
  if (qlen <= 0 || tlen <= 0) return;
  for (int t = lane; t <= tlen; t += 32) { su[t] = 0; sv[t] = 0; sx[t] = 0; sy[t] = 0; }
  __syncwarp();
  const int n_rows = qlen + tlen - 1, ncol = qlen < tlen ? qlen : tlen;
  int H0 = 0, last_H0_t = 0, emax = 0, emax_t = -1, emax_q = -1;
  bool zdropped = false;
  for (int r = 0; r < n_rows; ++r) {
    int st = r - qlen + 1; if (st < 0) st = 0;
    int en = r < tlen - 1 ? r : tlen - 1;
    if (en >= r && lane == 0) { sy[r] = 0; su[r] = r ? q : 0; }     // new cell (t == r) boundary, lines 121-137
    __syncwarp();
    uint8_t* pr = pmat + (size_t)r * ncol;
    const int ncell = en - st + 1;
    // chunks of 32 cells from HIGH t to low t: a cell reads the previous diagonal's x,v at t-1, which a lower chunk overwrites later
    for (int c0 = ((ncell - 1) >> 5) << 5; c0 >= 0; c0 -= 32) {
      const int t = st + c0 + lane;
      const bool act = t <= en;
      int xt1 = 0, vt1 = 0, ut = 0, yt = 0;
      if (act) {
        if (t > 0) { xt1 = sx[t - 1]; vt1 = sv[t - 1]; } else { xt1 = 0; vt1 = r ? q : 0; }
        ut = su[t]; yt = sy[t];
      }
      __syncwarp();
      if (act) {
        const int sq = target[t], sr = query[r - t];
        const int s = (sq == 4 || sr == 4) ? -e : (sq == sr ? 1 : -4);
        int z = s + qe * 2;
        int a = xt1 + vt1, b = yt + ut;
        int d = a > z ? 1 : 0;
        z = z > a ? z : a;
        d = b > z ? 2 : d;
        z = z > b ? z : b;
        z = z < max_sc ? z : max_sc;
        su[t] = (int8_t)(z - vt1); sv[t] = (int8_t)(z - ut);
        z -= q; a -= z; b -= z;
        sx[t] = (int8_t)(a > 0 ? a : 0); if (a > 0) d |= 0x08;
        sy[t] = (int8_t)(b > 0 ? b : 0); if (b > 0) d |= 0x10;
        pr[t - st] = (uint8_t)d;
      }
      __syncwarp();
    }
    // approximate max / z-drop along the greedy diagonal cursor (lines 284-300); warp-uniform scalars
    if (r > 0) {
      const bool in0 = last_H0_t >= st && last_H0_t <= en, in1 = last_H0_t + 1 >= st && last_H0_t + 1 <= en;
      if (in0 && in1) {
        const int d0 = (int)(uint8_t)sv[last_H0_t] - qe, d1 = (int)(uint8_t)su[last_H0_t + 1] - qe;
        if (d0 > d1) H0 += d0; else { H0 += d1; ++last_H0_t; }
      } else if (in0) {
        H0 += (int)(uint8_t)sv[last_H0_t] - qe;
      } else {
        ++last_H0_t; H0 += (int)(uint8_t)su[last_H0_t] - qe;
      }
      const int t = last_H0_t;                                       // ksw_apply_zdrop
      if (H0 > emax) { emax = H0; emax_t = t; emax_q = r - t; }
      else if (t >= emax_t && r - t >= emax_q) {
        const int tl = t - emax_t, ql = (r - t) - emax_q, l = tl > ql ? tl - ql : ql - tl;
        if (emax - H0 > zdrop + l * e) { zdropped = true; break; }
      }
    } else { H0 = (int)(uint8_t)sv[0] - qe - qe; last_H0_t = 0; }
    if (r == qlen + tlen - 2 && en == tlen - 1) out.score = H0;
  }
  out.max = emax; out.max_t = emax_t; out.max_q = emax_q; out.zdropped = zdropped ? 1u : 0u;
  __syncwarp();
  __threadfence_block();
  uint32_t n = 0;
  if (lane == 0 && emax_t >= 0 && emax_q >= 0) {                     // ksw_backtrack(is_rot=1) from (max_t, max_q)
    int i = emax_t, j = emax_q, state = 0;
    auto push = [&](uint32_t op, int len) {
      if (n == 0 || op != (tb[n - 1] & 0xfu)) tb[n++] = ((uint32_t)len << 4) | op; else tb[n - 1] += (uint32_t)len << 4;
    };
    while (i >= 0 && j >= 0) {
      const int r = i + j;
      int off = r - qlen + 1; if (off < 0) off = 0;
      const uint32_t tmp = pmat[(size_t)r * ncol + i - off];
      if (state == 0) state = tmp & 7;
      else if (!((tmp >> (state + 2)) & 1)) state = 0;
      if (state == 0) state = tmp & 7;
      if (state == 0) { push(0, 1); --i; --j; }
      else if (state == 1 || state == 3) { push(2, 1); --i; }
      else { push(1, 1); --j; }
    }
    if (i >= 0) push(2, i + 1);
    if (j >= 0) push(1, j + 1);
  }
  out.n_cig = __shfl_sync(0xffffffffu, n, 0);
}

struct KswProblems {               // flat problem arrays (n problems)
  uint32_t n;
  const uint32_t* q_off; const uint32_t* t_off; const uint32_t* tb_off;   // n+1 each
  const uint8_t* q; const uint8_t* t;
  uint32_t* tb; KswOut* out;
};

// Persistent warps: warp w takes problems w, w+W, ...; pmat/state scratch is per warp.
__global__ void __launch_bounds__(32 * KSW_WARPS) k_ksw(KswProblems P, uint8_t* __restrict__ pmat_all, size_t pmat_stride,
                                                        int8_t* __restrict__ state_all, uint32_t state_stride) {
  __shared__ int8_t sh[KSW_WARPS][4][KSW_SH_T];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t w = blockIdx.x * KSW_WARPS + wib, W = gridDim.x * KSW_WARPS;
  uint8_t* pmat = pmat_all + (size_t)w * pmat_stride;
  for (uint32_t p = w; p < P.n; p += W) {
    const int qlen = (int)(P.q_off[p + 1] - P.q_off[p]), tlen = (int)(P.t_off[p + 1] - P.t_off[p]);
    int8_t *su, *sv, *sx, *sy;
    if (tlen + 1 <= KSW_SH_T) { su = sh[wib][0]; sv = sh[wib][1]; sx = sh[wib][2]; sy = sh[wib][3]; }
    else { int8_t* g = state_all + (size_t)w * state_stride * 4; su = g; sv = g + state_stride; sx = g + 2 * state_stride; sy = g + 3 * state_stride; }
    KswOut o;
    ksw_extend_warp(P.q + P.q_off[p], qlen, P.t + P.t_off[p], tlen, su, sv, sx, sy, pmat, P.tb + P.tb_off[p], o);
    if (lane == 0) P.out[p] = o;
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------- rescue driver kernels
__device__ __forceinline__ void get_clips(const uint32_t* cg, uint32_t n, bool& hl, bool& hr, uint32_t& nl, uint32_t& nr) {   // :69-109
  hl = hr = false; nl = nr = 0;
  if (n == 0) return;
  if ((cg[0] & 0xfu) == CH) { if (n > 1 && (cg[1] & 0xfu) == CS) { hl = true; nl = cg[1] >> 4; } }
  else if ((cg[0] & 0xfu) == CS) { hl = true; nl = cg[0] >> 4; }
  if ((cg[n - 1] & 0xfu) == CH) { if (n >= 2 && (cg[n - 2] & 0xfu) == CS) { hr = true; nr = cg[n - 2] >> 4; } }
  else if ((cg[n - 1] & 0xfu) == CS) { hr = true; nr = cg[n - 1] >> 4; }
}

// Walks the neighbour exons of `rec` away from the read (collect_left_exons :333-365 / collect_right_exons
// :513-546).  WRITE=false: returns the total collected length (0 = no rescue).  WRITE=true: emits `tlen`
// target codes into `dst` (left: already reversed, i.e. nearest base first going leftwards).
template <bool WRITE>
__device__ uint32_t collect_exons(const DevIndex& ix, int32_t refid, bool left, uint32_t rec, uint32_t qlen,
                                  uint8_t* dst, uint32_t tlen) {
  uint32_t gathered = 0, n_ex = 0, i = 0, cur = rec, w = 0;
  const uint64_t base = ix.chrom_off[refid], lim = ix.chrom_off[refid + 1];
  while (qlen > gathered) {
    i++;
    // transcript-order neighbour away from the read = genomic neighbour on that side; getGuideExonForTid on the
    // neighbour's own coordinates returns that very record (exons of a tid are disjoint)
    const uint32_t nx = left ? __ldg(&ix.gprev[cur]) : rec_link(ix, cur).gnext;
    if (nx == NONE) { if (i == 1) return 0; break; }
    const RecKey k = rec_key(ix, nx);
    const uint32_t len = (uint32_t)(k.en - k.st);
    if (WRITE) {
      if (left) { for (uint32_t pos = (uint32_t)k.en; pos > (uint32_t)k.st && w < tlen; pos--) { const uint64_t o = base + (pos - 1) - 1; dst[w++] = o < lim ? ix.genome[o] : 4; } }
      else { for (uint32_t pos = (uint32_t)k.st; pos < (uint32_t)k.en && w < tlen; pos++) { const uint64_t o = base + pos - 1; dst[w++] = o < lim ? ix.genome[o] : 4; } }
    }
    gathered += len; n_ex++; cur = nx;
  }
  if (n_ex == 0) return 0;
  return gathered;
}

struct RescuePlan {                 // per (candidate, side) slot s = 2c + side (0 left, 1 right)
  uint32_t* qlen; uint32_t* tlen; uint32_t* tbcap;
};

__device__ __forceinline__ bool rescue_inputs(const DevIndex& ix, const DevBatch& b, const brb_params& p, uint32_t a, int s, uint32_t first_rec,
                                              uint32_t last_rec, int side, uint32_t& rec, uint32_t& total) {
  const uint32_t co = b.cigar_off[a], nc = b.cigar_off[a + 1] - co;
  bool hl, hr; uint32_t nl, nr; get_clips(b.cigar + co, nc, hl, hr, nl, nr);
  const uint32_t g = b.aln_group[a];
  const uint32_t seq_len = (uint32_t)(b.group_seq_off[g + 1] - b.group_seq_off[g]);
  const uint32_t e0 = b.exon_off[a], n_ex = b.exon_off[a + 1] - e0;
  if (side == 0) {
    if (!hl || nl < 5) return false;
    const RecKey k = rec_key(ix, first_rec); Hit h;
    filter_hit(p, s, exon_status(n_ex, 0), b.exon_st[e0], b.exon_en[e0], (uint32_t)k.st, (uint32_t)k.en, k.pos_start, h);
    if (h.left_gap > 0) return false;
    total = nl + h.left_ins; rec = first_rec;
  } else {
    if (!hr || nr < 5) return false;
    const RecKey k = rec_key(ix, last_rec); Hit h;
    filter_hit(p, s, exon_status(n_ex, n_ex - 1), b.exon_st[e0 + n_ex - 1], b.exon_en[e0 + n_ex - 1], (uint32_t)k.st, (uint32_t)k.en, k.pos_start, h);
    if (h.right_gap > 0) return false;
    total = h.right_ins + nr; rec = last_rec;
  }
  if ((int32_t)total > (int32_t)seq_len) total = seq_len;
  return true;
}

__global__ void __launch_bounds__(128) k_rescue_plan(DevIndex ix, DevBatch b, brb_params p, uint32_t n_cand, const uint32_t* __restrict__ c_task,
                                                     const uint32_t* __restrict__ c_first, const uint32_t* __restrict__ c_last,
                                                     const uint32_t* __restrict__ c_cap, RescuePlan R,
                                                     Ctl* __restrict__ ctl) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= 2 * n_cand) return;
  const uint32_t c = slot >> 1; const int side = slot & 1;
  uint32_t qlen = 0, tlen = 0;
  if (c_cap[c]) {
    const uint32_t task = c_task[c], a = task >> 1; const int s = task & 1;
    uint32_t rec, total;
    if (rescue_inputs(ix, b, p, a, s, c_first[c], c_last[c], side, rec, total)) {
      const uint32_t got = collect_exons<false>(ix, b.refid[a], side == 0, rec, total, nullptr, 0);
      if (got) { qlen = total; tlen = got < total + 40 ? got : total + 40; }   // align_reversed keeps the last qlen+40; right truncates to qlen+40
    }
  }
  R.qlen[slot] = qlen; R.tlen[slot] = tlen; R.tbcap[slot] = qlen ? qlen + tlen + 2 : 0;
  if (qlen) {       // scratch sizing for k_ksw: largest direction matrix and longest target
    const unsigned long long mat = (unsigned long long)(qlen + tlen - 1) * (qlen < tlen ? qlen : tlen);
    atomicMax(&ctl->ksw_max_mat, mat); atomicMax(&ctl->ksw_max_tlen, (unsigned long long)tlen);
  }
}

__global__ void __launch_bounds__(128) k_rescue_gather(DevIndex ix, DevBatch b, brb_params p, uint32_t n_cand, const uint32_t* __restrict__ c_task,
                                                       const uint32_t* __restrict__ c_first, const uint32_t* __restrict__ c_last,
                                                       const uint32_t* __restrict__ q_off, const uint32_t* __restrict__ t_off,
                                                       uint8_t* __restrict__ qbuf, uint8_t* __restrict__ tbuf) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= 2 * n_cand) return;
  const uint32_t qlen = q_off[slot + 1] - q_off[slot], tlen = t_off[slot + 1] - t_off[slot];
  if (!qlen) return;
  const uint32_t c = slot >> 1; const int side = slot & 1;
  const uint32_t task = c_task[c], a = task >> 1; const int s = task & 1;
  const uint32_t g = b.aln_group[a];
  const uint8_t* seq = b.seq + b.group_seq_off[g];
  const uint32_t seq_len = (uint32_t)(b.group_seq_off[g + 1] - b.group_seq_off[g]);
  uint8_t* qd = qbuf + q_off[slot];
  if (side == 0) for (uint32_t i = 0; i < qlen; i++) qd[i] = seq[qlen - 1 - i];           // get_left_rescue_query + reversal (:320-330, 368-395)
  else for (uint32_t i = 0; i < qlen; i++) qd[i] = seq[seq_len - qlen + i];               // get_right_rescue_query (:500-511)
  const uint32_t rec = side == 0 ? c_first[c] : c_last[c];
  collect_exons<true>(ix, b.refid[a], side == 0, rec, qlen, tbuf + t_off[slot], tlen);
}

// build_left_clip_segment / build_right_clip_segment: KswOut + traceback-order ops -> RescueSeg + override-coded CIGAR.
__global__ void __launch_bounds__(128) k_rescue_finish(uint32_t n_cand, const uint32_t* __restrict__ q_off, const uint32_t* __restrict__ tb_off,
                                                       const KswOut* __restrict__ ko, const uint32_t* __restrict__ tb, uint32_t* __restrict__ rpool,
                                                       RescueSeg* __restrict__ r_left, RescueSeg* __restrict__ r_right, uint32_t* __restrict__ c_cap,
                                                       Ctl* __restrict__ ctl) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= 2 * n_cand) return;
  const uint32_t c = slot >> 1; const int side = slot & 1;
  RescueSeg sgm; sgm.ok = 0; sgm.score = 0; sgm.rc = 0; sgm.cig_off = 0; sgm.cig_len = 0;
  const uint32_t qlen = q_off[slot + 1] - q_off[slot];
  if (qlen) {
    const KswOut o = ko[slot];
    if (o.max >= 10 && o.score != BRB_KSW_NEG_INF) {
      const uint32_t* T = tb + tb_off[slot]; const uint32_t n = o.n_cig;     // T[0] is the LAST op of the forward ksw cigar
      int32_t qc = 0, rc = 0;
      for (uint32_t k = 0; k < n; k++) { const uint32_t op = T[k] & 0xfu, len = T[k] >> 4; if (op == 0 || op == 1) qc += len; if (op == 0 || op == 2) rc += len; }
      uint32_t* out = rpool + tb_off[slot]; uint32_t m = 0;
      auto add = [&](uint32_t len, uint32_t op) { if (m > 0 && (out[m - 1] & 0xfu) == op) out[m - 1] = (((out[m - 1] >> 4) + len) << 4) | op; else out[m++] = (len << 4) | op; };
      auto conv = [&](uint32_t w, bool last) {
        const uint32_t op = w & 0xfu, len = w >> 4;
        if (last && op == 2) return;
        if (last && op == 1) { add(len, OV_S); return; }
        add(len, op == 2 ? OV_D : (op == 1 ? OV_I : OV_M));
      };
      const int32_t clip = (int32_t)qlen - qc;
      if (side == 0) {
        if (clip > 0) add((uint32_t)clip, OV_S);
        for (uint32_t k = 0; k < n; k++) conv(T[k], k == 0);                // forward index n-1 .. 0
      } else {
        for (uint32_t k = n; k-- > 0;) conv(T[k], k == 0);                  // forward index 0 .. n-1
        if (clip > 0) add((uint32_t)clip, OV_S);
      }
      sgm.ok = 1; sgm.score = o.max; sgm.rc = rc; sgm.cig_off = tb_off[slot]; sgm.cig_len = m;
      atomicAdd(&c_cap[c], m);
    }
    atomicAdd(&ctl->n_rescue, 1ull);
  }
  (side == 0 ? r_left : r_right)[c] = sgm;
}

__global__ void k_aln_group(DevBatch b, uint32_t* __restrict__ aln_group) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= b.n_groups) return;
  for (uint32_t a = b.group_off[g]; a < b.group_off[g + 1]; a++) aln_group[a] = g;
}

}  // namespace brb
