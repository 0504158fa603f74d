// Kernel group 5: long-read soft-clip rescue.
//   * k_rescue_plan : left_clip_rescue / right_clip_rescue drivers of the reference (src/evaluate.cpp:320-365, 451-546,
//     600-656): which (candidate, side) slots get a rescue, the query slice of the group-shared read sequence and the length
//     of the neighbour-exon target; every problem is appended to the queue of its size class (device counters, no host
//     round trip) and gets its share of the traceback pool.
//   * k_ksw_group<G, W> : ksw_extz2_sse (subprojects/packagefiles/ksw2/ksw2_extz2_sse.cpp:37-318) with bramble's fixed
//     arguments (src/evaluate.cpp:285-317: match 1, mismatch -4, N -1, q=4, e=1, full band, zdrop 40,
//     EXTZ_ONLY|APPROX_MAX|APPROX_DROP) + ksw_backtrack (ksw2.h).  A GROUP of G lanes (8, 16 or 32: a warp works on 4, 2
//     or 1 problems at once -- the typical clip is 5-40 bases) owns one problem.  The difference state u, v, x, y lives in
//     REGISTERS, indexed by QUERY position j (cell (r, t) of anti-diagonal r sits at j = r - t): 2W cells per lane as W
//     words of two 16-bit lanes, updated with the DPX instructions of sm_90+/sm_100 (VIADD.16x2, VIMNMX.S16x2/U16x2 -- one
//     instruction per two cells; the byte-wide video intrinsics are emulated on this architecture, 16x2 is native).  With
//     that indexing x and v of the previous diagonal are already in place, y, u and the target base arrive from cell j-1:
//     a 16-bit funnel shift inside the lane plus one shuffle per array across lanes.  No shared-memory state, no
//     __syncwarp inside a diagonal.  One direction byte per cell goes to a per-group scratch (fixed size per class); the
//     greedy H0 cursor and z-drop are evaluated per group after every diagonal; traceback by the group's first lane.
//     tools/ksw_proto.cpp is this formulation as plain loops, checked against the oracle on 20k random problems.
//   * k_ksw_big : problems beyond 512 query bases: one warp each, int8 state in shared / global memory (ksw_extend_warp).
//   * k_rescue_finish : build_left/right_clip_segment (:397-448, 548-598).
#pragma once
#include "dev_common.cuh"
#include "k_chain.cuh"

namespace brb {

struct KswOut { int32_t max, score, max_q, max_t; uint32_t zdropped, n_cig; };

constexpr int KSW_SH_T = 768;          // k_ksw_big: cells of difference state kept in shared memory per warp
constexpr int KSW_WARPS = 4;           // warps per CTA
constexpr int KSW_CLASSES = 6;         // query length <= 32, 64, 128, 256, 512, beyond
constexpr int KSW_BIG = KSW_CLASSES - 1;
constexpr int KSW_TAIL = 40;           // the target is at most the query length + 40 (src/evaluate.cpp:368-395, 600-656)
__host__ __device__ __forceinline__ int ksw_class_of(uint32_t qlen) { return qlen <= 32 ? 0 : qlen <= 64 ? 1 : qlen <= 128 ? 2 : qlen <= 256 ? 3 : qlen <= 512 ? 4 : 5; }

// DP + traceback for one problem by one warp (big problems).  `tb` receives the run-length ops in TRACEBACK order
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

// ---------------------------------------------------------------------------------------------- rescue inputs
__device__ __forceinline__ void get_clips(const uint32_t* cg, uint32_t n, bool& hl, bool& hr, uint32_t& nl, uint32_t& nr) {   // :69-109
  hl = hr = false; nl = nr = 0;
  if (n == 0) return;
  if ((cg[0] & 0xfu) == CH) { if (n > 1 && (cg[1] & 0xfu) == CS) { hl = true; nl = cg[1] >> 4; } }
  else if ((cg[0] & 0xfu) == CS) { hl = true; nl = cg[0] >> 4; }
  if ((cg[n - 1] & 0xfu) == CH) { if (n >= 2 && (cg[n - 2] & 0xfu) == CS) { hr = true; nr = cg[n - 2] >> 4; } }
  else if ((cg[n - 1] & 0xfu) == CS) { hr = true; nr = cg[n - 1] >> 4; }
}

// Total length of the neighbour exons of `rec` walking away from the read until the query is covered (collect_left_exons
// :333-365 / collect_right_exons :513-546); 0 = no rescue (no neighbour at the first step).
__device__ uint32_t collect_len(const DevIndex& ix, bool left, uint32_t rec, uint32_t qlen, unsigned long long* key = nullptr) {
  uint32_t gathered = 0, i = 0, cur = rec;
  unsigned long long h = 0x9E3779B97F4A7C15ull;          // (key: a mix of the coordinates of the exons walked -- they ARE the target)
  while (qlen > gathered) {
    i++;
    // transcript-order neighbour away from the read = genomic neighbour on that side; getGuideExonForTid on the
    // neighbour's own coordinates returns that very record (exons of a tid are disjoint)
    const uint32_t nx = left ? __ldg(&ix.gprev[cur]) : rec_link(ix, cur).gnext;
    if (nx == NONE) { if (i == 1) return 0; break; }
    const RecKey k = rec_key(ix, nx);
    gathered += (uint32_t)(k.en - k.st); cur = nx;
    h = (h ^ (((unsigned long long)(uint32_t)k.st << 32) | (uint32_t)k.en)) * 0xC2B2AE3D27D4EB4Full; h ^= h >> 29;
  }
  if (key) *key = h;
  return gathered;
}
// Do two rescues of the same read and side read the same target?  Both chains are walked the way collect_len walks them.
__device__ bool same_target(const DevIndex& ix, bool left, uint32_t rec1, uint32_t rec2, uint32_t qlen) {
  uint32_t gathered = 0, c1 = rec1, c2 = rec2;
  while (qlen > gathered) {
    const uint32_t n1 = left ? __ldg(&ix.gprev[c1]) : rec_link(ix, c1).gnext, n2 = left ? __ldg(&ix.gprev[c2]) : rec_link(ix, c2).gnext;
    if (n1 == NONE || n2 == NONE) return n1 == n2;
    const RecKey k1 = rec_key(ix, n1), k2 = rec_key(ix, n2);
    if (k1.st != k2.st || k1.en != k2.en) return false;
    gathered += (uint32_t)(k1.en - k1.st); c1 = n1; c2 = n2;
  }
  return true;
}

// The target of a rescue, written by the `nl` lanes of a group (lane gl of them): tlen codes into dst -- left: already
// reversed, i.e. nearest base first going leftwards (align_reversed keeps the last qlen+40 bases and reverses them).
__device__ __forceinline__ void gather_target(const DevIndex& ix, int32_t refid, bool left, uint32_t rec, uint32_t tlen, uint8_t* dst, unsigned gl, unsigned nl) {
  const uint64_t base = ix.chrom_off[refid], lim = ix.chrom_off[refid + 1];
  uint32_t w = 0, cur = rec;
  while (w < tlen) {
    const uint32_t nx = left ? __ldg(&ix.gprev[cur]) : rec_link(ix, cur).gnext;
    if (nx == NONE) break;
    const RecKey k = rec_key(ix, nx);
    const uint32_t len = (uint32_t)(k.en - k.st), take = len < tlen - w ? len : tlen - w;
    for (uint32_t i = gl; i < take; i += nl) {
      const uint64_t o = left ? base + ((uint32_t)k.en - 1 - i) - 1 : base + ((uint32_t)k.st + i) - 1;      // 1-based coordinate -> 0-based offset
      dst[w + i] = o < lim ? (uint8_t)genome_code(ix, o) : 4;
    }
    w += take; cur = nx;
  }
}

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

// Problems of one bundle.  Per slot s = 2 * candidate + side (0 left, 1 right): query / target length (0 = no rescue) and the
// offset of its traceback ops (tb) = of its override-coded clip CIGAR (rpool).  Per class: the slots of its problems.
struct KswWork {
  uint32_t* qlen; uint32_t* tlen; uint32_t* tb_off;       // [2 * cap_cand]
  uint32_t* queue;                                        // [KSW_CLASSES][2 * cap_cand]
  uint32_t queue_stride;
  uint32_t* tb; KswOut* out;                              // traceback pool [cap_tb], results per slot
  unsigned long long* key; uint32_t* rec; uint32_t* rep;  // [2 * cap_cand] target key, start record, and the slot whose DP this slot shares (itself: it runs its own)
};

// Pass 1: per slot the problem's sizes, its start record and the key of its target.
__global__ void __launch_bounds__(256) k_rescue_sizes(DevIndex ix, DevBatch b, brb_params p, const uint32_t* __restrict__ c_task,
                                                      const uint32_t* __restrict__ c_first, const uint32_t* __restrict__ c_last,
                                                      const uint32_t* __restrict__ c_cap, KswWork K, Ctl* __restrict__ ctl) {
  if (ctl_overflow(ctl)) return;
  const uint32_t slots = 2 * ctl->n_cand;
  for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < slots; slot += gridDim.x * blockDim.x) {
    const uint32_t c = slot >> 1; const int side = slot & 1;
    uint32_t qlen = 0, tlen = 0, rec = NONE; unsigned long long key = 0;
    if (c_cap[c]) {
      const uint32_t task = c_task[c], a = task >> 1; const int s = task & 1;
      uint32_t total;
      if (rescue_inputs(ix, b, p, a, s, c_first[c], c_last[c], side, rec, total)) {
        const uint32_t got = collect_len(ix, side == 0, rec, total, &key);
        if (got) { qlen = total; tlen = got < total + KSW_TAIL ? got : total + KSW_TAIL; }   // align_reversed keeps the last qlen+40; right truncates to qlen+40
      }
    }
    K.qlen[slot] = qlen; K.tlen[slot] = tlen; K.key[slot] = key; K.rec[slot] = rec;
  }
}

// Pass 2: identical problems of one read (same side, same clip, same neighbour exons -- the isoforms of a gene mostly share them) run
// ONE DP: a slot looks back over the earlier candidates of its alignment for the first one with the same sizes and key, checks the
// exon coordinates themselves, and shares that slot's result (K.rep).  Every slot keeps its own room for the clip CIGAR; only
// representatives get traceback work and a place in the queue of their size class.
__global__ void __launch_bounds__(256) k_rescue_plan(DevIndex ix, DevBatch b, const uint32_t* __restrict__ c_task, const uint4* __restrict__ a_cand,
                                                     KswWork K, Ctl* __restrict__ ctl) {
  if (ctl_overflow(ctl)) return;
  const uint32_t slots = 2 * ctl->n_cand, cap_tb = ctl->cap_tb;
  const unsigned lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  unsigned long long cells = 0, cells_run = 0;
  for (uint32_t base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; base < slots; base += gridDim.x * blockDim.x) {
    const uint32_t slot = base + lane;
    uint32_t qlen = 0, tlen = 0, rep = slot;
    if (slot < slots) {
      qlen = K.qlen[slot]; tlen = K.tlen[slot];
      if (qlen) {
        const uint32_t c = slot >> 1, side = slot & 1u, a = c_task[c] >> 1;
        const unsigned long long key = K.key[slot];
        for (uint32_t c2 = a_cand[a].x; c2 < c; c2++) {
          const uint32_t s2 = 2 * c2 + side;
          if (K.qlen[s2] == qlen && K.tlen[s2] == tlen && K.key[s2] == key && (c_task[c2] >> 1) == a &&
              same_target(ix, side == 0, K.rec[slot], K.rec[s2], qlen)) { rep = s2; break; }
        }
      }
    }
    const bool runs = qlen && rep == slot;
    // traceback / clip-CIGAR room: one atomic per warp
    const uint32_t want = qlen ? qlen + tlen + 2 : 0;
    const uint32_t incl = warp_incl_scan(want, lane), tot = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t tb0 = 0;
    if (lane == 31 && tot) tb0 = atomicAdd(&ctl->n_tb, tot);
    tb0 = __shfl_sync(0xffffffffu, tb0, 31);
    const bool fits = (unsigned long long)tb0 + tot <= cap_tb;
    if (slot < slots) { K.tb_off[slot] = tb0 + incl - want; K.rep[slot] = rep; }
    const int cls = (runs && fits) ? ksw_class_of(qlen) : -1;                    // (no room: the host grows the pool and replays)
#pragma unroll
    for (int k = 0; k < KSW_CLASSES; k++) {                // queue of its class: one atomic per warp and class
      const uint32_t B = __ballot_sync(0xffffffffu, cls == k);
      if (!B) continue;
      uint32_t q0 = 0;
      if (lane == (unsigned)(__ffs(B) - 1)) q0 = atomicAdd(&ctl->kq_n[k], (uint32_t)__popc(B));
      q0 = __shfl_sync(0xffffffffu, q0, __ffs(B) - 1);
      if (cls == k) K.queue[(size_t)k * K.queue_stride + q0 + __popc(B & lt_mask)] = slot;
    }
    if (cls == KSW_BIG) {                                  // scratch sizing of k_ksw_big: largest direction matrix, longest query and target
      atomicMax(&ctl->ksw_max_mat, (unsigned long long)(qlen + tlen - 1) * (qlen < tlen ? qlen : tlen));
      atomicMax(&ctl->ksw_max_tlen, (unsigned long long)tlen); atomicMax(&ctl->ksw_max_qlen, (unsigned long long)qlen);
    }
    if (qlen) cells += (unsigned long long)qlen * tlen;
    if (runs) cells_run += (unsigned long long)qlen * tlen;
  }
  for (int d = 16; d; d >>= 1) { cells += __shfl_down_sync(0xffffffffu, cells, d); cells_run += __shfl_down_sync(0xffffffffu, cells_run, d); }
  if (lane == 0 && cells) { atomicAdd(&ctl->ksw_cells, cells); atomicAdd(&ctl->ksw_cells_run, cells_run); }
}

// Where a problem's sequences come from: the batch (rescue) or explicit arrays (brb_ksw_extend, the parity tests).
struct KswSrc {
  // rescue
  const uint32_t* c_task; const uint32_t* c_first; const uint32_t* c_last;
  // direct (q != nullptr): problem p = slot
  const uint8_t* q; const uint8_t* t; const uint32_t* q_off; const uint32_t* t_off;
};

__device__ __forceinline__ uint32_t onehot_q(uint32_t c) { return c < 4 ? (1u << c) : 0x10u; }   // query N and target N never "match"
__device__ __forceinline__ uint32_t onehot_t(uint32_t c) { return c < 4 ? (1u << c) : 0x20u; }

// One queue of problems (class `cls`), G lanes per problem, 2 * W cells per lane (G * 2 * W >= the class's query length).
// Per-group scratch: `pitch` = G * 2 * W direction bytes per anti-diagonal, rows_cap anti-diagonals.
template <int G, int W>
__global__ void __launch_bounds__(32 * KSW_WARPS) k_ksw_group(DevIndex ix, DevBatch b, KswSrc S, KswWork K, int cls, uint32_t n_direct, Ctl* __restrict__ ctl,
                                                              uint8_t* __restrict__ pmat_all, uint32_t rows_cap) {
  constexpr int NG = 32 / G, CPL = 2 * W, QMAX = G * CPL, TMAX = QMAX + KSW_TAIL, PITCH = QMAX;
  __shared__ uint8_t s_t[KSW_WARPS][NG][TMAX + 8];
  if (!n_direct && ctl_overflow(ctl)) return;
  const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5, grp = lane / G, gl = lane % G;
  const uint32_t n_prob = n_direct ? n_direct : ctl->kq_n[cls];
  const uint32_t* queue = K.queue + (size_t)cls * K.queue_stride;
  uint8_t* pmat = pmat_all + ((size_t)(blockIdx.x * KSW_WARPS + wib) * NG + grp) * (size_t)rows_cap * PITCH;
  uint8_t* tg = s_t[wib][grp];
  const int q = 4, e = 1, qe = q + e, zdrop = 40;
  const uint32_t j0 = gl * CPL;
  for (;;) {
    // ---- next problem of every group
    uint32_t pi = 0;
    if (gl == 0) pi = atomicAdd(&ctl->kq_t[cls], 1u);
    pi = __shfl_sync(0xffffffffu, pi, grp * G);
    const bool have = pi < n_prob;
    if (!__any_sync(0xffffffffu, have)) break;
    uint32_t slot = 0; int qlen = 0, tlen = 0;
    uint32_t Q[W], T[W], X[W], V[W], Y[W], U[W], JW[W];
#pragma unroll
    for (int w = 0; w < W; w++) { Q[w] = 0x00100010u; T[w] = 0x00200020u; X[w] = 0; Y[w] = 0; U[w] = 0; V[w] = 0x00040004u; JW[w] = (j0 + 2 * w) | ((j0 + 2 * w + 1) << 16); }
    if (gl == 0) V[0] = 0x00040000u;                                   // v(r-1, -1) = r ? q : 0 for the cell t = 0: q at j > 0, 0 at j = 0
    if (have) {
      slot = queue[pi];
      if (S.q) {
        qlen = (int)(S.q_off[slot + 1] - S.q_off[slot]); tlen = (int)(S.t_off[slot + 1] - S.t_off[slot]);
        const uint8_t* qs = S.q + S.q_off[slot]; const uint8_t* ts = S.t + S.t_off[slot];
#pragma unroll
        for (int w = 0; w < W; w++) {
          const int ja = j0 + 2 * w, jb = ja + 1;
          Q[w] = (ja < qlen ? onehot_q(qs[ja]) : 0x10u) | ((jb < qlen ? onehot_q(qs[jb]) : 0x10u) << 16);
        }
        for (int i = gl; i < tlen; i += G) tg[i] = ts[i];
      } else {
        qlen = (int)K.qlen[slot]; tlen = (int)K.tlen[slot];
        const uint32_t c = slot >> 1; const int side = slot & 1;
        const uint32_t a = S.c_task[c] >> 1;
        const uint32_t g = b.aln_group[a];
        const uint8_t* seq = b.seq + b.group_seq_off[g];
        const int seq_len = (int)(b.group_seq_off[g + 1] - b.group_seq_off[g]);
#pragma unroll
        for (int w = 0; w < W; w++) {                                  // get_left_rescue_query + reversal (:320-330, 368-395) / get_right_rescue_query (:500-511)
          const int ja = j0 + 2 * w, jb = ja + 1;
          const uint32_t ca = ja < qlen ? seq[side == 0 ? qlen - 1 - ja : seq_len - qlen + ja] : 4u;
          const uint32_t cb = jb < qlen ? seq[side == 0 ? qlen - 1 - jb : seq_len - qlen + jb] : 4u;
          Q[w] = onehot_q(ca) | (onehot_q(cb) << 16);
        }
        gather_target(ix, b.refid[a], side == 0, side == 0 ? S.c_first[c] : S.c_last[c], (uint32_t)tlen, tg, gl, G);
      }
    }
    __syncwarp();
    // ---- the anti-diagonals
    const int n_rows = have && qlen > 0 && tlen > 0 ? qlen + tlen - 1 : 0;
    int H0 = 0, last_H0_t = 0, emax = 0, emax_t = -1, emax_q = -1, score = BRB_KSW_NEG_INF;
    bool done = n_rows == 0, zdropped = false;
    for (int r = 0; __any_sync(0xffffffffu, !done); ++r) {
      // what arrives from cell j-1: y, u of the previous diagonal and the target base (the group's first lane: the boundary)
      uint32_t cy = __shfl_up_sync(0xffffffffu, Y[W - 1], 1, G), cu = __shfl_up_sync(0xffffffffu, U[W - 1], 1, G), ct = __shfl_up_sync(0xffffffffu, T[W - 1], 1, G);
      if (gl == 0) { cy = 0; cu = r ? (uint32_t)q << 16 : 0u; ct = onehot_t(r < tlen ? tg[r] : 4u) << 16; }
      const uint32_t rp = (uint32_t)r * 0x00010001u;
      uint32_t D[W];
#pragma unroll
      for (int w = W - 1; w >= 0; w--) {                               // high words first: word w shifts in the old word w-1
        const uint32_t ys = __funnelshift_l(w ? Y[w - 1] : cy, Y[w], 16), us = __funnelshift_l(w ? U[w - 1] : cu, U[w], 16);
        T[w] = __funnelshift_l(w ? T[w - 1] : ct, T[w], 16);
        // score + 2(q+e): 11 match, 6 mismatch, 9 when either base is N (sc_N = -e)
        const uint32_t mt = __vminu2(T[w] & Q[w], 0x00010001u), nn = __vminu2((T[w] | Q[w]) & 0x00300030u, 0x00010001u);
        const uint32_t zs = 0x00060006u + mt * 5u + nn * 3u;
        const uint32_t a = __vadd2(X[w], V[w]), bb = __vadd2(ys, us);
        const uint32_t z1 = __vmaxs2(zs, a), d1 = __vminu2(z1 - zs, 0x00010001u);          // a > z
        const uint32_t z2 = __vmaxs2(z1, bb), d2 = __vminu2(z2 - z1, 0x00010001u);         // b > max(z, a)
        const uint32_t z = __vmins2(z2, 0x000b000bu);
        const uint32_t un = z - V[w], vn = z - us;                     // (non-negative by construction: no borrow between the halves)
        const uint32_t zq = z - 0x00040004u;
        const uint32_t xn = __vmaxs2(a, zq) - zq, yn = __vmaxs2(bb, zq) - zq;
        D[w] = (d1 & ~d2) + d2 * 2u + __vminu2(xn, 0x00010001u) * 8u + __vminu2(yn, 0x00010001u) * 16u;
        // a cell exists from anti-diagonal r = j on: until then x, v keep their initial values
        const uint32_t nst = __vminu2(__vmaxu2(JW[w], rp) - rp, 0x00010001u) * 0xffffu;    // 0xffff where j > r
        X[w] = (xn & ~nst) | (X[w] & nst); V[w] = (vn & ~nst) | (V[w] & nst);
        U[w] = un; Y[w] = yn;
      }
      if (r < (int)rows_cap) {
        if (W == 2) *reinterpret_cast<uint32_t*>(pmat + (size_t)r * PITCH + 4 * gl) = __byte_perm(D[0], D[1], 0x6420);
        else {
#pragma unroll
          for (int w = 0; w < W; w += 2) *reinterpret_cast<uint32_t*>(pmat + (size_t)r * PITCH + CPL * gl + 2 * w) = __byte_perm(D[w], D[w + 1], 0x6420);
        }
      }
      // approximate max / z-drop along the greedy diagonal cursor (ksw2_extz2_sse.cpp:284-300), per group
      const int st = r - qlen + 1 > 0 ? r - qlen + 1 : 0, en = r < tlen - 1 ? r : tlen - 1;
      const int jv = r - last_H0_t, ju = jv - 1;                       // v(r, last_H0_t), u(r, last_H0_t + 1)
      uint32_t pv = 0, pu = 0;
#pragma unroll
      for (int w = 0; w < W; w++) { if (((jv >> 1) & (W - 1)) == w) pv = V[w]; if (((ju >> 1) & (W - 1)) == w) pu = U[w]; }
      pv = (jv & 1) ? pv >> 16 : pv & 0xffffu; pu = (ju & 1) ? pu >> 16 : pu & 0xffffu;
      const int v_at = (int)__shfl_sync(0xffffffffu, pv, (jv >= 0 ? jv / CPL : 0) & (G - 1), G);
      const int u_at = (int)__shfl_sync(0xffffffffu, pu, (ju >= 0 ? ju / CPL : 0) & (G - 1), G);
      if (!done) {
        if (r > 0) {
          const bool in0 = last_H0_t >= st && last_H0_t <= en, in1 = last_H0_t + 1 >= st && last_H0_t + 1 <= en;
          if (in0 && in1) {
            const int d0 = v_at - qe, d1 = u_at - qe;
            if (d0 > d1) H0 += d0; else { H0 += d1; ++last_H0_t; }
          } else if (in0) H0 += v_at - qe;
          else { ++last_H0_t; H0 += u_at - qe; }                        // (then u(r, last_H0_t + 1) is the cell fetched above)
          const int t = last_H0_t;                                     // ksw_apply_zdrop
          if (H0 > emax) { emax = H0; emax_t = t; emax_q = r - t; }
          else if (t >= emax_t && r - t >= emax_q) {
            const int tl = t - emax_t, ql = (r - t) - emax_q, l = tl > ql ? tl - ql : ql - tl;
            if (emax - H0 > zdrop + l * e) { zdropped = true; done = true; }
          }
        } else { H0 = v_at - qe - qe; last_H0_t = 0; }
        if (r == n_rows - 1 && en == tlen - 1 && !zdropped) score = H0;
        if (r == n_rows - 1) done = true;
      }
    }
    __syncwarp();
    // ---- ksw_backtrack(is_rot=1) from (max_t, max_q) by the group's first lane; run-length ops in traceback order
    if (have && gl == 0) {
      uint32_t n = 0;
      uint32_t* tb = K.tb + K.tb_off[slot];
      if (emax_t >= 0 && emax_q >= 0) {
        int i = emax_t, j = emax_q, state = 0;
        uint32_t cur = 0;                                              // the open run: (len << 4) | op
        auto push = [&](uint32_t op, int len) {
          if (n == 0 || op != (cur & 0xfu)) { if (n) tb[n - 1] = cur; cur = ((uint32_t)len << 4) | op; n++; } else cur += (uint32_t)len << 4;
        };
        while (i >= 0 && j >= 0) {
          const uint32_t tmp = pmat[(size_t)(i + j) * PITCH + j];
          if (state == 0) state = tmp & 7;
          else if (!((tmp >> (state + 2)) & 1)) state = 0;
          if (state == 0) state = tmp & 7;
          if (state == 0) { push(0, 1); --i; --j; }
          else if (state == 1 || state == 3) { push(2, 1); --i; }
          else { push(1, 1); --j; }
        }
        if (i >= 0) push(2, i + 1);
        if (j >= 0) push(1, j + 1);
        if (n) tb[n - 1] = cur;
      }
      KswOut o; o.max = emax; o.score = score; o.max_q = emax_q; o.max_t = emax_t; o.zdropped = zdropped ? 1u : 0u; o.n_cig = n;
      K.out[slot] = o;
    }
    __syncwarp();
  }
}

// Problems beyond the register kernel (query > 512 bases): one warp each, sequences and state in per-warp global scratch sized
// for the largest such problem of the PREVIOUS sequence (capacities persist; a bundle that needs more makes the host grow them
// and replay, like every other pool).
struct KswBig { uint8_t* pmat; size_t pmat_stride; uint8_t* seq; uint32_t seq_stride; int8_t* state; uint32_t state_stride; };

__global__ void __launch_bounds__(32 * KSW_WARPS) k_ksw_big(DevIndex ix, DevBatch b, KswSrc S, KswWork K, uint32_t n_direct, Ctl* __restrict__ ctl, KswBig B) {
  __shared__ int8_t sh[KSW_WARPS][4][KSW_SH_T];
  if (!n_direct && ctl_overflow(ctl)) return;
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t w = blockIdx.x * KSW_WARPS + wib;
  const uint32_t n_prob = n_direct ? n_direct : ctl->kq_n[KSW_BIG];
  const uint32_t* queue = K.queue + (size_t)KSW_BIG * K.queue_stride;
  uint8_t* pmat = B.pmat + (size_t)w * B.pmat_stride;
  uint8_t* qbuf = B.seq + (size_t)w * B.seq_stride * 2; uint8_t* tbuf = qbuf + B.seq_stride;
  for (;;) {
    uint32_t pi = 0;
    if (lane == 0) pi = atomicAdd(&ctl->kq_t[KSW_BIG], 1u);
    pi = __shfl_sync(0xffffffffu, pi, 0);
    if (pi >= n_prob) break;
    const uint32_t slot = queue[pi];
    int qlen, tlen;
    if (S.q) { qlen = (int)(S.q_off[slot + 1] - S.q_off[slot]); tlen = (int)(S.t_off[slot + 1] - S.t_off[slot]); }
    else { qlen = (int)K.qlen[slot]; tlen = (int)K.tlen[slot]; }
    const unsigned long long mat = (unsigned long long)(qlen + tlen - 1) * (qlen < tlen ? qlen : tlen);
    if ((uint32_t)qlen > B.seq_stride || (uint32_t)tlen > B.seq_stride || mat > B.pmat_stride ||
        (tlen + 1 > KSW_SH_T && (uint32_t)(tlen + 1) > B.state_stride)) {              // scratch too small: counted, the host grows it and replays
      if (lane == 0) { atomicAdd(&ctl->ksw_big_skipped, 1u); atomicMax(&ctl->ksw_max_mat, mat); atomicMax(&ctl->ksw_max_tlen, (unsigned long long)tlen);
                       atomicMax(&ctl->ksw_max_qlen, (unsigned long long)qlen); }
      continue;
    }
    const uint8_t* qp; const uint8_t* tp;
    if (S.q) { qp = S.q + S.q_off[slot]; tp = S.t + S.t_off[slot]; }
    else {
      const uint32_t c = slot >> 1; const int side = slot & 1;
      const uint32_t a = S.c_task[c] >> 1, g = b.aln_group[a];
      const uint8_t* seq = b.seq + b.group_seq_off[g];
      const int seq_len = (int)(b.group_seq_off[g + 1] - b.group_seq_off[g]);
      for (int i = lane; i < qlen; i += 32) qbuf[i] = seq[side == 0 ? qlen - 1 - i : seq_len - qlen + i];
      gather_target(ix, b.refid[a], side == 0, side == 0 ? S.c_first[c] : S.c_last[c], (uint32_t)tlen, tbuf, lane, 32);
      __syncwarp();
      qp = qbuf; tp = tbuf;
    }
    int8_t *su, *sv, *sx, *sy;
    if (tlen + 1 <= KSW_SH_T) { su = sh[wib][0]; sv = sh[wib][1]; sx = sh[wib][2]; sy = sh[wib][3]; }
    else { int8_t* g = B.state + (size_t)w * B.state_stride * 4; su = g; sv = g + B.state_stride; sx = g + 2 * B.state_stride; sy = g + 3 * B.state_stride; }
    KswOut o;
    ksw_extend_warp(qp, qlen, tp, tlen, su, sv, sx, sy, pmat, K.tb + K.tb_off[slot], o);
    if (lane == 0) K.out[slot] = o;
    __syncwarp();
  }
}

// build_left_clip_segment / build_right_clip_segment: KswOut + traceback-order ops -> RescueSeg + override-coded CIGAR.
__global__ void __launch_bounds__(128) k_rescue_finish(KswWork K, uint32_t* __restrict__ rpool, RescueSeg* __restrict__ r_left,
                                                       RescueSeg* __restrict__ r_right, uint32_t* __restrict__ c_cap, Ctl* __restrict__ ctl) {
  if (ctl_overflow(ctl)) return;
  const uint32_t slots = 2 * ctl->n_cand;
  unsigned long long n_res = 0, bytes = 0;
  for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < slots; slot += gridDim.x * blockDim.x) {
    const uint32_t c = slot >> 1; const int side = slot & 1;
    RescueSeg sgm; sgm.ok = 0; sgm.score = 0; sgm.rc = 0; sgm.cig_off = 0; sgm.cig_len = 0;
    const uint32_t qlen = K.qlen[slot];
    if (qlen) {
      const uint32_t rs = K.rep ? K.rep[slot] : slot;                       // the slot that ran the DP of this problem
      const KswOut o = K.out[rs];
      const uint32_t tbo = K.tb_off[slot];
      if (o.max >= 10 && o.score != BRB_KSW_NEG_INF) {
        const uint32_t* T = K.tb + K.tb_off[rs]; const uint32_t n = o.n_cig;          // T[0] is the LAST op of the forward ksw cigar
        int32_t qc = 0, rc = 0;
        for (uint32_t k = 0; k < n; k++) { const uint32_t op = T[k] & 0xfu, len = T[k] >> 4; if (op == 0 || op == 1) qc += len; if (op == 0 || op == 2) rc += len; }
        uint32_t* out = rpool + tbo; uint32_t m = 0;
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
        sgm.ok = 1; sgm.score = o.max; sgm.rc = rc; sgm.cig_off = tbo; sgm.cig_len = m;
        atomicAdd(&c_cap[c], m);
      }
      n_res++; bytes += (qlen + 1) / 2 + (K.tlen[slot] + 1) / 2;
    }
    (side == 0 ? r_left : r_right)[c] = sgm;
  }
  if (n_res) { atomicAdd(&ctl->n_rescue, n_res); atomicAdd(&ctl->rescue_bytes, bytes); }
}

__global__ void k_aln_group(DevBatch b, uint32_t* __restrict__ aln_group) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= b.n_groups) return;
  for (uint32_t a = b.group_off[g]; a < b.group_off[g + 1]; a++) aln_group[a] = g;
}

}  // namespace brb
