// Kernel group 4: CIGAR projection -- merge the alignment's real CIGAR with the ideal (transcript-side)
// CIGAR of the match.  Replaces merge_ops / merge_cigars / get_new_cigar (reference src/bam.cpp:22-315,
// 443-472).  Device function, called by k_back for every emitted row that is not the trivial one-M case.
#pragma once
#include "dev_common.cuh"

namespace brb {

constexpr uint32_t OP_SKIP = 0x5f;   // '_' : "emit nothing" marker of merge_ops

__device__ __forceinline__ uint32_t merge_ops(uint32_t r, uint32_t i) {        // src/bam.cpp:22-111, first match wins
  const bool rMS = (r == CM || r == CS);
  if (rMS && i == OV_S) return CS;
  if (rMS && i == OV_M) return CM;
  if (rMS && i == OV_I) return CI;
  if (rMS && i == OV_D) return CD;
  if (r == CD && (i == CS || i == OV_S)) return OP_SKIP;
  if (r == CD && i == OV_M) return CD;
  if (r == CI && i == OV_S) return CS;
  if (r == CI && i == OV_M) return CI;
  if (i == OV_S) return CS;
  if (i == OV_M) return CM;
  if (i == OV_I) return CI;
  if (i == OV_D) return CD;
  if (r == CP) return i;
  if (r == CH) return CH;
  if (r == CI && i == CS) return CS;
  if (i == CS || i == CD || i == CI) return i;
  if (r == CS || r == CD || r == CI) return r;
  if (i == CM || i == CEQ) return CM;
  if (i == CX) return CX;
  if (r == CM || r == CEQ) return CM;
  if (r == CX) return CX;
  return i;
}

struct CigOut {                 // coalescing appender; the open (last) op lives in a register
  uint32_t* p; uint32_t n; uint32_t last; bool saw_i;
  __device__ __forceinline__ void add(uint32_t op, uint32_t len) {
    if (len == 0 || op == OP_SKIP) return;
    if (n > 0 && (last & 0xfu) == op) last += (len << 4);
    else { if (n > 0) p[n - 1] = last; last = (len << 4) | op; n++; saw_i |= (op == CI); }
  }
  __device__ __forceinline__ void finish() { if (n > 0) p[n - 1] = last; }
};

// The common short-read row: the real CIGAR only has M/=/X and N, the ideal CIGAR is the single word `iword` = one
// M of the same aligned length.  merge_cigars then skips every N, turns every chunk into M (merge_ops) and both
// sides end together: the result is the ideal CIGAR itself.  Returns false when the row needs the general merge.
__device__ __forceinline__ bool merge_is_trivial(const uint32_t* __restrict__ real, uint32_t n_real, uint32_t iword) {
  if (n_real == 0 || (iword & 0xfu) != CM) return false;
  uint32_t sum = 0;
  for (uint32_t i = 0; i < n_real; i++) {
    const uint32_t rw = __ldg(&real[i]), op = rw & 0xfu;
    if (op == CM || op == CEQ || op == CX) sum += rw >> 4; else if (op != CN) return false;
  }
  return sum > 0 && sum == (iword >> 4);
}

// Returns the number of words written to `out` (capacity >= n_real + n_ideal + 2).
// A one-word ideal CIGAR is passed by value in `iword` (match records keep it inline; `ideal` is then unused).
__device__ __noinline__ uint32_t merge_cigars(const uint32_t* __restrict__ real, uint32_t n_real, const uint32_t* __restrict__ ideal_,
                                 uint32_t n_ideal, uint32_t iword, uint32_t* __restrict__ out) {
  auto ideal = [&](uint32_t k) -> uint32_t { return n_ideal == 1 ? iword : ideal_[k]; };
  CigOut res; res.p = out; res.n = 0; res.last = 0; res.saw_i = false;
  uint32_t front_hard = 0, front_soft = 0, ci = 0;
  if (n_real > 0 && (real[0] & 0xfu) == CH) { front_hard = real[0] >> 4; ci++; }
  if (ci < n_real && (real[ci] & 0xfu) == CS) front_soft = real[ci] >> 4;
  uint32_t ri = 0, ii = 0, rpos = 0, ipos = 0;
  uint32_t clips = front_hard;
  while (clips > 0 && ri < n_real) {                                            // (1) leading hard clip
    const uint32_t rw = real[ri], avail = (rw >> 4) - rpos, chunk = clips < avail ? clips : avail;
    res.add(rw & 0xfu, chunk);
    clips -= chunk; rpos += chunk;
    if (rpos >= (rw >> 4)) { ri++; rpos = 0; }
  }
  clips = front_soft;
  while (clips > 0 && ri < n_real) {                                            // (2) leading soft clip vs override ops
    const uint32_t rw = real[ri], rop = rw & 0xfu, rrem = (rw >> 4) - rpos;
    const bool have_i = ii < n_ideal;
    const uint32_t iw = have_i ? ideal(ii) : 0, iop = have_i ? (iw & 0xfu) : 0xffu, irem = have_i ? (iw >> 4) - ipos : 0xffffffffu;
    const bool ov = have_i && (iop == OV_M || iop == OV_D || iop == OV_I || iop == OV_S);
    if (ov) {
      if (iop == OV_D) {
        res.add(merge_ops(rop, iop), irem);
        ipos += irem;
        if (ipos >= (iw >> 4)) { ii++; ipos = 0; }
      } else {
        uint32_t chunk = clips; if (chunk > rrem) chunk = rrem; if (chunk > irem) chunk = irem;
        res.add(merge_ops(rop, iop), chunk);
        clips -= chunk; rpos += chunk; ipos += chunk;
        if (rpos >= (rw >> 4)) { ri++; rpos = 0; }
        if (ipos >= (iw >> 4)) { ii++; ipos = 0; }
      }
    } else {
      uint32_t chunk = clips; if (chunk > rrem) chunk = rrem;
      res.add(merge_ops(rop, iop), chunk);
      clips -= chunk; rpos += chunk;
      if (rpos >= (rw >> 4)) { ri++; rpos = 0; }
    }
  }
  while (ri < n_real || ii < n_ideal) {                                         // (3) main two-cursor merge
    if (ri >= n_real) { const uint32_t iw = ideal(ii); res.add(iw & 0xfu, (iw >> 4) - ipos); ii++; ipos = 0; continue; }
    if (ii >= n_ideal) { const uint32_t rw = real[ri]; res.add(rw & 0xfu, (rw >> 4) - rpos); ri++; rpos = 0; continue; }
    const uint32_t rw = real[ri], iw = ideal(ii), rop = rw & 0xfu, iop = iw & 0xfu;
    const uint32_t rrem = (rw >> 4) - rpos, irem = (iw >> 4) - ipos;
    if (rop == CN) { ri++; rpos = 0; }
    else if (rop == CD && (iop == CS || iop == OV_S || iop == CI || iop == OV_I)) {
      const uint32_t chunk = rrem < irem ? rrem : irem;
      rpos += chunk; ipos += chunk;
      if (rpos >= (rw >> 4)) { ri++; rpos = 0; }
      if (ipos >= (iw >> 4)) { ii++; ipos = 0; }
    } else if (rop == CI) { res.add(CI, rrem); ri++; rpos = 0; }
    else if (iop == CD || iop == OV_D) { res.add(CD, irem); ii++; ipos = 0; }
    else {
      const uint32_t chunk = rrem < irem ? rrem : irem;
      res.add(merge_ops(rop, iop), chunk);
      rpos += chunk; ipos += chunk;
      if (rpos >= (rw >> 4)) { ri++; rpos = 0; }
      if (ipos >= (iw >> 4)) { ii++; ipos = 0; }
    }
  }
  res.finish();
  if (!res.saw_i) return res.n;                                                 // (4) and (5) only ever change a CIGAR that contains an I
  for (uint32_t i = 1; i + 1 < res.n; i++) {                                    // (4) I between clips becomes that clip
    if ((out[i] & 0xfu) != CI) continue;
    const uint32_t prev = out[i - 1] & 0xfu, next = out[i + 1] & 0xfu;
    if ((prev == CS || prev == CH) && (next == CS || next == CH)) out[i] = ((out[i] >> 4) << 4) | prev;
  }
  uint32_t n = 0;                                                               // (5) re-coalesce
  for (uint32_t i = 0; i < res.n; i++) {
    const uint32_t op = out[i] & 0xfu, len = out[i] >> 4;
    if (n > 0 && (out[n - 1] & 0xfu) == op) out[n - 1] += (len << 4); else out[n++] = (len << 4) | op;
  }
  return n;
}

// ---------------------------------------------------------------------------------------------------------------------
// Warp-cooperative merge_cigars for long CIGARs (long reads: 10^2..10^4 ops).  Same result as the serial code above, word
// for word; the serial two-cursor loop is re-expressed so that every real op can work out its own output pieces:
//   * both CIGARs advance along one axis: a real op consumes its length unless it is N or I, an ideal op unless it is D or
//     the override D.  Inclusive prefix sums give every op its interval [s, e) on that axis.
//   * a consuming real op emits one piece per ideal op it overlaps (merge_ops of the pair, nothing for real D under an
//     ideal S/I), the ideal Ds located in [s, e) in between, and -- once the ideal CIGAR is used up -- the rest of itself
//     verbatim.  Real I emits itself; real N nothing (or itself when the ideal CIGAR is already used up).  Whatever the
//     ideal CIGAR has left when the real one ends is copied verbatim.  Ties on one coordinate resolve like the serial
//     loop's if-chain: real N/I first, then the ideal Ds, then the next chunk.
//   * the pieces, in real-op order, are exactly the serial loop's add() calls; one run-length pass coalesces them.
// The leading-clip phases (1) and (2) run as in the serial code (a handful of iterations, every lane redundantly); the
// I-between-clips fix-up (4)/(5) runs on lane 0 in the rare case it applies.  A zero-length real op (not valid BAM)
// makes the serial loop's tie order depend on it: such a CIGAR is handed back (MERGE_SERIAL) and merged by the serial code.
// tools/merge_proto.cpp is this formulation as plain loops, checked against the oracle on 3M random CIGAR pairs.
constexpr uint32_t MERGE_SERIAL = 0xffffffffu;
constexpr uint32_t MERGE_COOP_MIN = 48;          // real + ideal ops from which a row is merged by the whole warp

// merge_ops as a table for the cooperative code (the if-chain, inlined at several places, was a tenth of its instructions
// and a good part of its code size): row r = 16 nibbles, nibble i = merge_ops(r, i), 15 = nothing is emitted.  The swallow
// rule of the main loop (real D under an ideal S / I) is folded in.  Built at compile time from the chain itself.
constexpr uint32_t merge_ops_c(uint32_t r, uint32_t i) {
  const bool rMS = (r == CM || r == CS);
  if (rMS && i == OV_S) return CS;
  if (rMS && i == OV_M) return CM;
  if (rMS && i == OV_I) return CI;
  if (rMS && i == OV_D) return CD;
  if (r == CD && (i == CS || i == OV_S)) return 15;
  if (r == CD && i == OV_M) return CD;
  if (r == CI && i == OV_S) return CS;
  if (r == CI && i == OV_M) return CI;
  if (i == OV_S) return CS;
  if (i == OV_M) return CM;
  if (i == OV_I) return CI;
  if (i == OV_D) return CD;
  if (r == CP) return i;
  if (r == CH) return CH;
  if (r == CI && i == CS) return CS;
  if (i == CS || i == CD || i == CI) return i;
  if (r == CS || r == CD || r == CI) return r;
  if (i == CM || i == CEQ) return CM;
  if (i == CX) return CX;
  if (r == CM || r == CEQ) return CM;
  if (r == CX) return CX;
  return i;
}
constexpr unsigned long long merge_ops_row(uint32_t r) {
  unsigned long long w = 0;
  for (uint32_t i = 0; i < 16; i++) {
    uint32_t v = merge_ops_c(r, i);
    if (r == CD && (i == CS || i == OV_S || i == CI || i == OV_I)) v = 15;      // main loop: both sides advance, nothing is emitted
    w |= (unsigned long long)(v & 15u) << (4 * i);
  }
  return w;
}
__constant__ unsigned long long c_merge_rows[16] = {
  merge_ops_row(0), merge_ops_row(1), merge_ops_row(2), merge_ops_row(3), merge_ops_row(4), merge_ops_row(5), merge_ops_row(6), merge_ops_row(7),
  merge_ops_row(8), merge_ops_row(9), merge_ops_row(10), merge_ops_row(11), merge_ops_row(12), merge_ops_row(13), merge_ops_row(14), merge_ops_row(15)};
// (an ideal op that passes through unchanged can only be 0..13: the value 15 is free to mean "nothing")
__device__ __forceinline__ uint32_t merge_ops_tab(uint32_t r, uint32_t i) { return (uint32_t)(c_merge_rows[r & 15u] >> (4 * (i & 15u))) & 15u; }

// scratch words merge_cigars_warp needs besides `out` (capacity n_real + n_ideal + 2)
__host__ __device__ __forceinline__ uint32_t merge_warp_scratch(uint32_t n_real, uint32_t n_ideal) { return n_real + 4 * n_ideal + 16; }

__device__ __noinline__ uint32_t merge_cigars_warp(const uint32_t* __restrict__ real, uint32_t nR, const uint32_t* __restrict__ ideal_, uint32_t nI,
                                                   uint32_t iword, uint32_t* __restrict__ out, uint32_t* __restrict__ scr, unsigned lane) {
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint32_t* EI = scr;                                  // [nI] inclusive prefix sums of the ideal ops' consumed lengths
  uint32_t* IW = scr + nI + 1;                         // [nI] the ideal words from the cursor on (first one already shortened by ipos)
  uint32_t* pc = scr + 2 * nI + 2;                     // pieces, (len << 4) | op
  uint32_t np = 0;
  auto ideal = [&](uint32_t k) -> uint32_t { return nI == 1 ? iword : ideal_[k]; };
  auto add = [&](uint32_t op, uint32_t len) { if (len && op != OP_SKIP) { if (lane == 0) pc[np] = (len << 4) | op; np++; } };
  // ---- phases (1), (2): leading hard / soft clip, as in merge_cigars (uniform: every lane runs them, lane 0 writes)
  uint32_t front_hard = 0, front_soft = 0, ci = 0;
  if (nR > 0 && (real[0] & 0xfu) == CH) { front_hard = real[0] >> 4; ci++; }
  if (ci < nR && (real[ci] & 0xfu) == CS) front_soft = real[ci] >> 4;
  uint32_t ri = 0, ii = 0, rpos = 0, ipos = 0, clips = front_hard;
  while (clips > 0 && ri < nR) {
    const uint32_t rw = real[ri], avail = (rw >> 4) - rpos, chunk = clips < avail ? clips : avail;
    add(rw & 0xfu, chunk); clips -= chunk; rpos += chunk;
    if (rpos >= (rw >> 4)) { ri++; rpos = 0; }
  }
  clips = front_soft;
  while (clips > 0 && ri < nR) {
    const uint32_t rw = real[ri], rop = rw & 0xfu, rrem = (rw >> 4) - rpos;
    const bool have_i = ii < nI;
    const uint32_t iw = have_i ? ideal(ii) : 0, iop = have_i ? (iw & 0xfu) : 0xffu, irem = have_i ? (iw >> 4) - ipos : 0xffffffffu;
    const bool ov = have_i && (iop == OV_M || iop == OV_D || iop == OV_I || iop == OV_S);
    if (ov && iop == OV_D) { add(merge_ops(rop, iop), irem); ipos += irem; if (ipos >= (iw >> 4)) { ii++; ipos = 0; } }
    else {
      uint32_t chunk = clips; if (chunk > rrem) chunk = rrem; if (ov && chunk > irem) chunk = irem;
      add(merge_ops(rop, iop), chunk); clips -= chunk; rpos += chunk;
      if (rpos >= (rw >> 4)) { ri++; rpos = 0; }
      if (ov) { ipos += chunk; if (ipos >= (iw >> 4)) { ii++; ipos = 0; } }
    }
  }
  const uint32_t nr = nR - ri, ni = nI - ii;
  // ---- the ideal side from the cursor on: words and prefix sums of the consumed lengths
  uint32_t SI = 0;
  for (uint32_t base = 0; base < ni; base += 32) {
    const uint32_t m = base + lane;
    uint32_t c = 0, w = 0;
    if (m < ni) { w = ideal(ii + m); if (m == 0) w -= ipos << 4; const uint32_t op = w & 0xfu; c = (op == CD || op == OV_D) ? 0u : (w >> 4); }
    const uint32_t incl = warp_incl_scan(c, lane) + SI;
    if (m < ni) { EI[m] = incl; IW[m] = w; }
    SI = __shfl_sync(0xffffffffu, incl, 31);
  }
  __syncwarp();
  // An ideal op is DONE at real coordinate x when it ends before x, or ends at x after consuming something; zero-consumers
  // (D, override D, empty ops) located at x are still pending there.  `adv` moves a cursor past everything done at x.
  auto done_at = [&](uint32_t m, uint32_t x) { const uint32_t em = EI[m]; if (em < x) return true; if (em > x) return false; const uint32_t w = IW[m], op = w & 0xfu; return op != CD && op != OV_D && (w >> 4) > 0; };
  // ---- the real side, 32 ops at a time (lane = op); the pieces of a tile go to `pc` in real-op order
  uint32_t S0 = 0, m_tile = 0;                         // coordinate and ideal cursor at the start of the tile
  for (uint32_t base = 0; base < nr; base += 32) {
    const uint32_t k = base + lane;
    const bool act = k < nr;
    uint32_t rop = CN, rlen = 0;
    if (act) { const uint32_t rw = real[ri + k]; rop = rw & 0xfu; rlen = (rw >> 4) - (k == 0 ? rpos : 0u); }
    if (__any_sync(0xffffffffu, act && rlen == 0)) return MERGE_SERIAL;
    const bool zero_op = rop == CN || rop == CI;
    const uint32_t cons = (act && !zero_op) ? rlen : 0u;
    const uint32_t e = warp_incl_scan(cons, lane) + S0, s = e - cons;
    while (m_tile < ni && done_at(m_tile, S0)) m_tile++;          // (uniform)
    S0 = __shfl_sync(0xffffffffu, e, 31);
    uint32_t m = m_tile;
    if (act) while (m < ni && done_at(m, s)) m++;                 // the ideal cursor when this op is at the front
    // the op's pieces (at most 3 are kept; more -- several ideal Ds under one real op -- is left to the serial code)
    uint32_t cnt = 0, p0 = 0, p1 = 0, p2 = 0;
    auto emit = [&](uint32_t op, uint32_t len) {
      if (len == 0 || op == 15u) return;
      const uint32_t w = (len << 4) | op;
      if (cnt == 0) p0 = w; else if (cnt == 1) p1 = w; else p2 = w;
      cnt++;
    };
    if (act) {
      if (zero_op) {
        if (rop == CI) emit(CI, rlen);
        else if (m >= ni) emit(CN, rlen);                         // the ideal CIGAR is used up: real ops are copied verbatim
      } else {
        uint32_t cur = s;
        for (; m < ni; m++) {
          const uint32_t iw = IW[m], iop = iw & 0xfu, il = iw >> 4;
          const uint32_t ic = (iop == CD || iop == OV_D) ? 0u : il, em = EI[m], sm = em - ic;
          if (ic == 0) {                                          // zero-consumer at coordinate em
            if (em >= e) break;                                   // it waits for the real op that starts there
            if (il) emit(CD, il);
            continue;
          }
          if (sm >= e) break;
          const uint32_t a = sm > cur ? sm : cur, bnd = em < e ? em : e;
          emit(merge_ops_tab(rop, iop), bnd - a);
          cur = bnd;
          if (em > e) break;                                      // the ideal op continues under the next real op
        }
        if (m >= ni) { const uint32_t from = SI > s ? SI : s; if (e > from) emit(rop, e - from); }     // ideal used up: the rest verbatim
      }
    }
    if (__any_sync(0xffffffffu, cnt > 3)) return MERGE_SERIAL;
    // exclusive prefix of counts 0..3: two ballots instead of a shuffle scan
    const uint32_t B0 = __ballot_sync(0xffffffffu, cnt & 1u), B1 = __ballot_sync(0xffffffffu, cnt & 2u);
    const uint32_t off = np + __popc(B0 & lt_mask) + 2 * __popc(B1 & lt_mask);
    if (cnt > 0) pc[off] = p0;
    if (cnt > 1) pc[off + 1] = p1;
    if (cnt > 2) pc[off + 2] = p2;
    np += __popc(B0) + 2 * __popc(B1);
  }
  // ---- real side used up: the rest of the ideal CIGAR verbatim
  {
    const uint32_t SR = S0;
    uint32_t m0 = m_tile;
    while (m0 < ni && done_at(m0, SR)) m0++;
    for (uint32_t base = m0; base < ni; base += 32) {
      const uint32_t m = base + lane;
      uint32_t w = 0;
      if (m < ni) {
        const uint32_t iw = IW[m], iop = iw & 0xfu, il = iw >> 4;
        const uint32_t ic = (iop == CD || iop == OV_D) ? 0u : il, em = EI[m], sm = em - ic;
        const uint32_t len = ic == 0 ? il : em - (sm > SR ? sm : SR);
        if (len) w = (len << 4) | iop;
      }
      const uint32_t V = __ballot_sync(0xffffffffu, w != 0);
      if (w) pc[np + __popc(V & lt_mask)] = w;
      np += __popc(V);
    }
  }
  __syncwarp();
  // ---- run-length coalescing of the pieces into `out`; on the way: is there an I run between two clip runs (step (4))?
  uint32_t n_out = 0, open_w = 0; bool have_open = false, fix = false, open_susp = false;
  for (uint32_t base = 0; base < np; base += 32) {
    const uint32_t i = base + lane;
    const bool v = i < np;
    const uint32_t w = v ? pc[i] : 0u, op = w & 0xfu, len = w >> 4;
    uint32_t prev_op = __shfl_up_sync(0xffffffffu, op, 1);
    if (lane == 0) prev_op = have_open ? (open_w & 0xfu) : 0xffu;
    const bool head = v && op != prev_op;
    const uint32_t H = __ballot_sync(0xffffffffu, head), V = __ballot_sync(0xffffffffu, v);
    const uint32_t L = warp_incl_scan(len, lane);                       // (invalid lanes add 0)
    const int last_v = 31 - __clz(V);
    if (H == 0) { open_w += __shfl_sync(0xffffffffu, L, last_v) << 4; continue; }
    const int first_h = __ffs(H) - 1;
    // the pieces before the first head extend the open run, which is then closed
    const uint32_t before = first_h > 0 ? __shfl_sync(0xffffffffu, L, first_h - 1) : 0u;
    if (have_open && lane == 0) out[n_out - 1] = open_w + (before << 4);
    // every head but the last closes inside this tile
    const uint32_t above = H & ~((2u << lane) - 1u);                    // heads after me
    const int next_h = above ? __ffs(above) - 1 : -1;
    const uint32_t run_end = __shfl_sync(0xffffffffu, L, next_h > 0 ? next_h - 1 : last_v);
    const uint32_t next_op = __shfl_sync(0xffffffffu, op, next_h > 0 ? next_h : 0);
    const uint32_t L_before_me = L - len;
    // an I run whose neighbours are both clip runs: (previous run = prev_op of its head; next run = the next head's op)
    const bool susp = head && op == CI && (prev_op == CS || prev_op == CH);
    if (susp && next_h >= 0 && (next_op == CS || next_op == CH)) fix = true;
    if (open_susp && (int)lane == first_h && (op == CS || op == CH)) fix = true;
    if (head) {
      const uint32_t word = ((run_end - L_before_me) << 4) | op;
      const uint32_t idx = n_out + __popc(H & lt_mask);
      if (next_h >= 0) out[idx] = word;
    }
    const int last_h = 31 - __clz(H);
    const uint32_t last_word = (((__shfl_sync(0xffffffffu, L, last_v) - __shfl_sync(0xffffffffu, L_before_me, last_h)) << 4)) | __shfl_sync(0xffffffffu, op, last_h);
    open_susp = __shfl_sync(0xffffffffu, (uint32_t)susp, last_h) != 0;
    n_out += __popc(H); open_w = last_word; have_open = true;
  }
  if (have_open && lane == 0) out[n_out - 1] = open_w;
  __syncwarp();
  if (!__any_sync(0xffffffffu, fix)) return n_out;
  // ---- (4) an I between clips becomes that clip, (5) re-coalesce (rare: lane 0)
  uint32_t n = 0;
  if (lane == 0) {
    for (uint32_t i = 1; i + 1 < n_out; i++) {
      if ((out[i] & 0xfu) != CI) continue;
      const uint32_t prev = out[i - 1] & 0xfu, next = out[i + 1] & 0xfu;
      if ((prev == CS || prev == CH) && (next == CS || next == CH)) out[i] = ((out[i] >> 4) << 4) | prev;
    }
    for (uint32_t i = 0; i < n_out; i++) {
      const uint32_t op = out[i] & 0xfu, len = out[i] >> 4;
      if (n > 0 && (out[n - 1] & 0xfu) == op) out[n - 1] += (len << 4); else out[n++] = (len << 4) | op;
    }
  }
  __syncwarp();
  return __shfl_sync(0xffffffffu, n, 0);
}

// One merge by the whole warp, whatever its shape (cooperative code, serial code on lane 0 where that is required).
__device__ __forceinline__ uint32_t merge_cigars_any_warp(const uint32_t* __restrict__ real, uint32_t nR, const uint32_t* __restrict__ ideal, uint32_t nI,
                                                          uint32_t iword, uint32_t* __restrict__ out, uint32_t* __restrict__ scr, unsigned lane) {
  uint32_t n = merge_cigars_warp(real, nR, ideal, nI, iword, out, scr, lane);
  if (n == MERGE_SERIAL) {
    __syncwarp();
    if (lane == 0) n = merge_cigars(real, nR, ideal, nI, iword, out);
    n = __shfl_sync(0xffffffffu, n, 0);
  }
  return n;
}

// Standalone entry for the parity tests (brb_merge_cigars).
__global__ void k_merge_pairs(uint32_t n, const uint32_t* real_off, const uint32_t* real, const uint32_t* ideal_off,
                              const uint32_t* ideal, const uint32_t* out_off, uint32_t* out_len, uint32_t* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t ni = ideal_off[i + 1] - ideal_off[i];
  out_len[i] = merge_cigars(real + real_off[i], real_off[i + 1] - real_off[i], ideal + ideal_off[i], ni, ni == 1 ? ideal[ideal_off[i]] : 0,
                            out + out_off[i]);
}

// The same pairs, one warp each, through the cooperative code.
__global__ void k_merge_pairs_warp(uint32_t n, const uint32_t* real_off, const uint32_t* real, const uint32_t* ideal_off,
                                   const uint32_t* ideal, const uint32_t* out_off, const uint32_t* scr_off, uint32_t* out_len, uint32_t* out, uint32_t* scr) {
  const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned lane = threadIdx.x & 31;
  if (i >= n) return;
  const uint32_t ni = ideal_off[i + 1] - ideal_off[i];
  const uint32_t len = merge_cigars_any_warp(real + real_off[i], real_off[i + 1] - real_off[i], ideal + ideal_off[i], ni, ni == 1 ? ideal[ideal_off[i]] : 0,
                                             out + out_off[i], scr + scr_off[i], lane);
  if (lane == 0) out_len[i] = len;
}

}  // namespace brb
