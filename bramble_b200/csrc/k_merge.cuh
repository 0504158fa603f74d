// Kernel group 4: CIGAR projection -- merge the alignment's real CIGAR with the ideal (transcript-side)
// CIGAR of the match.  Replaces merge_ops / merge_cigars / get_new_cigar (reference src/bam.cpp:22-315,
// 443-472) and the CIGAR reversal of reverse_complement_bam (:636-702).  One thread per output row.
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

struct CigOut {
  uint32_t* p; uint32_t n;
  __device__ __forceinline__ void add(uint32_t op, uint32_t len) {
    if (len == 0 || op == OP_SKIP) return;
    if (n > 0 && (p[n - 1] & 0xfu) == op) p[n - 1] += (len << 4);
    else p[n++] = (len << 4) | op;
  }
};

// Returns the number of words written to `out` (capacity >= n_real + n_ideal + 2).
__device__ uint32_t merge_cigars(const uint32_t* __restrict__ real, uint32_t n_real, const uint32_t* __restrict__ ideal,
                                 uint32_t n_ideal, uint32_t* __restrict__ out) {
  CigOut res; res.p = out; res.n = 0;
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
    const uint32_t iw = have_i ? ideal[ii] : 0, iop = have_i ? (iw & 0xfu) : 0xffu, irem = have_i ? (iw >> 4) - ipos : 0xffffffffu;
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
    if (ri >= n_real) { const uint32_t iw = ideal[ii]; res.add(iw & 0xfu, (iw >> 4) - ipos); ii++; ipos = 0; continue; }
    if (ii >= n_ideal) { const uint32_t rw = real[ri]; res.add(rw & 0xfu, (rw >> 4) - rpos); ri++; rpos = 0; continue; }
    const uint32_t rw = real[ri], iw = ideal[ii], rop = rw & 0xfu, iop = iw & 0xfu;
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

// One thread per output row: merged CIGAR into a bump-allocated scratch pool (capacity n_real + n_ideal + 2
// words per row, one atomic per warp), its exact length into o_clen.
__global__ void __launch_bounds__(256) k_merge(DevBatch b, Ctl* __restrict__ ctl, const uint32_t* __restrict__ o_src,
                                               const uint32_t* __restrict__ o_cand, const uint32_t* __restrict__ m_ioff,
                                               const uint32_t* __restrict__ m_ilen, const uint32_t* __restrict__ ipool,
                                               uint32_t* __restrict__ o_tmpoff, uint32_t* __restrict__ o_clen, uint32_t* __restrict__ tpool) {
  if (ctl_overflow(ctl)) return;
  const unsigned lane = threadIdx.x & 31;
  const uint32_t n_rec = ctl->n_rows;
  const uint32_t stride = gridDim.x * blockDim.x;
  WarpPool wp; wp.cur = 0; wp.end = 0;
  const uint32_t cap_tmpw = ctl->cap_tmpw;
  for (uint32_t base = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); base < n_rec; base += stride) {
    const uint32_t r = base + lane;
    const bool act = r < n_rec;
    uint32_t ro = 0, nr = 0, c = 0, ni = 0;
    if (act) { const uint32_t a = o_src[r]; c = o_cand[r]; ro = b.cigar_off[a]; nr = b.cigar_off[a + 1] - ro; ni = m_ilen[c]; }
    bool fits;
    const uint32_t off = warp_pool_alloc(wp, &ctl->n_tmpw, cap_tmpw, act ? nr + ni + 2 : 0, lane, fits);
    if (act && fits) {
      o_tmpoff[r] = off;
      o_clen[r] = merge_cigars(b.cigar + ro, nr, ipool + m_ioff[c], ni, tpool + off);
    }
  }
}

// Dense, row-ordered CIGAR pool: row r's words go to opool[o_coff[r] .. o_coff[r+1]) (o_coff = exclusive scan of
// o_clen), reversed for '-' transcripts (reverse_complement_bam, src/bam.cpp:636-702).
__global__ void __launch_bounds__(256) k_compact(Ctl* __restrict__ ctl, const int8_t* __restrict__ o_strand,
                                                 const uint32_t* __restrict__ o_tmpoff, const uint32_t* __restrict__ o_coff,
                                                 const uint32_t* __restrict__ tpool, uint32_t* __restrict__ opool) {
  if (ctl_overflow(ctl)) return;
  const uint32_t n_rec = ctl->n_rows;
  const uint32_t W = o_coff[n_rec];
  if (blockIdx.x == 0 && threadIdx.x == 0) { ctl->n_outw = W; ctl->out_words = W; }
  if (W > ctl->cap_outw) return;
  for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rec; r += gridDim.x * blockDim.x) {
    const uint32_t d0 = o_coff[r], n = o_coff[r + 1] - d0;
    const uint32_t* src = tpool + o_tmpoff[r];
    if (o_strand[r] == '-') for (uint32_t i = 0; i < n; i++) opool[d0 + i] = src[n - 1 - i];
    else for (uint32_t i = 0; i < n; i++) opool[d0 + i] = src[i];
  }
}

// Standalone entry for the parity tests (brb_merge_cigars).
__global__ void k_merge_pairs(uint32_t n, const uint32_t* real_off, const uint32_t* real, const uint32_t* ideal_off,
                              const uint32_t* ideal, const uint32_t* out_off, uint32_t* out_len, uint32_t* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out_len[i] = merge_cigars(real + real_off[i], real_off[i + 1] - real_off[i], ideal + ideal_off[i], ideal_off[i + 1] - ideal_off[i],
                            out + out_off[i]);
}

}  // namespace brb
