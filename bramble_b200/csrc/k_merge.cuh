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

// Standalone entry for the parity tests (brb_merge_cigars).
__global__ void k_merge_pairs(uint32_t n, const uint32_t* real_off, const uint32_t* real, const uint32_t* ideal_off,
                              const uint32_t* ideal, const uint32_t* out_off, uint32_t* out_len, uint32_t* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t ni = ideal_off[i + 1] - ideal_off[i];
  out_len[i] = merge_cigars(real + real_off[i], real_off[i + 1] - real_off[i], ideal + ideal_off[i], ni, ni == 1 ? ideal[ideal_off[i]] : 0,
                            out + out_off[i]);
}

}  // namespace brb
