// Development harness (not part of the product or of the tests): the data-parallel formulation of merge_cigars that
// k_merge.cuh's warp-cooperative version implements -- every real CIGAR op derives its output pieces from prefix sums
// alone, then one run-length coalescing pass -- written as plain loops and checked against the serial restatement
// (oracle/libbrb_port.so: brb_port_merge_cigars, reference src/bam.cpp:113-315) on random CIGAR pairs.
//   g++ -O2 -std=c++17 -o /tmp/merge_proto tools/merge_proto.cpp oracle/libbrb_port.so -Wl,-rpath,$PWD/oracle && /tmp/merge_proto
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
#include "../oracle/brb_port.h"

enum { CM = 0, CI = 1, CD = 2, CN = 3, CS = 4, CH = 5, CP = 6, CEQ = 7, CX = 8, OV_M = 10, OV_D = 11, OV_I = 12, OV_S = 13, SKIP = 0x5f };

static uint32_t merge_ops(uint32_t r, uint32_t i) {
  const bool rMS = (r == CM || r == CS);
  if (rMS && i == OV_S) return CS;
  if (rMS && i == OV_M) return CM;
  if (rMS && i == OV_I) return CI;
  if (rMS && i == OV_D) return CD;
  if (r == CD && (i == CS || i == OV_S)) return SKIP;
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

struct Piece { uint32_t op, len; };

static void parallel_model(const uint32_t* real, uint32_t nR, const uint32_t* ideal, uint32_t nI, std::vector<uint32_t>& out) {
  std::vector<Piece> pieces;
  auto add = [&](uint32_t op, uint32_t len) { pieces.push_back({op, len}); };
  // ---- prologue: phases (1) and (2) of the serial algorithm, unchanged (a handful of iterations)
  uint32_t front_hard = 0, front_soft = 0, ci = 0;
  if (nR > 0 && (real[0] & 0xf) == CH) { front_hard = real[0] >> 4; ci++; }
  if (ci < nR && (real[ci] & 0xf) == CS) front_soft = real[ci] >> 4;
  uint32_t ri = 0, ii = 0, rpos = 0, ipos = 0, clips = front_hard;
  while (clips > 0 && ri < nR) {
    const uint32_t rw = real[ri], avail = (rw >> 4) - rpos, chunk = clips < avail ? clips : avail;
    add(rw & 0xf, chunk); clips -= chunk; rpos += chunk;
    if (rpos >= (rw >> 4)) { ri++; rpos = 0; }
  }
  clips = front_soft;
  while (clips > 0 && ri < nR) {
    const uint32_t rw = real[ri], rop = rw & 0xf, rrem = (rw >> 4) - rpos;
    const bool have_i = ii < nI;
    const uint32_t iw = have_i ? ideal[ii] : 0, iop = have_i ? (iw & 0xf) : 0xff, irem = have_i ? (iw >> 4) - ipos : 0xffffffffu;
    const bool ov = have_i && (iop == OV_M || iop == OV_D || iop == OV_I || iop == OV_S);
    if (ov) {
      if (iop == OV_D) { add(merge_ops(rop, iop), irem); ipos += irem; if (ipos >= (iw >> 4)) { ii++; ipos = 0; } }
      else {
        uint32_t chunk = clips; if (chunk > rrem) chunk = rrem; if (chunk > irem) chunk = irem;
        add(merge_ops(rop, iop), chunk); clips -= chunk; rpos += chunk; ipos += chunk;
        if (rpos >= (rw >> 4)) { ri++; rpos = 0; }
        if (ipos >= (iw >> 4)) { ii++; ipos = 0; }
      }
    } else {
      uint32_t chunk = clips; if (chunk > rrem) chunk = rrem;
      add(merge_ops(rop, iop), chunk); clips -= chunk; rpos += chunk;
      if (rpos >= (rw >> 4)) { ri++; rpos = 0; }
    }
  }
  // ---- main loop (3) as a data-parallel computation.  Cursor state: (ri, rpos, ii, ipos).
  const uint32_t nr = nR - ri, ni = nI - ii;
  auto rop_ = [&](uint32_t k) { return real[ri + k] & 0xf; };
  auto rlen_ = [&](uint32_t k) { return (real[ri + k] >> 4) - (k == 0 ? rpos : 0); };
  auto iop_ = [&](uint32_t m) { return ideal[ii + m] & 0xf; };
  auto ilen_ = [&](uint32_t m) { return (ideal[ii + m] >> 4) - (m == 0 ? ipos : 0); };
  auto icons = [&](uint32_t m) { const uint32_t op = iop_(m); return (op == CD || op == OV_D) ? 0u : ilen_(m); };
  std::vector<uint64_t> EI(ni), ER(nr);                 // inclusive prefix sums of the consumed lengths
  uint64_t acc = 0;
  for (uint32_t m = 0; m < ni; m++) { acc += icons(m); EI[m] = acc; }
  const uint64_t SI = acc;
  acc = 0;
  for (uint32_t k = 0; k < nr; k++) { const uint32_t op = rop_(k); acc += (op == CN || op == CI) ? 0 : rlen_(k); ER[k] = acc; }
  const uint64_t SR = acc;
  // ideal cursor when the real side stands at coordinate x with a real op at its front: everything that ends before x, and a
  // positive-length consumer that ends exactly at x, is done; zero-consumers located at x are still pending
  auto ii_at = [&](uint64_t x) {
    uint32_t lo = 0, hi = ni;
    while (lo < hi) { const uint32_t mid = (lo + hi) / 2; if (EI[mid] < x) lo = mid + 1; else hi = mid; }
    if (lo < ni && EI[lo] == x && icons(lo) > 0) lo++;
    return lo;
  };
  std::vector<std::vector<Piece>> per_op(nr);
  for (uint32_t k = 0; k < nr; k++) {                  // independent per real op ("lane = real op")
    auto& P = per_op[k];
    const uint32_t rop = rop_(k), rlen = rlen_(k);
    const uint64_t e = ER[k], s = e - ((rop == CN || rop == CI) ? 0 : rlen);
    if (rop == CN || rop == CI) {
      const bool overlay = ii_at(s) < ni;
      if (rop == CI) P.push_back({CI, rlen});
      else if (!overlay) P.push_back({CN, rlen});
      continue;
    }
    uint32_t m = ii_at(s);
    uint64_t cur = s;                                   // real coordinate consumed so far inside this op
    for (; m < ni; m++) {
      const uint32_t iop = iop_(m), il = ilen_(m), ic = icons(m);
      const uint64_t em = EI[m], sm = em - ic;
      if (ic == 0) {                                    // zero-consumer at coordinate em
        if (em >= e && !(e == s && em == s)) break;     // belongs to a later real op (an empty real op handles nothing at all -> see below)
        if (e == s) break;                              // zero-length consuming real op: advances nothing, emits nothing
        if ((iop == CD || iop == OV_D) && il > 0) P.push_back({CD, il});
        continue;
      }
      if (e == s) break;
      if (sm >= e) break;
      const uint64_t a = sm > cur ? sm : cur, b = em < e ? em : e;
      const uint32_t len = (uint32_t)(b - a);
      const bool swallow = rop == CD && (iop == CS || iop == OV_S || iop == CI || iop == OV_I);
      if (!swallow) P.push_back({merge_ops(rop, iop), len});
      cur = b;
      if (em > e) break;                                // the ideal op continues under the next real op
    }
    if (m >= ni && e > s) {                             // ideal exhausted inside (or before) this op: the rest of it verbatim
      const uint64_t from = SI > s ? SI : s;
      if (e > from) P.push_back({rop, (uint32_t)(e - from)});
    }
  }
  for (auto& P : per_op) for (auto& p : P) pieces.push_back(p);
  // real exhausted: what is left of the ideal CIGAR, verbatim
  {
    uint32_t m = ii_at(SR);
    // (a zero-length consuming real op at the very end does not change the cursor either)
    for (; m < ni; m++) {
      const uint32_t ic = icons(m), il = ilen_(m);
      if (ic == 0) { pieces.push_back({iop_(m), il}); continue; }
      const uint64_t em = EI[m], sm = em - ic, from = sm > SR ? sm : SR;
      pieces.push_back({iop_(m), (uint32_t)(em - from)});
    }
  }
  // ---- coalesce (drop empty pieces and SKIPs, merge equal neighbours), then steps (4) and (5)
  out.clear();
  for (auto& p : pieces) {
    if (p.len == 0 || p.op == SKIP) continue;
    if (!out.empty() && (out.back() & 0xf) == p.op) out.back() += p.len << 4; else out.push_back(p.len << 4 | p.op);
  }
  for (size_t i = 1; i + 1 < out.size(); i++) {
    if ((out[i] & 0xf) != CI) continue;
    const uint32_t prev = out[i - 1] & 0xf, next = out[i + 1] & 0xf;
    if ((prev == CS || prev == CH) && (next == CS || next == CH)) out[i] = (out[i] >> 4) << 4 | prev;
  }
  size_t n = 0;
  for (size_t i = 0; i < out.size(); i++) {
    const uint32_t op = out[i] & 0xf, len = out[i] >> 4;
    if (n > 0 && (out[n - 1] & 0xf) == op) out[n - 1] += len << 4; else out[n++] = len << 4 | op;
  }
  out.resize(n);
}

int main(int argc, char** argv) {
  const long N = argc > 1 ? atol(argv[1]) : 2000000;
  std::mt19937_64 rng(12345);
  const uint32_t ops_r[] = {0, 1, 2, 3, 4, 5, 7, 8, 0, 0, 1, 2, 6}, ops_i[] = {0, 1, 2, 4, 10, 11, 12, 13, 0, 0, 2, 7, 8};
  long bad = 0, n_fallback = 0;
  for (long t = 0; t < N; t++) {
    const int mode = (int)(rng() % 4);
    const uint32_t maxr = mode == 0 ? 5 : mode == 1 ? 12 : mode == 2 ? 60 : 300, maxi = mode == 3 ? 40 : 10;
    const uint32_t nR = (uint32_t)(rng() % (maxr + 1)), nI = (uint32_t)(rng() % (maxi + 1));
    std::vector<uint32_t> real(nR), ideal(nI);
    const bool zeros = rng() % 8 == 0, structured = rng() % 2 == 0;
    for (uint32_t k = 0; k < nR; k++) {
      uint32_t op = ops_r[rng() % 13], len = (uint32_t)(rng() % 40) + 1;
      if (structured) { if (k == 0 && rng() % 3 == 0) op = CH; else if (k <= 1 && rng() % 2 == 0) op = CS; else if (k + 1 == nR && rng() % 3 == 0) op = CS; }
      if (zeros && rng() % 6 == 0) len = 0;
      real[k] = len << 4 | op;
    }
    for (uint32_t m = 0; m < nI; m++) {
      uint32_t op = ops_i[rng() % 13], len = (uint32_t)(rng() % 60) + 1;
      if (structured && m < 3 && rng() % 2 == 0) op = 10 + (uint32_t)(rng() % 4);
      if (zeros && rng() % 6 == 0) len = 0;
      ideal[m] = len << 4 | op;
    }
    if (rng() % 3 == 0 && nR && nI) {      // make the two sides consume (nearly) the same length, as real data does
      uint64_t sr = 0, si = 0;
      for (auto w : real) if ((w & 0xf) != CN && (w & 0xf) != CI) sr += w >> 4;
      for (auto w : ideal) if ((w & 0xf) != CD && (w & 0xf) != OV_D) si += w >> 4;
      if (sr > si && (ideal[nI - 1] & 0xf) != CD && (ideal[nI - 1] & 0xf) != OV_D) ideal[nI - 1] += (uint32_t)((sr - si + (rng() % 3) - 1) << 4);
    }
    std::vector<uint32_t> want(nR + nI + 8), got;
    uint32_t roff[2] = {0, nR}, ioff[2] = {0, nI}, ooff[2] = {0, nR + nI + 8}, olen = 0;
    static const uint32_t dummy = 0;
    brb_port_merge_cigars(1, roff, nR ? real.data() : &dummy, ioff, nI ? ideal.data() : &dummy, ooff, &olen, want.data());
    want.resize(olen);
    parallel_model(real.data(), nR, ideal.data(), nI, got);
    bool real_zero = false;
    for (auto w : real) if ((w >> 4) == 0) real_zero = true;
    if (real_zero) { n_fallback++; continue; }          // the cooperative kernel hands these to the serial code (a zero-length real op is not valid BAM)
    if (got != want) {
      if (bad < 5) {
        printf("MISMATCH case %ld\n real :", t); for (auto w : real) printf(" %u%c", w >> 4, "MIDNSHP=X?mdis??"[w & 0xf]);
        printf("\n ideal:"); for (auto w : ideal) printf(" %u%c", w >> 4, "MIDNSHP=X?mdis??"[w & 0xf]);
        printf("\n want :"); for (auto w : want) printf(" %u%c", w >> 4, "MIDNSHP=X?mdis??"[w & 0xf]);
        printf("\n got  :"); for (auto w : got) printf(" %u%c", w >> 4, "MIDNSHP=X?mdis??"[w & 0xf]);
        printf("\n");
      }
      bad++;
    }
  }
  printf("%ld cases (%ld with a zero-length real op left to the serial code), %ld mismatches\n", N, n_fallback, bad);
  return bad != 0;
}
