// Development harness (not part of the product or of the tests): the query-indexed formulation of ksw_extz2 that k_ksw.cuh's
// sub-warp kernel implements -- state per QUERY position j (cell (r, t) lives at j = r - t), 16-bit lanes, the target as a
// shift register -- written as plain loops and checked against the restatement of the reference's SSE code
// (oracle/libbrb_port.so: brb_port_ksw_extend, itself pinned on the reference's own translation unit).
//   g++ -O2 -std=c++17 -o /tmp/ksw_proto tools/ksw_proto.cpp oracle/libbrb_port.so -Wl,-rpath,$PWD/oracle && /tmp/ksw_proto
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
#include "../oracle/brb_port.h"

struct Res { int max, score, max_q, max_t, zdropped; std::vector<uint32_t> cigar; };

static Res model(const uint8_t* query, int qlen, const uint8_t* target, int tlen) {
  Res R; R.max = 0; R.score = -0x40000000; R.max_q = R.max_t = -1; R.zdropped = 0;
  if (qlen <= 0 || tlen <= 0) return R;
  const int q = 4, e = 1, qe = q + e, zdrop = 40, max_sc = 1 + 2 * qe;
  std::vector<int16_t> X(qlen, 0), V(qlen), Y(qlen, 0), U(qlen, 0), T(qlen, 4);
  for (int j = 0; j < qlen; j++) V[j] = j ? q : 0;
  const int n_rows = qlen + tlen - 1, pitch = qlen;
  std::vector<uint8_t> pmat((size_t)n_rows * pitch, 0);
  int H0 = 0, last_H0_t = 0, emax = 0, emax_t = -1, emax_q = -1; bool zdropped = false;
  for (int r = 0; r < n_rows; r++) {
    const int st = r - qlen + 1 > 0 ? r - qlen + 1 : 0, en = r < tlen - 1 ? r : tlen - 1;
    // shift registers (from j-1 to j): Y, U of the previous diagonal and the target base
    std::vector<int16_t> Ys(qlen), Us(qlen);
    for (int j = qlen - 1; j >= 1; j--) { Ys[j] = Y[j - 1]; Us[j] = U[j - 1]; T[j] = T[j - 1]; }
    Ys[0] = 0; Us[0] = r ? q : 0; T[0] = r < tlen ? target[r] : 4;
    for (int j = 0; j < qlen; j++) {
      const bool started = j <= r;                       // cell j exists from diagonal r = j on; afterwards it may run past the target (garbage, never read)
      const int sq = T[j], sr = query[j];
      const int s = (sq == 4 || sr == 4) ? -e : (sq == sr ? 1 : -4);
      int z = s + 2 * qe;
      int a = X[j] + V[j], b = Ys[j] + Us[j];
      int d = a > z ? 1 : 0; z = z > a ? z : a;
      d = b > z ? 2 : d; z = z > b ? z : b;
      z = z < max_sc ? z : max_sc;
      const int16_t u2 = (int16_t)(z - V[j]), v2 = (int16_t)(z - Us[j]);
      z -= q; a -= z; b -= z;
      const int16_t x2 = (int16_t)(a > 0 ? a : 0), y2 = (int16_t)(b > 0 ? b : 0);
      if (a > 0) d |= 8; if (b > 0) d |= 16;
      U[j] = u2; Y[j] = y2;                              // U, Y of a cell that has not started are never read
      if (started) { X[j] = x2; V[j] = v2; }
      pmat[(size_t)r * pitch + j] = (uint8_t)d;
    }
    if (r > 0) {
      const bool in0 = last_H0_t >= st && last_H0_t <= en, in1 = last_H0_t + 1 >= st && last_H0_t + 1 <= en;
      if (in0 && in1) {
        const int d0 = V[r - last_H0_t] - qe, d1 = U[r - last_H0_t - 1] - qe;
        if (d0 > d1) H0 += d0; else { H0 += d1; ++last_H0_t; }
      } else if (in0) H0 += V[r - last_H0_t] - qe;
      else { ++last_H0_t; H0 += U[r - last_H0_t] - qe; }
      const int t = last_H0_t;
      if (H0 > emax) { emax = H0; emax_t = t; emax_q = r - t; }
      else if (t >= emax_t && r - t >= emax_q) {
        const int tl = t - emax_t, ql = (r - t) - emax_q, l = tl > ql ? tl - ql : ql - tl;
        if (emax - H0 > zdrop + l * e) { zdropped = true; break; }
      }
    } else { H0 = V[0] - qe - qe; last_H0_t = 0; }
    if (r == qlen + tlen - 2 && en == tlen - 1) R.score = H0;
  }
  R.max = emax; R.max_t = emax_t; R.max_q = emax_q; R.zdropped = zdropped;
  if (emax_t >= 0 && emax_q >= 0) {
    std::vector<uint32_t> tb;
    auto push = [&](uint32_t op, int len) { if (tb.empty() || op != (tb.back() & 0xf)) tb.push_back((uint32_t)len << 4 | op); else tb.back() += (uint32_t)len << 4; };
    int i = emax_t, j = emax_q, state = 0;
    while (i >= 0 && j >= 0) {
      const uint32_t tmp = pmat[(size_t)(i + j) * pitch + j];
      if (state == 0) state = tmp & 7; else if (!((tmp >> (state + 2)) & 1)) state = 0;
      if (state == 0) state = tmp & 7;
      if (state == 0) { push(0, 1); --i; --j; } else if (state == 1 || state == 3) { push(2, 1); --i; } else { push(1, 1); --j; }
    }
    if (i >= 0) push(2, i + 1);
    if (j >= 0) push(1, j + 1);
    R.cigar.assign(tb.rbegin(), tb.rend());
  }
  return R;
}

int main(int argc, char** argv) {
  const long N = argc > 1 ? atol(argv[1]) : 20000;
  std::mt19937_64 rng(777);
  long bad = 0, nz = 0, nend = 0;
  for (long k = 0; k < N; k++) {
    const int ql = 1 + (int)(rng() % (k % 10 ? 120 : 500));
    std::vector<uint8_t> q(ql), t;
    for (auto& c : q) c = rng() % 4;
    if (k % 4 == 0) { t.resize(1 + rng() % (ql + 40)); for (auto& c : t) c = rng() % 4; }
    else {
      t.assign(q.begin(), q.end());
      const int ne = (int)(rng() % (ql / 12 + 2));
      for (int x = 0; x < ne; x++) {
        const size_t pos = rng() % t.size(); const int r = (int)(rng() % 10);
        if (r < 4) t[pos] = rng() % 5; else if (r < 7) { if (t.size() > 1) t.erase(t.begin() + pos); } else t.insert(t.begin() + pos, (uint8_t)(rng() % 4));
      }
      const int tail = (int)(rng() % 41);
      for (int x = 0; x < tail; x++) t.push_back(rng() % 4);
      if ((int)t.size() > ql + 40) t.resize(ql + 40);
    }
    const uint64_t qo[2] = {0, (uint64_t)ql}, to[2] = {0, (uint64_t)t.size()};
    int32_t mx, sc, mq, mt; uint8_t zd; uint32_t nc; std::vector<uint32_t> cg(4096);
    brb_port_ksw_extend(1, qo, q.data(), to, t.data(), &mx, &sc, &mq, &mt, &zd, 4096, &nc, cg.data());
    cg.resize(nc);
    const Res R = model(q.data(), ql, t.data(), (int)t.size());
    nz += zd; nend += sc != -0x40000000;
    if (R.max != mx || R.score != sc || R.max_q != mq || R.max_t != mt || R.zdropped != zd || R.cigar != cg) {
      if (bad < 5) printf("MISMATCH %ld ql=%d tl=%zu: max %d/%d score %d/%d mq %d/%d mt %d/%d zd %d/%d ncig %zu/%u\n", k, ql, t.size(), R.max, mx, R.score, sc, R.max_q, mq, R.max_t, mt, R.zdropped, zd, R.cigar.size(), nc);
      bad++;
    }
  }
  printf("%ld problems (%ld z-dropped, %ld reached the end), %ld mismatches\n", N, nz, nend, bad);
  return bad != 0;
}
