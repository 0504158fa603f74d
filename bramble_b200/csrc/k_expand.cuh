// Kernel group 8: compact input columns expanded on the device (brb_batch_in::pos0 / cigar_cnt16 / mate_delta16 / refid16).
// The host ships fewer bytes per alignment; what the kernels of the path read (exon CSR, CIGAR offsets, mate index, refid) is
// rebuilt here once per upload:
// in ONE launch (k_expand_all: single-pass scans with decoupled look-back), so that an upload is a handful of copies and one kernel with
// no host round trip in between -- on the end-to-end path several contexts share the GPU, and every extra launch of one context waits behind
// the GPU-filling kernels of the others:
//   16-bit CIGAR word counts -> cigar_off (scan);  mate_delta16 -> mate_idx, refid16 -> refid;
//   pos0 + CIGAR -> number of read exons per alignment -> exon_off (second scan) -> the exons themselves
// Read exons from pos0 + CIGAR: the reference's ingest (GSamRecord::setupCoordinates, gclib/GSam.cpp:197-291, and the `end++` of
// process_exons, src/bramble.cpp:246-255), as the host half of the drop-in restates it (host/bamrec.hpp read_exons): an exon is closed
// at an N that follows aligned bases (unless an I sits between two Ns), the last one at the end; coordinates 1-based, half-open.
#pragma once
#include "dev_common.cuh"

namespace brb {

// The exons of one alignment: calls emit(start, end) per exon, returns whether the CIGAR is valid (the reference's check: a last
// exon end of 0 is "invalid CIGAR").
template <typename Emit>
__device__ __forceinline__ bool walk_read_exons(int32_t pos0, const uint32_t* __restrict__ cg, uint32_t n, Emit&& emit) {
  uint32_t l = 0; int32_t exstart = pos0; uint32_t ex_e = 0;
  bool started = false, intron = false, ins = false;
  for (uint32_t i = 0; i < n; i++) {
    const uint32_t w = __ldg(&cg[i]), op = w & 0xfu, len = w >> 4;
    switch (op) {
      case 0: case 7: case 8: started = true; l += len; intron = false; ins = false; break;
      case 2: l += len; ins = false; break;
      case 1: ins = true; break;
      case 3:
        if (!started) break;
        if (!ins || !intron) { ex_e = (uint32_t)(pos0 + (int32_t)l); emit((uint32_t)(exstart + 1), ex_e + 1); }
        l += len; exstart = pos0 + (int32_t)l; intron = true; break;
      case 4: case 5: ins = false; break;
      default: break;
    }
  }
  if (!intron) { ex_e = (uint32_t)(pos0 + (int32_t)l); emit((uint32_t)(exstart + 1), ex_e + 1); }
  return ex_e != 0;
}

struct ExpandJob {
  uint32_t n;
  const uint16_t* cnt16; uint32_t n_cigar;           // CIGAR words per alignment (or null: cigar_off came from the host) and the size of the word array
  uint32_t* cigar_off; const uint32_t* cigar;
  const int32_t* pos0;                                 // or null: the exon columns came from the host
  uint32_t* exon_off; uint32_t* exon_st; uint32_t* exon_en; uint32_t exon_cap;
  const int16_t* mate_delta16; int32_t* mate_idx; const int16_t* refid16; int32_t* refid;
  unsigned long long* look;                            // 2 * tiles look-back words (zeroed), then the tile ticket
  uint32_t* res;                                       // [0] first invalid alignment + 1, [1] read exons, [2] CIGAR words (sum of the counts), [3] counts that ran past the word array
};

// Single-pass scan bookkeeping (decoupled look-back): word = flag << 62 | value, flag 1 = the tile's own sum, 2 = inclusive prefix.
__device__ __forceinline__ unsigned long long lookback(volatile unsigned long long* st, uint32_t t, unsigned long long own) {
  unsigned long long prefix = 0;
  if (t) {
    st[t] = (1ull << 62) | own; __threadfence();
    for (uint32_t j = t; j-- > 0;) {
      unsigned long long w;
      do { w = st[j]; } while ((w >> 62) == 0);
      prefix += w & ((1ull << 62) - 1);
      if ((w >> 62) == 2) break;
    }
  }
  st[t] = (2ull << 62) | (prefix + own); __threadfence();
  return prefix;
}

// ONE launch per upload: tiles of 256 alignments taken by ticket (a tile only ever waits for tiles before it, which are running);
// per tile: CIGAR offsets (scan of the 16-bit counts), 32-bit links, exon counts from pos0 + CIGAR, exon offsets (second scan), exons.
__global__ void __launch_bounds__(256) k_expand_all(ExpandJob J) {
  __shared__ uint32_t s_warp[8]; __shared__ unsigned long long s_pref; __shared__ uint32_t s_tile;
  const uint32_t tiles = (J.n + 255) / 256;
  volatile unsigned long long* look_c = J.look; volatile unsigned long long* look_e = J.look + tiles;
  uint32_t* ticket = reinterpret_cast<uint32_t*>(J.look + 2 * (size_t)tiles);
  const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  auto block_scan = [&](uint32_t v, uint32_t& excl, uint32_t& total) {         // exclusive scan over the 256 threads
    const uint32_t incl = warp_incl_scan(v, lane);
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    uint32_t base = 0, tot = 0;
    for (unsigned w = 0; w < 8; w++) { const uint32_t x = s_warp[w]; if (w < wid) base += x; tot += x; }
    __syncthreads();
    excl = base + incl - v; total = tot;
  };
  for (;;) {
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t t = s_tile;
    if (t >= tiles) break;
    const uint32_t a = t * 256 + threadIdx.x;
    const bool in = a < J.n;
    if (in) {
      if (J.mate_delta16) { const int32_t d = J.mate_delta16[a]; J.mate_idx[a] = d ? (int32_t)a + d : -1; }
      if (J.refid16) J.refid[a] = J.refid16[a];
    }
    // ---- CIGAR offsets
    uint32_t c0 = 0, nc = 0;
    if (J.cnt16) {
      nc = in ? J.cnt16[a] : 0u;
      uint32_t excl, total; block_scan(nc, excl, total);
      if (threadIdx.x == 0) s_pref = lookback(look_c, t, total);
      __syncthreads();
      const unsigned long long start = s_pref + excl;
      if (start + nc > J.n_cigar) { if (nc) atomicAdd(&J.res[3], 1u); nc = 0; }                 // the counts say more words than were handed over: nothing may be read there
      c0 = (uint32_t)(start < 0xffffffffull ? start : 0xffffffffull);
      if (in) J.cigar_off[a] = c0;
      if (t == tiles - 1 && threadIdx.x == 255) { const unsigned long long tot = s_pref + total; J.cigar_off[J.n] = (uint32_t)(tot < 0xffffffffull ? tot : 0xffffffffull); J.res[2] = J.cigar_off[J.n]; }
      __syncthreads();
    } else if (in) {
      c0 = J.cigar_off[a]; const uint32_t c1 = J.cigar_off[a + 1];
      nc = (c0 <= c1 && c1 <= J.n_cigar) ? c1 - c0 : 0u;       // (offsets from the host have not been through k_validate_batch yet)
      if (a == J.n - 1) J.res[2] = c1;
    }
    // ---- read exons from pos0 + CIGAR
    if (J.pos0) {
      uint32_t k = 0; bool ok = true;
      if (in) ok = walk_read_exons(J.pos0[a], J.cigar + c0, nc, [&](uint32_t, uint32_t) { k++; });
      if (in && !ok) atomicMax(&J.res[0], a + 1u);
      uint32_t excl, total; block_scan(k, excl, total);
      if (threadIdx.x == 0) s_pref = lookback(look_e, t, total);
      __syncthreads();
      uint32_t e0 = (uint32_t)(s_pref + excl);
      if (in) {
        J.exon_off[a] = e0;
        if ((unsigned long long)e0 + k <= J.exon_cap) walk_read_exons(J.pos0[a], J.cigar + c0, nc, [&](uint32_t s, uint32_t e) { J.exon_st[e0] = s; J.exon_en[e0] = e; e0++; });
      }
      if (t == tiles - 1 && threadIdx.x == 255) { const uint32_t tot = (uint32_t)(s_pref + total); J.exon_off[J.n] = tot; J.res[1] = tot; }
      __syncthreads();
    }
  }
}

}  // namespace brb
