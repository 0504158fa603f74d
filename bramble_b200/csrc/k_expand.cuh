// Kernel group 8: compact input columns expanded on the device (brb_batch_in::pos0 / cigar_cnt16 / mate_delta16 / refid16).
// The host ships fewer bytes per alignment; what the kernels of the path read (exon CSR, CIGAR offsets, mate index, refid) is
// rebuilt here once per upload:
//   k_expand_counts   16-bit CIGAR word counts -> 32-bit (input of the offset scan)
//   k_expand_links    mate_delta16 -> mate_idx, refid16 -> refid
//   k_exon_count      thread per alignment: number of read exons its CIGAR yields
//   k_exon_fill       ... and the exons themselves, behind the scan of the counts
// Read exons from pos0 + CIGAR: the reference's ingest (GSamRecord::setupCoordinates, gclib/GSam.cpp:197-291, and the `end++` of
// process_exons, src/bramble.cpp:246-255), as the host half of the drop-in restates it (host/bamrec.hpp read_exons): an exon is closed
// at an N that follows aligned bases (unless an I sits between two Ns), the last one at the end; coordinates 1-based, half-open.
#pragma once
#include "dev_common.cuh"

namespace brb {

__global__ void __launch_bounds__(256) k_expand_counts(uint32_t n, const uint16_t* __restrict__ cnt16, uint32_t* __restrict__ cnt32) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a < n) cnt32[a] = cnt16[a];
}

__global__ void __launch_bounds__(256) k_expand_links(uint32_t n, const int16_t* __restrict__ mate_delta16, int32_t* __restrict__ mate_idx,
                                                      const int16_t* __restrict__ refid16, int32_t* __restrict__ refid) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= n) return;
  if (mate_delta16) { const int32_t d = mate_delta16[a]; mate_idx[a] = d ? (int32_t)a + d : -1; }
  if (refid16) refid[a] = refid16[a];
}

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

__global__ void __launch_bounds__(256) k_exon_count(uint32_t n, const int32_t* __restrict__ pos0, const uint32_t* __restrict__ cigar_off,
                                                    const uint32_t* __restrict__ cigar, uint32_t n_cigar, uint32_t* __restrict__ cnt, uint32_t* __restrict__ err) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= n) return;
  uint32_t k = 0;
  const uint32_t c0 = cigar_off[a], c1 = cigar_off[a + 1];
  const bool sane = c0 <= c1 && c1 <= n_cigar;              // (offsets that came from the host have not been through k_validate_batch yet)
  const bool ok = sane && walk_read_exons(pos0[a], cigar + c0, c1 - c0, [&](uint32_t, uint32_t) { k++; });
  cnt[a] = k;
  if (!ok) atomicMax(err, a + 1u);                          // (an) alignment whose CIGAR yields no exon
}

__global__ void __launch_bounds__(256) k_exon_fill(uint32_t n, const int32_t* __restrict__ pos0, const uint32_t* __restrict__ cigar_off,
                                                   const uint32_t* __restrict__ cigar, const uint32_t* __restrict__ exon_off,
                                                   uint32_t* __restrict__ exon_st, uint32_t* __restrict__ exon_en) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= n) return;
  uint32_t k = exon_off[a];
  const uint32_t c0 = cigar_off[a];
  walk_read_exons(pos0[a], cigar + c0, cigar_off[a + 1] - c0, [&](uint32_t s, uint32_t e) { exon_st[k] = s; exon_en[k] = e; k++; });
}

}  // namespace brb
