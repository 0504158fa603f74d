// Kernel group 3: mate pairing and record emission (one thread per alignment "unit").
// Replaces process_mate_pair / add_mate_info (reference src/mates.cpp:28-261) and the group loop of
// convert_reads (src/core.cpp:384-415).  process_pairs (src/bramble.cpp:272-311) links every
// alignment to at most one mate and the link is symmetric, so inside a name group the units
// {i unpaired} and {i < mate(i)} are independent: the lower index is `this_read` (read1 of the
// BamInfo), the higher one the mate.  Rows of a unit come out ascending by tid (sorted tid lists +
// set_intersection in the reference), the read row immediately followed by its mate row.
#pragma once
#include "dev_common.cuh"

namespace brb {

struct PairView {
  const uint32_t* t_coff; const uint8_t* m_valid; const uint32_t* m_tid;
};

__device__ __forceinline__ uint32_t count_valid(const PairView& v, uint32_t a, uint32_t& only) {
  uint32_t n = 0;
  for (uint32_t c = v.t_coff[2 * a], e = v.t_coff[2 * a + 2]; c < e; c++) if (v.m_valid[c]) { n++; only = c; }
  return n;
}

// Pass A: rows per unit.  Also records for every valid match of the read its partner match on the
// mate (c_partner; NONE = not emitted), which pass B uses to rank the rows.
constexpr uint32_t NONE = 0xffffffffu;
constexpr uint32_t SELF = 0xfffffffeu;   // emitted without a mate row

__global__ void __launch_bounds__(256) k_pair_count(DevBatch b, PairView v, uint32_t* __restrict__ c_partner,
                                                    uint32_t* __restrict__ u_nrows, uint8_t* __restrict__ u_case) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= b.n_aln) return;
  const int32_t mate = b.mate_idx[a];
  uint32_t rows = 0; uint8_t ucase = 255;
  if (mate < 0 || (uint32_t)mate > a) {
    uint32_t r_only = 0, m_only = 0;
    const uint32_t nr = count_valid(v, a, r_only);
    const uint32_t nm = (mate >= 0 && (uint32_t)mate < b.n_aln) ? count_valid(v, (uint32_t)mate, m_only) : 0;
    if (nr > 0) {
      if (nm == 0) {                                   // case 0: mate absent or without a match (src/mates.cpp:150-176)
        ucase = 0;
        for (uint32_t c = v.t_coff[2 * a], e = v.t_coff[2 * a + 2]; c < e; c++) if (v.m_valid[c]) { c_partner[c] = SELF; rows++; }
      } else {
        uint32_t common = 0;
        const uint32_t mb = v.t_coff[2 * (uint32_t)mate], me = v.t_coff[2 * (uint32_t)mate + 2];
        for (uint32_t c = v.t_coff[2 * a], e = v.t_coff[2 * a + 2]; c < e; c++) {
          if (!v.m_valid[c]) continue;
          const uint32_t tid = v.m_tid[c]; uint32_t part = NONE;
          for (uint32_t d = mb; d < me; d++) if (v.m_valid[d] && v.m_tid[d] == tid) part = d;   // last one wins
          c_partner[c] = part;
          if (part != NONE) common++;
        }
        if (common > 0) { ucase = 1; rows = 2 * common; }                                       // same transcript
        else if (nr == 1 && nm == 1) { ucase = 2; rows = 2; c_partner[r_only] = m_only; }       // different transcripts
      }
    }
  }
  u_nrows[a] = rows; u_case[a] = ucase;
}

struct Rows {
  uint32_t* src; uint32_t* tid; uint32_t* pos; int8_t* strand; uint8_t* flags; double* sim; int32_t* clip;
  uint32_t* cand; uint32_t* ccap;
};

__global__ void __launch_bounds__(256) k_pair_emit(DevBatch b, PairView v, const uint32_t* __restrict__ c_partner,
                                                   const uint32_t* __restrict__ u_roff, const uint8_t* __restrict__ u_case,
                                                   const uint32_t* __restrict__ c_task, const uint32_t* __restrict__ m_pos,
                                                   const double* __restrict__ m_sim, const int32_t* __restrict__ m_clip,
                                                   const uint32_t* __restrict__ m_ilen, Rows o) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= b.n_aln) return;
  const uint32_t base = u_roff[a], rows = u_roff[a + 1] - base;
  if (rows == 0) return;
  const uint8_t ucase = u_case[a];
  const uint32_t cb = v.t_coff[2 * a], ce = v.t_coff[2 * a + 2];
  auto put = [&](uint32_t row, uint32_t c, uint32_t aln, uint8_t fl) {
    o.src[row] = aln; o.tid[row] = v.m_tid[c]; o.pos[row] = m_pos[c]; o.strand[row] = (c_task[c] & 1) ? '-' : '+';
    o.flags[row] = fl; o.sim[row] = m_sim[c]; o.clip[row] = m_clip[c]; o.cand[row] = c;
    o.ccap[row] = (b.cigar_off[aln + 1] - b.cigar_off[aln]) + m_ilen[c] + 2;
  };
  for (uint32_t c = cb; c < ce; c++) {
    if (!v.m_valid[c]) continue;
    const uint32_t part = c_partner[c];
    if (ucase != 0 && part == NONE) continue;
    const uint32_t tid = v.m_tid[c];
    uint32_t rank = 0; bool dup_later = false;
    for (uint32_t d = cb; d < ce; d++) {
      if (d == c || !v.m_valid[d]) continue;
      if (ucase != 0 && c_partner[d] == NONE) continue;
      const uint32_t t2 = v.m_tid[d];
      if (t2 < tid) rank++;
      else if (t2 == tid) { if (d < c) rank++; dup_later |= (d > c); }
    }
    (void)dup_later;
    if (ucase == 0) put(base + rank, c, a, BRB_REC_FIRST);
    else {
      const uint8_t fl = BRB_REC_PAIRED | (ucase == 1 ? BRB_REC_SAMETX : 0);
      put(base + 2 * rank, c, a, fl | BRB_REC_FIRST);
      put(base + 2 * rank + 1, part, (uint32_t)b.mate_idx[a], fl);
    }
  }
}

__global__ void k_group_rows(DevBatch b, const uint32_t* __restrict__ u_roff, uint32_t* __restrict__ g_roff) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g <= b.n_groups) g_roff[g] = u_roff[g < b.n_groups ? b.group_off[g] : b.n_aln];
}

}  // namespace brb
