// Kernel group 3: mate pairing and record emission.
// Replaces process_mate_pair / add_mate_info (reference src/mates.cpp:28-261) and the group loop of
// convert_reads (src/core.cpp:384-415).  process_pairs (src/bramble.cpp:272-311) links every
// alignment to at most one mate and the link is symmetric, so inside a name group the units
// {i unpaired} and {i < mate(i)} are independent: the lower index is `this_read` (read1 of the
// BamInfo), the higher one the mate.  Rows of a unit come out ascending by tid (sorted tid lists +
// set_intersection in the reference), the read row immediately followed by its mate row.
//
// Work mapping: one thread per CANDIDATE SLOT (not per alignment), so the lanes of a warp that belong to
// the same read run the same trip count over that read's (or its mate's) match list and a 200-isoform
// gene does not leave 31 lanes idle.  Per-alignment totals are folded with match.any + one atomic per run.
#pragma once
#include "dev_common.cuh"

namespace brb {

constexpr uint32_t SELF = 0xfffffffeu;   // emitted without a mate row

struct PairView {
  const uint4* a_cand;          // per alignment: first slot, #slots '+', #slots '-'
  const uint32_t* c_task;       // per slot: 2a+s
  const uint32_t* m_tid;        // per slot: tid of the match or NONE
};

// Pass A (per slot): partner match on the mate (last one wins, like the reference's map overwrite) and the
// per-alignment totals  a_cnt[a] = {valid matches, matches with a partner}.
__global__ void __launch_bounds__(256) k_partner(DevBatch b, const Ctl* __restrict__ ctl, PairView v,
                                                 uint32_t* __restrict__ c_partner, uint2* __restrict__ a_cnt) {
  if (ctl_overflow(ctl)) return;
  const unsigned lane = threadIdx.x & 31;
  const uint32_t n_cand = ctl->n_cand;
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t base = blockIdx.x * blockDim.x + (threadIdx.x & ~31u); base < n_cand; base += stride) {
    const uint32_t c = base + lane;
    const bool act = c < n_cand;
    uint32_t a = NONE, part = NONE; bool valid = false;
    if (act) {
      const uint32_t tid = v.m_tid[c];
      valid = tid != NONE;
      if (valid) {
        a = v.c_task[c] >> 1;                                // (c_task is only defined for slots that hold a match)
        const int32_t mate = __ldg(&b.mate_idx[a]);
        if (mate < 0) part = SELF;
        else if ((uint32_t)mate > a) {                       // only the unit owner (lower index) needs partners
          const uint4 mc = __ldg(&v.a_cand[mate]);
          for (uint32_t d = mc.x, e = mc.x + mc.y + mc.z; d < e; d++) if (v.m_tid[d] == tid) part = d;
        }
      }
      if (valid) c_partner[c] = part;
    }
    const uint32_t same = __match_any_sync(0xffffffffu, a);
    const uint32_t Bv = __ballot_sync(0xffffffffu, valid);
    const uint32_t Bp = __ballot_sync(0xffffffffu, valid && part != NONE && part != SELF);
    if (act && lane == (unsigned)__ffs(same) - 1) {
      const uint32_t nv = __popc(same & Bv), np = __popc(same & Bp);
      if (nv) atomicAdd(&a_cnt[a].x, nv);
      if (np) atomicAdd(&a_cnt[a].y, np);
    }
  }
}

// Pass B (per alignment): the pairing case and the number of rows of the unit (src/mates.cpp:150-261).
__global__ void __launch_bounds__(256) k_pair_count(DevBatch b, const Ctl* __restrict__ ctl, PairView v, const uint2* __restrict__ a_cnt,
                                                    uint32_t* __restrict__ c_partner, uint32_t* __restrict__ u_nrows,
                                                    uint8_t* __restrict__ u_case) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= b.n_aln) return;
  uint32_t rows = 0; uint8_t ucase = 255;
  if (!ctl_overflow(ctl)) {
    const int32_t mate = b.mate_idx[a];
    if (mate < 0 || (uint32_t)mate > a) {
      const uint2 cr = a_cnt[a];
      const uint32_t nr = cr.x, nm = mate >= 0 ? a_cnt[mate].x : 0;
      if (nr > 0) {
        if (nm == 0) { ucase = 0; rows = nr; }                             // case 0: mate absent or without a match
        else if (cr.y > 0) { ucase = 1; rows = 2 * cr.y; }                 // same transcript
        else if (nr == 1 && nm == 1) {                                     // different transcripts
          ucase = 2; rows = 2;
          const uint4 rc = v.a_cand[a], mc = v.a_cand[mate];
          uint32_t r_only = NONE, m_only = NONE;
          for (uint32_t c = rc.x, e = rc.x + rc.y + rc.z; c < e; c++) if (v.m_tid[c] != NONE) r_only = c;
          for (uint32_t c = mc.x, e = mc.x + mc.y + mc.z; c < e; c++) if (v.m_tid[c] != NONE) m_only = c;
          c_partner[r_only] = m_only;
        }
      }
    }
  }
  u_nrows[a] = rows; u_case[a] = ucase;
}

struct Rows {
  uint32_t* src; uint32_t* tid; uint32_t* pos; int8_t* strand; uint8_t* flags; double* sim; int32_t* clip;
  uint32_t* cand;
};

// Pass C (per slot of a unit owner): rank of the match among the emitted matches of its read (ascending tid,
// slot order among equals) and the row(s) themselves.
__global__ void __launch_bounds__(256) k_pair_emit(DevBatch b, Ctl* __restrict__ ctl, PairView v, const uint32_t* __restrict__ c_partner,
                                                   const uint32_t* __restrict__ u_roff, const uint8_t* __restrict__ u_case,
                                                   const uint32_t* __restrict__ m_pos, const double* __restrict__ m_sim,
                                                   const int32_t* __restrict__ m_clip, int has_scores, Rows o) {
  if (ctl_overflow(ctl)) return;
  const uint32_t R = u_roff[b.n_aln];
  if (blockIdx.x == 0 && threadIdx.x == 0) ctl->n_rows = R;
  if (R > ctl->cap_rows) return;
  const uint32_t n_cand = ctl->n_cand;
  for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < n_cand; c += gridDim.x * blockDim.x) {
    const uint32_t tid = v.m_tid[c];
    if (tid == NONE) continue;
    const uint32_t task = v.c_task[c], a = task >> 1;
    const uint8_t ucase = u_case[a];
    if (ucase == 255) continue;
    const uint32_t part = c_partner[c];
    if (ucase != 0 && (part == NONE || part == SELF)) continue;
    const uint4 ac = __ldg(&v.a_cand[a]);
    uint32_t rank = 0;
    for (uint32_t d = ac.x, e = ac.x + ac.y + ac.z; d < e; d++) {
      const uint32_t t2 = v.m_tid[d];
      if (d == c || t2 == NONE) continue;
      if (ucase != 0) { const uint32_t pd = c_partner[d]; if (pd == NONE || pd == SELF) continue; }
      if (t2 < tid || (t2 == tid && d < c)) rank++;
    }
    const uint32_t base = u_roff[a];
    auto put = [&](uint32_t row, uint32_t cc, uint32_t aln, uint32_t t, uint8_t fl) {
      o.src[row] = aln; o.tid[row] = t; o.pos[row] = m_pos[cc]; o.strand[row] = (v.c_task[cc] & 1) ? '-' : '+';
      o.flags[row] = fl; o.cand[row] = cc;
      if (has_scores) { o.sim[row] = m_sim[cc]; o.clip[row] = m_clip[cc]; }
    };
    if (ucase == 0) put(base + rank, c, a, tid, BRB_REC_FIRST);
    else {
      const uint8_t fl = BRB_REC_PAIRED | (ucase == 1 ? BRB_REC_SAMETX : 0);
      put(base + 2 * rank, c, a, tid, fl | BRB_REC_FIRST);
      put(base + 2 * rank + 1, part, (uint32_t)b.mate_idx[a], v.m_tid[part], fl);
    }
  }
}

__global__ void k_group_rows(DevBatch b, const uint32_t* __restrict__ u_roff, uint32_t* __restrict__ g_roff) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g <= b.n_groups) g_roff[g] = u_roff[g < b.n_groups ? b.group_off[g] : b.n_aln];
}

}  // namespace brb
