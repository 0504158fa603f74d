// Kernel group 3: mate pairing, CIGAR projection and record emission -- the back half of a bundle.
// Replaces process_mate_pair / add_mate_info (reference src/mates.cpp:28-261), the group loop of convert_reads
// (src/core.cpp:384-415) and update_cigar / merge_cigars (src/bam.cpp:113-315,443-528).
// process_pairs (src/bramble.cpp:272-311) links every alignment to at most one mate and the link is symmetric,
// so inside a name group the units {i unpaired} and {i < mate(i)} are independent: the lower index is
// `this_read` (read1 of the BamInfo), the higher one the mate.  Rows of a unit come out ascending by tid (sorted
// tid lists + set_intersection in the reference), the read row immediately followed by its mate row.
//
// Whether a match of the read is emitted can be decided per match, from the match counts of the two ends that
// the front kernel left in a_cand.w (nr, nm):
//   nm == 0                      -> every match, unpaired            (case 0, src/mates.cpp:150-176)
//   the mate has the same tid    -> paired, same transcript          (case 1: a non-empty intersection keeps only
//                                                                      the common tids)
//   no common tid, nr == nm == 1 -> paired, different transcripts    (case 2)
// Work mapping (both kernels): a warp takes batches of 32 alignments (lane = alignment) and walks the
// concatenated candidate slots of the batch's UNIT OWNERS 32 at a time (lane = slot), so a 200-isoform gene
// fills the warp instead of idling 31 lanes, and slots of mates / non-owners cost nothing.
//   k_pair_a : partner search on the mate's slot list, the merged CIGAR of each emitted row into a scratch
//              pool (exact length known afterwards), rows + CIGAR words per alignment.
//   (scan)   : one 64-bit exclusive scan turns those into row / word offsets in bundle order.
//   k_pair_b : rank of each emitted match inside its unit, the row columns, and the CIGAR words copied to their
//              final, dense, row-ordered place (reversed for '-' transcripts).
#pragma once
#include "dev_common.cuh"
#include "k_merge.cuh"

namespace brb {

constexpr uint32_t SELF = 0xfffffffeu;   // emitted without a mate row
constexpr uint32_t INLINE = 0xffffffffu; // SlotOut::rec.w: both merged CIGARs are one word each and live in the record
constexpr unsigned long long NONE64 = ~0ull;

struct PairIn {
  const uint4* a_cand;          // per alignment: first slot, #slots '+', #slots '-', #matches
  const uint32_t* m_tid;        // per slot: tid of the match or NONE
  const uint4* m_rec;           // per slot holding a match: pos, ideal-CIGAR length, pool offset or the word itself, task
  const double* m_sim; const int32_t* m_clip;      // null when every score is 0
  const uint32_t* ipool;        // ideal CIGARs longer than one word
};
struct SlotOut {                // per slot of a unit owner that has matches
  unsigned long long* key;      // emitted match: tid << 32 | CIGAR words of its row(s); otherwise NONE64
  uint4* rec;                   // emitted match: .x partner (SELF: unpaired row, else the mate's slot) and
                                //   .w == INLINE: .y / .z = the one-word merged CIGAR of the read row / mate row
                                //   otherwise   : .y / .z = CIGAR lengths, .w = scratch-pool offset (mate row after the read row's capacity)
};

// [lo,hi) = my lanes of this chunk; sum over them of a per-lane value given its inclusive warp prefix P
__device__ __forceinline__ uint32_t seg_sum(uint32_t P, uint32_t lo, uint32_t hi) {
  const uint32_t top = __shfl_sync(0xffffffffu, P, hi ? hi - 1 : 0), bot = __shfl_sync(0xffffffffu, P, lo ? lo - 1 : 0);
  return hi > lo ? top - (lo ? bot : 0) : 0;
}

__global__ void __launch_bounds__(256, 6) k_pair_a(DevBatch b, Ctl* __restrict__ ctl, PairIn v, SlotOut so, uint32_t* __restrict__ tpool,
                                                unsigned long long* __restrict__ u_cnt /* rows | words << 32 */) {
  if (ctl_overflow(ctl)) return;
  const unsigned lane = threadIdx.x & 31;
  const uint32_t n_batches = (b.n_aln + 31) >> 5;
  const uint32_t cap_tmpw = ctl->cap_tmpw;
  WarpPool wp; wp.cur = 0; wp.end = 0;
  for (;;) {
    uint32_t wb = 0;
    if (lane == 0) wb = atomicAdd(&ctl->ticket_a, 1u);
    wb = __shfl_sync(0xffffffffu, wb, 0);
    if (wb >= n_batches) break;
    const uint32_t a = wb * 32 + lane;
    uint32_t off = 0, n = 0, nr = 0, nm = 0, m_off = 0, m_n = 0, ro = 0, nrc = 0, mro = 0, mnrc = 0;
    if (a < b.n_aln) {
      const uint4 ac = __ldg(&v.a_cand[a]);
      const int32_t mate = __ldg(&b.mate_idx[a]);
      if ((mate < 0 || (uint32_t)mate > a) && ac.w) {         // unit owner with at least one match
        off = ac.x; n = ac.y + ac.z; nr = ac.w;
        ro = __ldg(&b.cigar_off[a]); nrc = __ldg(&b.cigar_off[a + 1]) - ro;
        if (mate >= 0) {
          const uint4 mc = __ldg(&v.a_cand[mate]);
          m_off = mc.x; m_n = mc.y + mc.z; nm = mc.w;
          mro = __ldg(&b.cigar_off[mate]); mnrc = __ldg(&b.cigar_off[mate + 1]) - mro;
        }
      }
    }
    const uint32_t incl = warp_incl_scan(n, lane), excl = incl - n, total = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t rows = 0, words = 0;
    for (uint32_t base = 0; base < total; base += 32) {
      const uint32_t item = base + lane;
      const bool actv = item < total;
      const unsigned L = warp_owner(incl, actv ? item : total - 1);
      const uint32_t c = __shfl_sync(0xffffffffu, off, L) + (item - __shfl_sync(0xffffffffu, excl, L));
      const uint32_t o_nr = __shfl_sync(0xffffffffu, nr, L), o_nm = __shfl_sync(0xffffffffu, nm, L);
      const uint32_t o_moff = __shfl_sync(0xffffffffu, m_off, L), o_mn = __shfl_sync(0xffffffffu, m_n, L);
      const uint32_t o_ro = __shfl_sync(0xffffffffu, ro, L), o_nrc = __shfl_sync(0xffffffffu, nrc, L);
      const uint32_t o_mro = __shfl_sync(0xffffffffu, mro, L), o_mnrc = __shfl_sync(0xffffffffu, mnrc, L);
      const uint32_t tid = actv ? v.m_tid[c] : NONE;
      uint32_t part = NONE;
      if (tid != NONE) {
        if (o_nm == 0) part = SELF;
        else {
          uint32_t only = NONE;
          for (uint32_t d = o_moff, e = o_moff + o_mn; d < e; d++) { const uint32_t t = v.m_tid[d]; if (t == tid) part = d; if (t != NONE) only = d; }
          if (part == NONE && o_nr == 1 && o_nm == 1) part = only;          // different transcripts
        }
      }
      const bool emitted = part != NONE, paired = emitted && part != SELF;
      uint4 mr = make_uint4(0, 0, 0, 0), pr = make_uint4(0, 0, 0, 0);
      if (emitted) mr = v.m_rec[c];
      if (paired) pr = v.m_rec[part];
      // one-word results stay in the slot record: no scratch traffic for the common short-read row
      const bool inl = emitted && mr.y == 1 && merge_is_trivial(b.cigar + o_ro, o_nrc, mr.z) &&
                       (!paired || (pr.y == 1 && merge_is_trivial(b.cigar + o_mro, o_mnrc, pr.z)));
      const uint32_t cap_r = o_nrc + mr.y + 2, want = (emitted && !inl) ? cap_r + (paired ? o_mnrc + pr.y + 2 : 0) : 0;
      bool fit;
      const uint32_t toff = warp_pool_alloc(wp, &ctl->n_tmpw, cap_tmpw, want, lane, fit);
      uint32_t clr = 0, clm = 0;
      if (inl) { clr = 1; clm = paired ? 1 : 0; }
      else if (emitted && fit) {
        // (one-word ideal CIGARs are not in the pool: .z is the word)
        clr = merge_cigars(b.cigar + o_ro, o_nrc, v.ipool + mr.z, mr.y, mr.z, tpool + toff);
        if (paired) clm = merge_cigars(b.cigar + o_mro, o_mnrc, v.ipool + pr.z, pr.y, pr.z, tpool + toff + cap_r);
      }
      if (actv) {
        so.key[c] = emitted ? ((unsigned long long)tid << 32) | (clr + clm) : NONE64;
        if (emitted) so.rec[c] = inl ? make_uint4(part, mr.z, pr.z, INLINE) : make_uint4(part, clr, clm, toff);
      }
      // rows / words of MY alignment in this chunk
      const uint32_t Pr = warp_incl_scan(emitted ? (paired ? 2u : 1u) : 0u, lane), Pw = warp_incl_scan(clr + clm, lane);
      uint32_t lo_b = 0, hi_b = 0;
      if (n && incl > base && excl < base + 32) { lo_b = excl > base ? excl - base : 0; hi_b = incl - base < 32 ? incl - base : 32; }
      rows += seg_sum(Pr, lo_b, hi_b); words += seg_sum(Pw, lo_b, hi_b);
    }
    if (a < b.n_aln) u_cnt[a] = (unsigned long long)rows | ((unsigned long long)words << 32);
  }
}

struct Rows {
  uint32_t* src; uint32_t* tid; uint32_t* pos; int8_t* strand; uint8_t* flags; double* sim; int32_t* clip;
  uint32_t* coff;               // n_rows+1: dense CIGAR pool offsets (CSR)
  uint32_t* pool;               // dense, row-ordered merged CIGARs
};

// u_off = exclusive scan of u_cnt (n_aln+1 entries; low half rows, high half CIGAR words).
__global__ void __launch_bounds__(256, 6) k_pair_b(DevBatch b, Ctl* __restrict__ ctl, PairIn v, SlotOut so, const uint32_t* __restrict__ tpool,
                                                const unsigned long long* __restrict__ u_cnt, const unsigned long long* __restrict__ u_off,
                                                Rows o) {
  if (ctl_overflow(ctl)) return;
  const unsigned long long tot = u_off[b.n_aln];
  const uint32_t R = (uint32_t)tot, W = (uint32_t)(tot >> 32);
  if (blockIdx.x == 0 && threadIdx.x == 0) { ctl->n_rows = R; ctl->n_outw = W; ctl->out_words = W; }
  if (R > ctl->cap_rows || W > ctl->cap_outw) return;
  if (blockIdx.x == 0 && threadIdx.x == 0) o.coff[R] = W;
  const unsigned lane = threadIdx.x & 31;
  const uint32_t n_batches = (b.n_aln + 31) >> 5;
  for (;;) {
    uint32_t wb = 0;
    if (lane == 0) wb = atomicAdd(&ctl->ticket_b, 1u);
    wb = __shfl_sync(0xffffffffu, wb, 0);
    if (wb >= n_batches) break;
    const uint32_t a = wb * 32 + lane;
    uint32_t off = 0, n = 0, roff = 0, woff = 0, nrc = 0; int32_t mate = -1;
    if (a < b.n_aln && (uint32_t)u_cnt[a]) {                    // the alignment emits rows
      const uint4 ac = __ldg(&v.a_cand[a]);
      off = ac.x; n = ac.y + ac.z;
      const unsigned long long uo = u_off[a];
      roff = (uint32_t)uo; woff = (uint32_t)(uo >> 32);
      nrc = __ldg(&b.cigar_off[a + 1]) - __ldg(&b.cigar_off[a]);
      mate = __ldg(&b.mate_idx[a]);
    }
    const uint32_t incl = warp_incl_scan(n, lane), excl = incl - n, total = __shfl_sync(0xffffffffu, incl, 31);
    for (uint32_t base = 0; base < total; base += 32) {
      const uint32_t item = base + lane;
      const bool actv = item < total;
      const unsigned L = warp_owner(incl, actv ? item : total - 1);
      const uint32_t o_off = __shfl_sync(0xffffffffu, off, L), o_n = __shfl_sync(0xffffffffu, n, L);
      const uint32_t c = o_off + (item - __shfl_sync(0xffffffffu, excl, L));
      const uint32_t o_roff = __shfl_sync(0xffffffffu, roff, L), o_woff = __shfl_sync(0xffffffffu, woff, L);
      const uint32_t o_nrc = __shfl_sync(0xffffffffu, nrc, L);
      const int32_t o_mate = __shfl_sync(0xffffffffu, mate, L);
      if (!actv) continue;
      const unsigned long long mykey = so.key[c];
      if (mykey == NONE64) continue;
      const uint32_t tid = (uint32_t)(mykey >> 32);
      // rank among the emitted matches of the read: ascending tid, slot order among equals; CIGAR words before me
      uint32_t rank = 0, wpre = 0;
      for (uint32_t d = o_off, e = o_off + o_n; d < e; d++) {
        const unsigned long long k = so.key[d];
        const uint32_t t2 = (uint32_t)(k >> 32);
        if (k != NONE64 && (t2 < tid || (t2 == tid && d < c))) { rank++; wpre += (uint32_t)k; }
      }
      const uint4 sr = so.rec[c], mr = v.m_rec[c];
      const uint32_t part = sr.x;
      const bool paired = part != SELF, inl = sr.w == INLINE;
      const uint32_t row = o_roff + (paired ? 2 * rank : rank), w0 = o_woff + wpre;
      const uint32_t clr = inl ? 1u : sr.y;
      auto put = [&](uint32_t r, uint32_t cc, const uint4& m, uint32_t aln, uint32_t t, uint8_t fl, uint32_t wpos, uint32_t word, const uint32_t* src, uint32_t len) {
        const bool minus = m.w & 1;
        o.src[r] = aln; o.tid[r] = t; o.pos[r] = m.x; o.strand[r] = minus ? '-' : '+'; o.flags[r] = fl; o.coff[r] = wpos;
        if (v.m_sim) { o.sim[r] = v.m_sim[cc]; o.clip[r] = v.m_clip[cc]; }
        if (inl) o.pool[wpos] = word;
        else if (minus) for (uint32_t i = 0; i < len; i++) o.pool[wpos + i] = src[len - 1 - i];      // reverse_complement_bam, src/bam.cpp:636-702
        else for (uint32_t i = 0; i < len; i++) o.pool[wpos + i] = src[i];
      };
      if (!paired) put(row, c, mr, wb * 32 + L, tid, BRB_REC_FIRST, w0, sr.y, tpool + sr.w, clr);
      else {
        const uint32_t t2 = v.m_tid[part];
        const uint4 pm = v.m_rec[part];
        const uint8_t fl = BRB_REC_PAIRED | (t2 == tid ? BRB_REC_SAMETX : 0);
        put(row, c, mr, wb * 32 + L, tid, fl | BRB_REC_FIRST, w0, sr.y, tpool + sr.w, clr);
        put(row + 1, part, pm, (uint32_t)o_mate, t2, fl, w0 + clr, sr.z, tpool + sr.w + (o_nrc + mr.y + 2), inl ? 1u : sr.z);
      }
    }
  }
}

// Clip-rescue path only: the front kernel cannot know which candidates survive k_ksw / k_build, so the match
// count of each alignment (a_cand.w) is filled in afterwards.
__global__ void __launch_bounds__(256) k_count_valid(DevBatch b, const Ctl* __restrict__ ctl, uint4* __restrict__ a_cand, const uint32_t* __restrict__ m_tid) {
  if (ctl_overflow(ctl)) return;
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= b.n_aln) return;
  uint4 ac = a_cand[a]; uint32_t nv = 0;
  for (uint32_t c = ac.x, e = ac.x + ac.y + ac.z; c < e; c++) nv += m_tid[c] != NONE;
  ac.w = nv; a_cand[a] = ac;
}

__global__ void k_group_rows(DevBatch b, const unsigned long long* __restrict__ u_off, uint32_t* __restrict__ g_roff) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g <= b.n_groups) g_roff[g] = (uint32_t)u_off[g < b.n_groups ? b.group_off[g] : b.n_aln];
}

}  // namespace brb
