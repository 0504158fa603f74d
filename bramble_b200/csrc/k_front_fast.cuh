// Short-read front kernel with a fast path (default short-read preset: no similarity filter).
//
// Same job and same results as k_front_short<false> (k_fused.cuh): exon-0 lookup, one candidate per record of the
// exon-0 range(s), evaluation along the transcript's exon chain (reference src/g2t.cpp:103-257,
// src/evaluate.cpp:111-282, 675-786, 888-1134).  What differs is HOW a surviving candidate is evaluated:
//
//   fast path   the perfectly spliced candidate -- read exon 0 starts inside guide exon 0 and (single exon) ends
//               inside it, or every junction of the read sits exactly on the junction between consecutive guide
//               exons of the transcript and the last read exon ends inside its guide exon.  With exact
//               boundaries every left/right insertion and gap of findOverlapping is 0, each later read exon
//               has exactly one surviving hit (the chain successor, exon_id + 1), and build_cigar_match only
//               ever appends M to M: the ideal CIGAR is the single word M(sum of exon lengths) and pos is the
//               offset inside the first ('+') or last ('-') guide exon.  A dozen integer compares per read
//               exon, no Hit bookkeeping, no CIGAR builder, no pool.
//   queue       every other surviving candidate (soft clips, shifted junctions, wrong isoforms that the walk
//               will eliminate, duplicate-tid neighbourhoods) is pushed to a per-warp queue in shared memory and
//               evaluated by the general code (dedupe_first + eval_short) 32 at a time.  In the flat loop one
//               such lane made the whole warp pay for the general path in ~60 % of the rounds; queued, the
//               general path runs with (nearly) all lanes busy.
//
// Slots: a surviving candidate reserves its slot when it is seen (ballot compaction keeps the slots of an
// alignment contiguous); a queued candidate that ends up eliminated leaves a hole (tid NONE), which the back
// kernel skips.  The order of an alignment's matches inside its slot range is immaterial (the back kernel ranks
// by tid and a tid occurs once per alignment).
#pragma once
#include "dev_common.cuh"
#include "k_lookup.cuh"
#include "k_chain.cuh"

#ifndef VAR_FAST_CTAS
#define VAR_FAST_CTAS 8
#endif

namespace brb {

// Verdict of the fast path on one surviving exon-0 candidate.
//   FAST_MATCH   the perfectly spliced case; pos / word are its match.
//   FAST_REJECT  (only with max_junction_insertion == max_junction_deletion == 0, the short-read preset) the walk is
//                certain to eliminate it: with zero junction tolerance a hit on read exon j >= 1 must start exactly at
//                the read exon's start, so only the immediate chain successor can survive findOverlapping, and it
//                has to be exon_id + 1 (correct_for_gaps) and, for a middle exon, end exactly at the read exon's end.
//   FAST_DEFER   anything else: the general path decides.
enum { FAST_DEFER = 0, FAST_MATCH = 1, FAST_REJECT = 2 };
// Only the KEY half of the exon-0 record is needed unless the read is spliced: the index build has checked that the exons of a
// transcript are disjoint and ascending (k_index_links, codes 3/4), so once read exon 0 is known to lie inside guide exon 0
// (st <= qs, qe <= en) no neighbour exon of the tid can overlap it -- the link half (exon_id, chain successor) is fetched for
// spliced reads only.  The same holds under the last read exon (qe <= en of its guide exon).
__device__ __forceinline__ int fast_short(const DevIndex& ix, const DevBatch& b, bool strict, const ReadExons& rx, int s, const RecKey& key0,
                                          uint32_t rec0, uint32_t& pos, uint32_t& word) {
  uint32_t qs = rx.qs0, qe = rx.qe0;
  if ((uint32_t)key0.st > qs || qe <= qs) return FAST_DEFER;                        // left soft clip / empty read exon
  const uint32_t first_pos = (qs - (uint32_t)key0.st) + key0.pos_start;             // src/g2t.cpp:160-163
  uint32_t total = qe - qs;
  int32_t k_en = key0.en; uint32_t k_pos_start = key0.pos_start;
  if (rx.n_ex > 1) {
    if (qe != (uint32_t)key0.en) return FAST_DEFER;
    const int miss = strict ? FAST_REJECT : FAST_DEFER;
    const RecLink lk0 = rec_link(ix, rec0);
    uint32_t prev_eid = lk0.meta & 0xffu, r = lk0.gnext;
    for (uint32_t j = 1; j < rx.n_ex; j++) {
      qs = __ldg(&b.exon_st[rx.e0 + j]); qe = __ldg(&b.exon_en[rx.e0 + j]);
      if (qe <= qs) return FAST_DEFER;
      if (r == NONE) return miss;
      const RecKey k = rec_key(ix, r); const uint2 l = rec_link_tail(ix, r);         // .x successor, .y meta
      if ((uint32_t)k.st != qs) return miss;
      const uint32_t eid = l.y & 0xffu;
      if (((eid - prev_eid) & 0xffu) != 1) return miss;                             // correct_for_gaps, src/evaluate.cpp:131-140
      if (j + 1 < rx.n_ex) { if ((uint32_t)k.en != qe) return miss; }
      else if (qe > (uint32_t)k.en) return FAST_DEFER;                               // right soft clip
      total += qe - qs; prev_eid = eid; r = l.x; k_en = k.en; k_pos_start = k.pos_start;
    }
  } else if (qe > (uint32_t)key0.en) return FAST_DEFER;                             // right soft clip
  pos = s ? ((uint32_t)k_en - qe) + k_pos_start : first_pos;                         // src/g2t.cpp:196-199; rcpos follows the last guide exon
  word = (total << 4) | CM;
  return FAST_MATCH;
}

constexpr uint32_t FQ_CAP = 64;          // queued candidates per warp (drained as soon as 32 are waiting)

__global__ void __launch_bounds__(256, VAR_FAST_CTAS) k_front_fast(DevIndex ix, DevBatch b, brb_params p, Ctl* __restrict__ ctl,
                                                       uint4* __restrict__ a_cand, uint32_t* __restrict__ ipool, Matches M) {
  __shared__ uint32_t s_all[8][9][32];                    // per alignment of the batch: qs0, qe0, e0, n_ex, excl, n0, lo0, lo1, matches from the queue
  __shared__ uint32_t s_q[8][3][FQ_CAP];                  // queue: record, owner lane | strand << 5, slot
  if (ctl->bad_input) return;                             // k_validate_batch rejected the CSR offsets: nothing may follow them
  const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint32_t (*s_al)[32] = s_all[wid];
  uint32_t* s_nv = s_al[8];
  uint32_t* q_rec = s_q[wid][0]; uint32_t* q_own = s_q[wid][1]; uint32_t* q_slot = s_q[wid][2];
  const uint32_t n_batches = (b.n_aln + 31) >> 5;
  const uint32_t cap_cand = ctl->cap_cand, cap_ideal = ctl->cap_ideal;       // constant during the launch
  // (exon_id differences are taken mod 256 like the reference's uint8_t: with a transcript of more than 256 exons a far
  //  successor could pass for exon_id + 1, so the shortcut is only taken when no chain is that long)
  const bool strict = p.max_junc_ins == 0 && p.max_junc_gap == 0 && ix.max_chain <= 256;
  WarpPool wp; wp.cur = 0; wp.end = 0;
  WarpPool sp; sp.cur = 0; sp.end = 0; (void)sp;
  unsigned long long my_hits = 0;
  uint32_t bad = 0, n_mates = 0, kept = 0;
  uint32_t wb_next = 0, wb_end = 0;
  for (;;) {
    if (wb_next == wb_end) {                                // dynamic batch tickets (batches differ a lot in cost: hot genes), kTicketFront batches each
      uint32_t t = 0;
      if (lane == 0) t = atomicAdd(&ctl->ticket, kTicketFront);
      wb_next = __shfl_sync(0xffffffffu, t, 0); wb_end = wb_next + kTicketFront;
    }
    const uint32_t wb = wb_next++;
    if (wb >= n_batches) break;
    const uint32_t a = wb * 32 + lane;
    ReadExons rx; rx.e0 = 0; rx.n_ex = 0; rx.qs0 = 0; rx.qe0 = 0;
    uint32_t lo0 = 0, n0 = 0, lo1 = 0, n1 = 0;
    if (a < b.n_aln) {
      rx.e0 = ld_stream<1>(&b.exon_off[a]); rx.n_ex = ld_stream<1>(&b.exon_off[a + 1]) - rx.e0;
      const int32_t refid = ld_stream<1>(&b.refid[a]);
      const int8_t sd = ld_stream<1>(&b.strand[a]);
      const int32_t m = ld_stream<1>(&b.mate_idx[a]);             // process_pairs links are single and symmetric (src/bramble.cpp:272-311)
      if (m >= 0 && ((uint32_t)m >= b.n_aln || (uint32_t)m == a || ld_stream<1>(&b.mate_idx[m]) != (int32_t)a)) bad = 1;
      if (m >= 0) n_mates++;
      if (rx.n_ex) {
        rx.qs0 = ld_stream<1>(&b.exon_st[rx.e0]); rx.qe0 = ld_stream<1>(&b.exon_en[rx.e0]);
        uint32_t pe = rx.qe0; if (rx.qe0 < rx.qs0) bad = 2;
        for (uint32_t j = 1; j < rx.n_ex; j++) {            // read exons ascending and disjoint (CIGAR order)
          const uint32_t qs = ld_stream<1>(&b.exon_st[rx.e0 + j]), qe = ld_stream<1>(&b.exon_en[rx.e0 + j]);
          if (qe < qs || qs < pe) bad = 2;
          pe = qe;
        }
        uint32_t hi;
        if (task_active(p, sd, 0)) { overlap_range(ix, refid, false, rx.qs0, rx.qe0, lo0, hi); n0 = hi - lo0; }
        if (task_active(p, sd, 1)) { overlap_range(ix, refid, true, rx.qs0, rx.qe0, lo1, hi); n1 = hi - lo1; }
      }
    }
    const uint32_t n = n0 + n1;
    const uint32_t incl = warp_incl_scan(n, lane), excl = incl - n, total = __shfl_sync(0xffffffffu, incl, 31);
    // candidate slots come out of a per-warp chunk of the slot space: one atomic per ~2048 slots instead of one per
    // batch (waiting for its return value was 10 % of the kernel's stall samples)
    if (total > sp.end - sp.cur) {
      const uint32_t grab = total > POOL_CHUNK ? total : POOL_CHUNK;
      uint32_t g0 = 0;
      if (lane == 0) g0 = atomicAdd(&ctl->n_cand, grab);
      g0 = __shfl_sync(0xffffffffu, g0, 0);
      sp.cur = g0; sp.end = g0 + grab;
    }
    const uint32_t sbase = sp.cur;
    const bool slots_fit = sp.end <= cap_cand && sp.end >= sp.cur;
    sp.cur += total;
    s_al[0][lane] = rx.qs0; s_al[1][lane] = rx.qe0; s_al[2][lane] = rx.e0; s_al[3][lane] = rx.n_ex;
    s_al[4][lane] = excl; s_al[5][lane] = n0; s_al[6][lane] = lo0; s_al[7][lane] = lo1; s_nv[lane] = 0;
    __syncwarp();
    uint32_t nres = 0, nvalid = 0, wr = 0, qn = 0;          // slots reserved / fast matches of MY alignment; slots reserved by the batch; queue length
    for (uint32_t base = 0; base < total || qn > 0;) {
      if (qn >= 32 || base >= total) {
        // ---- the general path for up to 32 queued candidates
        const uint32_t take = qn < 32 ? qn : 32;
        const bool go = lane < take;
        uint32_t rec0 = 0, L = 0, slot = 0; int s = 0;
        ReadExons ox; ox.qs0 = ox.qe0 = ox.e0 = ox.n_ex = 0;
        if (go) {
          rec0 = q_rec[lane]; L = q_own[lane] & 31u; s = (int)(q_own[lane] >> 5); slot = q_slot[lane];
          ox.qs0 = s_al[0][L]; ox.qe0 = s_al[1][L]; ox.e0 = s_al[2][L]; ox.n_ex = s_al[3][L];
        }
        bool pool_fit;
        const uint32_t ioff = warp_pool_alloc(wp, &ctl->n_ideal, cap_ideal, go ? 3 * ox.n_ex : 0, lane, pool_fit);   // at most S/I/D + M + S/I/D per read exon
        if (go) {
          bool ok = pool_fit; uint32_t pos = 0, ilen = 0, word = 0; double sim = 0.0;
          RecKey key; RecLink lk;
          if (ok) {
            ok = dedupe_first(ix, p, ox, s, rec0);
            if (ok) { key = rec_key(ix, rec0); lk = rec_link(ix, rec0); ok = eval_short<false>(ix, b, p, ox, s, key, lk, ipool + ioff, pos, ilen, word, sim); }
          }
          if (ok) { st_stream(&M.rec[slot], make_uint4(pos, (ilen << 1) | (uint32_t)s, ilen == 1 ? word : ioff, key.tid)); atomicAdd(&s_nv[L], 1u); }
          else st_stream(&M.rec[slot], make_uint4(0, 0, 0, NONE));
        }
        __syncwarp();
        if (lane + 32 < qn) { const uint32_t x = q_rec[lane + 32], y = q_own[lane + 32], z = q_slot[lane + 32]; q_rec[lane] = x; q_own[lane] = y; q_slot[lane] = z; }
        qn -= take;
        __syncwarp();
        continue;
      }
      // ---- one round of the flat walk: filter, fast path, queue the rest
      const uint32_t item = base + lane;
      const bool actv = item < total;
      const unsigned L = warp_owner(incl, actv ? item : total - 1);
      ReadExons ox;
      ox.qs0 = s_al[0][L]; ox.qe0 = s_al[1][L]; ox.e0 = s_al[2][L]; ox.n_ex = s_al[3][L];
      const uint32_t idx = item - s_al[4][L], o_n0 = s_al[5][L];
      const int s = idx >= o_n0;
      const uint32_t rec0 = s ? s_al[7][L] + (idx - o_n0) : s_al[6][L] + idx;
      bool sv = false, ov = false; RecKey key;
      if (actv) {
        key = rec_key(ix, rec0);
        ov = (int32_t)ox.qs0 < key.en;                     // st < qe holds for every record below hi
        sv = ov && slots_fit && passes_filter(p, s, exon_status(ox.n_ex, 0), ox.qs0, ox.qe0, (uint32_t)key.st, (uint32_t)key.en);
      }
      my_hits += __popc(__ballot_sync(0xffffffffu, ov)) * (lane == 0);
      bool fast = false; uint32_t pos = 0, word = 0;
      if (sv) { const int v = fast_short(ix, b, strict, ox, s, key, rec0, pos, word); fast = v == FAST_MATCH; sv = v != FAST_REJECT; }
      const uint32_t B_sv = __ballot_sync(0xffffffffu, sv), B_fast = __ballot_sync(0xffffffffu, fast), B_q = B_sv & ~B_fast;
      const uint32_t slot = sbase + wr + __popc(B_sv & lt_mask);
      if (fast) st_stream(&M.rec[slot], make_uint4(pos, 2u | (uint32_t)s, word, key.tid));
      else if (sv) { const uint32_t qi = qn + __popc(B_q & lt_mask); q_rec[qi] = rec0; q_own[qi] = L | ((uint32_t)s << 5); q_slot[qi] = slot; }
      wr += __popc(B_sv); qn += __popc(B_q);
      if (n && incl > base && excl < base + 32) {
        const uint32_t mine = bit_range(excl > base ? excl - base : 0, incl - base < 32 ? incl - base : 32);
        nres += __popc(B_sv & mine); nvalid += __popc(B_fast & mine);
      }
      base += 32;
      __syncwarp();                                         // queue entries are visible to the draining lanes
    }
    nvalid += s_nv[lane];
    kept += nvalid;
    const uint32_t first = sbase + warp_incl_scan(nres, lane) - nres;
    if (a < b.n_aln) st_stream(&a_cand[a], make_uint4(first, nres, nvalid, 0));
    __syncwarp();                                           // the per-alignment facts in shared memory are rewritten by the next batch
  }
  for (int d = 16; d; d >>= 1) { my_hits += __shfl_down_sync(0xffffffffu, my_hits, d); kept += __shfl_down_sync(0xffffffffu, kept, d);
                                 n_mates += __shfl_down_sync(0xffffffffu, n_mates, d); }
  if (lane == 0) {
    if (my_hits) atomicAdd(&ctl->n_hits, my_hits);
    if (kept) atomicAdd(&ctl->n_matches, (unsigned long long)kept);
    if (n_mates) atomicAdd(&ctl->n_mates, (unsigned long long)n_mates);
  }
  const uint32_t anybad = __reduce_max_sync(0xffffffffu, bad);
  if (lane == 0 && anybad) atomicMax(&ctl->bad_input, anybad);
}

}  // namespace brb
