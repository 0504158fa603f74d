// The front half of the projection as ONE persistent kernel: exon-overlap lookup, candidate enumeration and
// (FUSED) the whole per-candidate evaluation -- reference g2tTree::getGuideExons (src/g2t.cpp:103-257),
// get_intervals (src/evaluate.cpp:184-282) and evaluate_exon_chains (:888-1134) for a bundle.
//
// Work mapping.  A warp takes batches of 32 consecutive alignments (lane = alignment) in a grid-stride loop.
//   pass 1  for every (strand, exon index) round each lane brackets its query's index range through the bin
//           directory, then the warp walks the CONCATENATION of the 32 ranges 32 records at a time
//           (lane = record): a read hitting 200 isoforms of a hot gene costs 7 full rounds, not 200 single-lane
//           ones.  Ballots return survivors / first / last survivor to the owning lane.
//   alloc   one atomic per batch reserves the candidate slots of its 32 alignments ('+' task before '-' task,
//           slots of an alignment contiguous -- pairing relies on that).
//   pass 2  the tightened exon-0 ranges are walked again the same way; now lane = candidate transcript.  FUSED:
//           the lane evaluates its candidate in place (duplicate-tid rule, chain walk, ideal CIGAR, similarity)
//           with the record key already in registers and the read's exons hot in L1, and writes the match row.
//           !FUSED (clip-rescue path): the lane only records (record, task) for k_chain / k_ksw / k_build.
#pragma once
#include "dev_common.cuh"
#include "k_lookup.cuh"
#include "k_chain.cuh"

#ifndef VAR_FRONT_CTAS
#define VAR_FRONT_CTAS 5
#endif

// Resident CTAs per SM the long-read front kernel (k_front<false>) is compiled for: it waits on chains of dependent index loads, so it
// wants warps in flight more than registers (tools/variants.py + tools/time_variants_lr.sh: c4 front 0.45 ms at 4, 0.41 at 6 and 8).
#ifndef VAR_FRONT_SPLIT_CTAS
#define VAR_FRONT_SPLIT_CTAS 6
#endif

namespace brb {

// a_cand[a] = { first candidate slot, #slots ('+' task first, then '-'), #slots that hold a match, - }
// (on the clip-rescue path the match count is filled in later by k_count_valid)
template <bool FUSED>
__global__ void __launch_bounds__(256, FUSED ? 1 : VAR_FRONT_SPLIT_CTAS) k_front(DevIndex ix, DevBatch b, brb_params p, Ctl* __restrict__ ctl,
                                               uint8_t* q_any, uint4* __restrict__ a_cand,
                                               uint32_t* __restrict__ c_rec, uint32_t* __restrict__ c_task,
                                               uint32_t* __restrict__ ipool, Matches M) {
  if (ctl->bad_input) return;                             // k_validate_batch rejected the CSR offsets: nothing may follow them
  const unsigned lane = threadIdx.x & 31;
  // alignments per batch (b.front_shift): 32; for long reads 8, or 1 when a few reads carry most of the work (genes with hundreds of isoforms).  The kernel waits on dependent index loads (bin directory -> binary
  // search -> records), so it wants many warps in flight: a bundle of 10^5 long reads is only ~3000 batches of 32, fewer than the
  // resident warps.  The exon queries of a batch are flattened over the lanes, so small batches do not idle lanes.
  const uint32_t BS = b.front_shift, BA = 1u << BS;
  const uint32_t n_batches = (b.n_aln + BA - 1) >> BS;
  const uint32_t cap_cand = ctl->cap_cand, cap_ideal = ctl->cap_ideal;       // constant during the launch
  WarpPool wp; wp.cur = 0; wp.end = 0;
  unsigned long long my_hits = 0;
  uint32_t bad = 0, n_mates = 0, kept = 0;
  for (;;) {
    uint32_t wb = 0;                                        // dynamic batch tickets: batches differ a lot in cost (hot genes)
    if (lane == 0) wb = atomicAdd(&ctl->ticket, 1u);
    wb = __shfl_sync(0xffffffffu, wb, 0);
    if (wb >= n_batches) break;
    const uint32_t a = wb * BA + lane;
    const bool valid = lane < BA && a < b.n_aln;
    uint32_t e0 = 0, n_ex = 0; int32_t refid = -1; bool act[2] = {false, false};
    if (valid) {
      e0 = __ldg(&b.exon_off[a]); n_ex = __ldg(&b.exon_off[a + 1]) - e0; refid = __ldg(&b.refid[a]);
      const int8_t sd = b.strand[a];
      act[0] = task_active(p, sd, 0); act[1] = task_active(p, sd, 1);
      const int32_t m = __ldg(&b.mate_idx[a]);             // process_pairs links are single and symmetric (src/bramble.cpp:272-311)
      if (m >= 0 && ((uint32_t)m >= b.n_aln || (uint32_t)m == a || __ldg(&b.mate_idx[m]) != (int32_t)a)) bad = 1;
      if (m >= 0) n_mates++;
    }
    bool fail[2]; uint32_t ncand[2] = {0, 0}, clo[2] = {0, 0}, chi[2] = {0, 0};
    uint32_t st0 = 0, en0 = 0;                              // exon 0 (kept for pass 2)
    // ---- pass 1a: exon 0 of every alignment, both strands: the candidate lists (all hits enumerated)
    if (valid && n_ex) {
      st0 = __ldg(&b.exon_st[e0]); en0 = __ldg(&b.exon_en[e0]);
      uint32_t pe = en0; if (en0 < st0) bad = 2;
      for (uint32_t j = 1; j < n_ex; j++) {                 // read exons ascending and disjoint (CIGAR order)
        const uint32_t qs = __ldg(&b.exon_st[e0 + j]), qe = __ldg(&b.exon_en[e0 + j]);
        if (qe < qs || qs < pe) bad = 2;
        pe = qe;
      }
    }
#pragma unroll
    for (int s = 0; s < 2; s++) {
      fail[s] = (n_ex == 0);
      const bool have = valid && n_ex && act[s];
      uint32_t lo = 0, n = 0; const int status = exon_status(n_ex, 0);
      if (have) { uint32_t hi; overlap_range(ix, refid, s, st0, en0, lo, hi); n = hi - lo; }
      FlatOut f;
      warp_flat_count(ix, p, s, lane, lo, n, st0, en0, status, f);
      if (have) {
        my_hits += f.nhit;
        ncand[s] = f.nf; clo[s] = f.first; chi[s] = f.last;
        if (f.nf == 0) fail[s] = true;                      // exon 0 is FIRST or ONLY: never a skippable small middle exon
      }
      if (fail[s] || !act[s]) ncand[s] = 0;
    }
    // ---- pass 1b: the later exons of the strands still alive, as ONE flat list of queries (lane = query): a read of a 300-exon
    //      gene is 10-20 rounds of 32 queries instead of 300 rounds with one busy lane.  A later exon only has to have A surviving
    //      hit (else the strand fails, or the exon becomes an INS_EXON segment): unless the caller wants every hit counted
    //      (b.count_hits: the reference-defined figure for the roofline numerator) only the first PROBE records of a range are
    //      examined, the rest only for the queries that found nothing there.
    {
      constexpr uint32_t PROBE = 32;
      const uint32_t smask = (valid && n_ex > 1) ? ((act[0] && !fail[0]) ? 1u : 0u) | ((act[1] && !fail[1]) ? 2u : 0u) : 0u;
      const uint32_t nq = (n_ex > 1 ? n_ex - 1 : 0) * (uint32_t)__popc(smask);
      const uint32_t incl_q = warp_incl_scan(nq, lane), excl_q = incl_q - nq, total_q = __shfl_sync(0xffffffffu, incl_q, 31);
      uint32_t failed = 0;                                  // strands of MY alignment that a later exon failed (filled through the ballots below)
      for (uint32_t base = 0; base < total_q; base += 32) {
        const uint32_t item = base + lane;
        const bool actq = item < total_q;
        const unsigned L = warp_owner(incl_q, actq ? item : total_q - 1);
        const uint32_t o_ex = __shfl_sync(0xffffffffu, excl_q, L), o_e0 = __shfl_sync(0xffffffffu, e0, L), o_n = __shfl_sync(0xffffffffu, n_ex, L);
        const uint32_t o_sm = __shfl_sync(0xffffffffu, smask, L); const int32_t o_ref = __shfl_sync(0xffffffffu, refid, L);
        uint32_t lo = 0, n = 0, qs = 0, qe = 0, j = 0; int s = 0, status = 0;
        if (actq) {
          const uint32_t idx = item - o_ex, per = o_n - 1;
          s = o_sm == 3 ? (int)(idx / per) : (o_sm == 2 ? 1 : 0); j = 1 + idx % per;
          qs = __ldg(&b.exon_st[o_e0 + j]); qe = __ldg(&b.exon_en[o_e0 + j]);
          status = exon_status(o_n, j);
          uint32_t hi; overlap_range(ix, o_ref, s, qs, qe, lo, hi); n = hi - lo;
          // a later exon (MIDDLE or LAST) only accepts guide exons that START within the junction tolerances of its own start
          // (src/g2t.cpp:156-190): in a gene with hundreds of isoforms the overlap range is hundreds of records, that window a few
          if (!b.count_hits && n > PROBE) {
            const uint32_t w0 = st_lower_bound(ix, o_ref, s, qs > p.max_junc_gap ? qs - p.max_junc_gap : 0u), w1 = st_lower_bound(ix, o_ref, s, qs + p.max_junc_ins + 1);
            const uint32_t a0 = w0 > lo ? w0 : lo, a1 = w1 < hi ? w1 : hi;
            lo = a0; n = a1 > a0 ? a1 - a0 : 0;
          }
        }
        const uint32_t n1 = b.count_hits ? n : (n < PROBE ? n : PROBE);
        FlatOut f;
        warp_flat_count(ix, p, s, lane, lo, n1, qs, qe, status, f);
        uint32_t nf = f.nf, nhit = f.nhit;
        const bool more = actq && nf == 0 && n1 < n;
        if (__any_sync(0xffffffffu, more)) {
          FlatOut g;
          warp_flat_count(ix, p, s, lane, lo + n1, more ? n - n1 : 0, qs, qe, status, g);
          nf += g.nf; nhit += g.nhit;
        }
        bool fl = false;
        if (actq) {
          my_hits += nhit;
          if (status == MIDDLE_EXON) q_any[2 * (o_e0 + j) + s] = nf ? 1 : 0;     // only a MIDDLE exon can become an INS_EXON segment
          // src/evaluate.cpp:249-281: only a small MIDDLE exon may be skipped (data is non-empty for j>0)
          fl = nf == 0 && !(status == MIDDLE_EXON && p.ignore_small_exons && (qe - qs <= p.max_error_exon));
        }
        // tell the owners: for every failing query, its owner lane and strand
        for (uint32_t F = __ballot_sync(0xffffffffu, fl); F; F &= F - 1) {
          const int src = __ffs(F) - 1;
          const unsigned ownerL = __shfl_sync(0xffffffffu, L, src); const int fs = __shfl_sync(0xffffffffu, s, src);
          if (lane == ownerL) failed |= 1u << fs;
        }
      }
      if (failed & 1u) { fail[0] = true; ncand[0] = 0; }
      if (failed & 2u) { fail[1] = true; ncand[1] = 0; }
    }
    // ---- candidate slots of the batch
    const uint32_t tot = ncand[0] + ncand[1];
    const uint32_t incl_t = warp_incl_scan(tot, lane), wtot = __shfl_sync(0xffffffffu, incl_t, 31);
    uint32_t sbase = 0;
    if (lane == 31 && wtot) sbase = atomicAdd(&ctl->n_cand, wtot);
    sbase = __shfl_sync(0xffffffffu, sbase, 31);
    const uint32_t off0 = sbase + incl_t - tot, off1 = off0 + ncand[0];
    const bool slots_fit = (unsigned long long)sbase + wtot <= cap_cand;
    if (!wtot) { if (valid) a_cand[a] = make_uint4(off0, 0, 0, 0); continue; }
    uint32_t nvalid = 0;                                    // matches kept for MY alignment (pairing reads it from a_cand.w)
    __syncwarp();                                           // q_any of this batch is visible to its candidate lanes
    // ---- pass 2
#pragma unroll
    for (int s = 0; s < 2; s++) {
      const uint32_t lo = clo[s], n = ncand[s] ? chi[s] - clo[s] : 0, dst = s ? off1 : off0;
      const uint32_t incl = warp_incl_scan(n, lane), excl = incl - n;
      const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
      uint32_t written = 0;
      for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t item = base + lane;
        const bool actv = item < total;
        const unsigned L = warp_owner(incl, actv ? item : total - 1);
        const uint32_t o_lo = __shfl_sync(0xffffffffu, lo, L), o_ex = __shfl_sync(0xffffffffu, excl, L);
        ReadExons rx;
        rx.qs0 = __shfl_sync(0xffffffffu, st0, L); rx.qe0 = __shfl_sync(0xffffffffu, en0, L);
        rx.e0 = __shfl_sync(0xffffffffu, e0, L); rx.n_ex = __shfl_sync(0xffffffffu, n_ex, L);
        const uint32_t i = o_lo + (item - o_ex);
        bool sv = false;
        if (actv) {
          const RecKey k = rec_key(ix, i);
          sv = (int32_t)rx.qs0 < k.en && passes_filter(p, s, exon_status(rx.n_ex, 0), rx.qs0, rx.qe0, (uint32_t)k.st, (uint32_t)k.en);
        }
        const uint32_t B = __ballot_sync(0xffffffffu, sv);
        uint32_t mine = 0;
        if (n && incl > base && excl < base + 32) mine = bit_range(excl > base ? excl - base : 0, incl - base < 32 ? incl - base : 32);
        const uint32_t o_mine = __shfl_sync(0xffffffffu, mine, L), o_wr = __shfl_sync(0xffffffffu, written, L);
        const uint32_t o_dst = __shfl_sync(0xffffffffu, dst, L);
        const uint32_t slot = o_dst + o_wr + __popc(B & o_mine & ((1u << lane) - 1u));
        const uint32_t task = 2 * (wb * BA + L) + s;
        written += __popc(B & mine);
        if (!FUSED) {
          if (sv && slots_fit) { c_rec[slot] = i; c_task[slot] = task; }
        } else {
          bool ok = false; WalkOut w; w.cap = 0; w.first_rec = i; w.last_rec = i; w.n_seg = 0; w.n_guide = 0;
          if (sv) {
            uint32_t rec0 = i, fp, lp;
            ok = dedupe_first(ix, p, rx, s, rec0);
            if (ok) ok = walk<false>(ix, b, p, q_any, rx, s, rec0, w, nullptr, nullptr, nullptr, nullptr, fp, lp);
          }
          bool pool_fit;
          const uint32_t ioff = warp_pool_alloc(wp, &ctl->n_ideal, cap_ideal, ok ? w.cap : 0, lane, pool_fit);
          bool keep = false;
          if (sv && slots_fit) {
            if (ok && pool_fit) keep = build_one(ix, b, p, slot, task, rx, s, q_any, w, nullptr, nullptr, nullptr, ipool, ioff, M);
            else M.rec[slot] = make_uint4(0, 0, 0, NONE);
          }
          kept += keep ? 1u : 0u;
          nvalid += __popc(__ballot_sync(0xffffffffu, keep) & mine);
        }
      }
    }
    if (valid) a_cand[a] = make_uint4(off0, ncand[0] + ncand[1], nvalid, 0);
  }
  // work counters (algorithmic-bytes numerator) + input validation: one atomic per warp
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


// ---- short-read front kernel ---------------------------------------------------------------------------------
// Short-read presets (!long_reads, max_error_exon == 0, no clip rescue): a read exon without a surviving hit
// fails the strand (src/evaluate.cpp:249-281), and a candidate whose transcript misses an exon is eliminated by
// its own walk -- so only EXON 0 has to be looked up; every later exon is checked by following the candidate's
// exon chain.  That leaves one flat pass per batch of 32 alignments: bracket the exon-0 range(s) of each lane
// (both strands side by side), reserve one candidate slot per record of those ranges with a single atomic, then
// walk the concatenated ranges 32 records at a time; a lane whose record survives the overlap + tolerance
// filters evaluates that candidate on the spot (eval_short); matches are ballot-compacted so that those of an
// alignment are contiguous.  k_front_fast (k_front_fast.cuh) is this kernel with a fast path for perfectly spliced
// candidates and is what the default short-read preset runs; this one serves the similarity-filter overrides and
// the A/B tests (BRB_OPT_NO_FASTPATH).  (n_hits then only counts exon-0 hits; the reference-defined figure for the roofline numerator
// comes from the generic kernel, see BRB_OPT_GENERIC_FRONT.)
template <bool SIM>
__global__ void __launch_bounds__(256, VAR_FRONT_CTAS) k_front_short(DevIndex ix, DevBatch b, brb_params p, Ctl* __restrict__ ctl,
                                                        uint4* __restrict__ a_cand, uint32_t* __restrict__ ipool, Matches M) {
  if (ctl->bad_input) return;                             // k_validate_batch rejected the CSR offsets: nothing may follow them
  const unsigned lane = threadIdx.x & 31;
  __shared__ uint32_t s_all[8][9][32];
  uint32_t (*s_al)[32] = s_all[threadIdx.x >> 5];
  uint32_t* s_incl = s_al[8]; (void)s_incl;
  const uint32_t n_batches = (b.n_aln + 31) >> 5;
  const uint32_t cap_cand = ctl->cap_cand, cap_ideal = ctl->cap_ideal;       // constant during the launch
  WarpPool wp; wp.cur = 0; wp.end = 0;
  unsigned long long my_hits = 0;
  uint32_t bad = 0, n_mates = 0, kept = 0;
  for (;;) {
    uint32_t wb = 0;                                        // dynamic batch tickets: batches differ a lot in cost (hot genes)
    if (lane == 0) wb = atomicAdd(&ctl->ticket, 1u);
    wb = __shfl_sync(0xffffffffu, wb, 0);
    if (wb >= n_batches) break;
    const uint32_t a = wb * 32 + lane;
    ReadExons rx; rx.e0 = 0; rx.n_ex = 0; rx.qs0 = 0; rx.qe0 = 0;
    uint32_t lo0 = 0, n0 = 0, lo1 = 0, n1 = 0;
    if (a < b.n_aln) {
      rx.e0 = __ldg(&b.exon_off[a]); rx.n_ex = __ldg(&b.exon_off[a + 1]) - rx.e0;
      const int32_t refid = __ldg(&b.refid[a]);
      const int8_t sd = b.strand[a];
      const int32_t m = __ldg(&b.mate_idx[a]);             // process_pairs links are single and symmetric (src/bramble.cpp:272-311)
      if (m >= 0 && ((uint32_t)m >= b.n_aln || (uint32_t)m == a || __ldg(&b.mate_idx[m]) != (int32_t)a)) bad = 1;
      if (m >= 0) n_mates++;
      if (rx.n_ex) {
        rx.qs0 = __ldg(&b.exon_st[rx.e0]); rx.qe0 = __ldg(&b.exon_en[rx.e0]);
        uint32_t pe = rx.qe0; if (rx.qe0 < rx.qs0) bad = 2;
        for (uint32_t j = 1; j < rx.n_ex; j++) {            // read exons ascending and disjoint (CIGAR order)
          const uint32_t qs = __ldg(&b.exon_st[rx.e0 + j]), qe = __ldg(&b.exon_en[rx.e0 + j]);
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
    uint32_t sbase = 0;
    if (lane == 31 && total) sbase = atomicAdd(&ctl->n_cand, total);
    sbase = __shfl_sync(0xffffffffu, sbase, 31);
    const bool slots_fit = (unsigned long long)sbase + total <= cap_cand;
    // per-alignment facts of the batch go through shared memory: a lane fetches those of the alignment that owns
    // its candidate (one LDS each) and they do not sit in registers across the evaluation
    s_al[0][lane] = rx.qs0; s_al[1][lane] = rx.qe0; s_al[2][lane] = rx.e0; s_al[3][lane] = rx.n_ex;
    s_al[4][lane] = excl; s_al[5][lane] = n0; s_al[6][lane] = lo0; s_al[7][lane] = lo1;
    __syncwarp();
    uint32_t nvalid = 0, wr = 0;                            // matches kept for MY alignment / by the batch so far
    for (uint32_t base = 0; base < total; base += 32) {
      const uint32_t item = base + lane;
      const bool actv = item < total;
      const unsigned L = warp_owner(incl, actv ? item : total - 1);
      ReadExons ox;
      ox.qs0 = s_al[0][L]; ox.qe0 = s_al[1][L]; ox.e0 = s_al[2][L]; ox.n_ex = s_al[3][L];
      const uint32_t o_ex = s_al[4][L], o_n0 = s_al[5][L];
      const int o_status = exon_status(ox.n_ex, 0);
      const uint32_t idx = item - o_ex;
      const int s = idx >= o_n0;
      uint32_t rec0 = s ? s_al[7][L] + (idx - o_n0) : s_al[6][L] + idx;
      bool sv = false, ov = false; RecKey key;
      if (actv) {
        key = rec_key(ix, rec0);
        ov = (int32_t)ox.qs0 < key.en;                     // st < qe holds for every record below hi
        sv = ov && passes_filter(p, s, o_status, ox.qs0, ox.qe0, (uint32_t)key.st, (uint32_t)key.en);
      }
      my_hits += __popc(__ballot_sync(0xffffffffu, ov)) * (lane == 0);
      const uint32_t want = sv ? 3 * ox.n_ex : 0;           // at most S/I/D + M + S/I/D per read exon
      bool pool_fit;
      const uint32_t ioff = warp_pool_alloc(wp, &ctl->n_ideal, cap_ideal, want, lane, pool_fit);
      bool ok = false; uint32_t pos = 0, ilen = 0, word = 0; double sim = 0.0;
      if (sv && pool_fit && slots_fit) {
        RecLink lk = rec_link(ix, rec0);
        ok = true;
        if (lk.gprev_en > (int32_t)ox.qs0 || lk.gnext_st < (int32_t)ox.qe0) {      // a neighbour exon of this tid also overlaps read exon 0
          const uint32_t r1 = rec0;
          ok = dedupe_first(ix, p, ox, s, rec0);
          if (ok && rec0 != r1) { key = rec_key(ix, rec0); lk = rec_link(ix, rec0); }
        }
        if (ok) ok = eval_short<SIM>(ix, b, p, ox, s, key, lk, ipool + ioff, pos, ilen, word, sim);
      }
      // matches are written densely, in (alignment, record) order: slots of an alignment stay contiguous
      const uint32_t B_ok = __ballot_sync(0xffffffffu, ok);
      if (ok) {
        const uint32_t slot = sbase + wr + __popc(B_ok & ((1u << lane) - 1u));
        M.rec[slot] = make_uint4(pos, (ilen << 1) | (uint32_t)s, ilen == 1 ? word : ioff, key.tid);
        if (SIM) { M.sim[slot] = sim; M.clip[slot] = 0; }
      }
      wr += __popc(B_ok);
      if (n && incl > base && excl < base + 32) nvalid += __popc(B_ok & bit_range(excl > base ? excl - base : 0, incl - base < 32 ? incl - base : 32));
    }
    __syncwarp();                                           // the per-alignment facts in shared memory are rewritten by the next batch
    kept += nvalid;
    const uint32_t first = sbase + warp_incl_scan(nvalid, lane) - nvalid;
    if (a < b.n_aln) a_cand[a] = make_uint4(first, nvalid, nvalid, 0);
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
