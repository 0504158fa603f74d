// Kernel group 3: mate pairing, CIGAR projection and record emission -- the back half of a bundle, ONE kernel.
// Replaces process_mate_pair / add_mate_info (reference src/mates.cpp:28-261), the group loop of convert_reads
// (src/core.cpp:384-415) and update_cigar / merge_cigars (src/bam.cpp:113-315,443-528).
// process_pairs (src/bramble.cpp:272-311) links every alignment to at most one mate and the link is symmetric,
// so inside a name group the units {i unpaired} and {i < mate(i)} are independent: the lower index is
// `this_read` (read1 of the BamInfo), the higher one the mate.  Rows of a unit come out ascending by tid (sorted
// tid lists + set_intersection in the reference), the read row immediately followed by its mate row.
//
// Whether a match of the read is emitted can be decided per match, from the match counts of the two ends that
// the front kernel left in a_cand.z (nr, nm):
//   nm == 0                      -> every match, unpaired            (case 0, src/mates.cpp:150-176)
//   the mate has the same tid    -> paired, same transcript          (case 1: a non-empty intersection keeps only
//                                                                      the common tids)
//   no common tid, nr == nm == 1 -> paired, different transcripts    (case 2)
//
// Work mapping (k_back): persistent warps take batches of 32 alignments by ticket (lane = alignment) and walk the
// concatenated match slots of the batch's UNIT OWNERS 32 at a time (lane = slot), so a 200-isoform gene fills
// the warp instead of idling 31 lanes, and slots of mates / non-owners cost nothing.
//   phase 1   partner search on the mate's slot list; the common short-read row (real CIGAR = M/N only, ideal
//             CIGAR = one M of the same length) is recognised from a per-alignment sum and needs no merge; every
//             other emitted row is queued and merged afterwards with all lanes busy (merge_cigars into a scratch
//             pool).  Per-match results stay in shared memory (global scratch for batches beyond BACK_STAGE).
//   claim     one atomic reserves the batch's rows and CIGAR words in a dense buffer -- in COMPLETION order, so a
//             batch full of hot-gene reads never stalls the batches after it (an in-order single-pass scan did:
//             half of the issue slots went to the look-back spin).
//   phase 2   rank of each emitted match inside its unit (ascending tid), its row(s) and the CIGAR words (reversed
//             for '-' transcripts), written where the batch claimed room.
// Default: that place is final -- the rows of a unit are found through a_roff / a_rcnt (brb_batch_out::aln_rec_off /
// aln_rec_cnt).  BRB_OPT_ORDERED_ROWS: phase 2 writes 16-byte packed row records into a provisional buffer and k_place
// (after a scan of the 32-alignment batch totals) moves every batch to its place in bundle order while unpacking
// them into the output columns: streaming, fully coalesced.
#pragma once
#include "dev_common.cuh"
#include "k_merge.cuh"

// Tuning knobs (defaults = measured best on B200; tools/variants.py builds alternatives with -DVAR_...=x)
#ifndef VAR_UNROLL
#define VAR_UNROLL 4
#endif
#ifndef VAR_PLACE_CTAS
#define VAR_PLACE_CTAS 5
#endif
#ifndef VAR_PLACE_ROWS
#define VAR_PLACE_ROWS 4
#endif
#ifndef VAR_PARTNER_GUESS
#define VAR_PARTNER_GUESS 1          // measured: k_back 0.485 -> 0.477 ms on the C2 workload
#endif
#ifndef VAR_BACK_STAGE
#define VAR_BACK_STAGE 192
#endif
#ifndef VAR_BACK_CTAS
#define VAR_BACK_CTAS 5
#endif

namespace brb {

constexpr uint32_t SELF = 0xfffffffeu;     // partner: emitted without a mate row
constexpr int kUnroll = VAR_UNROLL;
constexpr uint32_t INLINE = 0xffffffffu;   // Stage.w: every merged CIGAR of the row(s) is the one-word ideal CIGAR itself
constexpr uint32_t PENDING = 0xfffffffeu;  // Stage.w: general merge queued
constexpr uint32_t NOFIT = 0xfffffffdu;    // Stage.w: the scratch pool is exhausted (the host grows it and replays)
constexpr uint32_t PREMERGED = 0xfffffffcu; // Stage.w (long reads): the merged CIGARs of the row(s) are where BackIn::mg says
constexpr uint32_t BACK_STAGE = VAR_BACK_STAGE;       // matches per batch staged in shared memory

// match record written by the front kernels, one uint4 per candidate slot:
//   .x pos   .y ideal-CIGAR length << 1 | transcript strand is '-'   .z its pool offset, or the word itself when the length is 1
//   .w tid (NONE: the slot holds no match)
struct BackIn {
  const uint4* a_cand;          // per alignment: first slot, #slots to walk, #matches among them
  const uint4* m_rec;
  const double* m_sim; const int32_t* m_clip;      // null when every score is 0
  const uint32_t* ipool;        // ideal CIGARs longer than one word
  const uint2* mg;              // long reads: merged CIGAR of every match, made by k_merge_long: .x offset in the scratch pool (NONE: the one-word case, the word is m_rec.z; NOFIT: pool exhausted), .y words
};
struct Rows {
  uint32_t* src; uint32_t* tid; uint32_t* pos; int8_t* strand; uint8_t* flags; double* sim; int32_t* clip;
  uint32_t* coff;               // n_rows+1: dense CIGAR pool offsets (CSR)
  uint32_t* pool;               // dense, row-ordered merged CIGARs
  uint32_t* a_roff;             // n_aln (+1 when ordered): first row of each alignment's unit
  uint32_t* a_rcnt;             // n_aln: rows of each alignment's unit (0 for mates and for units that emit nothing)
};

// Rows as k_back leaves them: dense, but batches in completion order.  One 16-byte record per row:
//   .x source alignment  .y tid  .z pos  .w first CIGAR word relative to the batch (28 bits) | '-' strand << 28 | BRB_REC_* flags << 29
struct ProvRows { uint4* row; uint32_t* pool; double* sim; int32_t* clip; };
// Per batch of 32 alignments: rows | CIGAR words << 32, and where k_back put them.
struct BatchTotals { unsigned long long* agg; unsigned long long* claim; };

// Sum of the M/=/X lengths when the real CIGAR holds nothing but M/=/X and N, else 0.  With an ideal CIGAR of one
// M word of that length merge_cigars skips every N, turns every chunk into M (merge_ops) and both sides end
// together: the merged CIGAR is the ideal word itself.
__device__ __forceinline__ uint32_t real_msum(const uint32_t* __restrict__ real, uint32_t n) {
  uint32_t sum = 0;
  for (uint32_t i = 0; i < n; i++) {
    const uint32_t rw = ld_stream<2>(&real[i]), op = rw & 0xfu;
    if (op == CM || op == CEQ || op == CX) sum += rw >> 4; else if (op != CN) return 0;
  }
  return sum;
}
__device__ __forceinline__ bool merge_trivial(uint32_t iword, uint32_t msum) { return (iword & 0xfu) == CM && msum && msum == (iword >> 4); }

__device__ __forceinline__ bool front_overflow(const Ctl* c) { return c->n_cand > c->cap_cand || c->n_ideal > c->cap_ideal || c->n_tb > c->cap_tb || c->bad_input != 0; }

// ORDERED: rows go to a provisional buffer as packed records and k_place moves them into bundle order afterwards.
// !ORDERED: the claimed place is final -- rows are written straight into the output columns, blocks of 32 alignments in
// completion order, and a_roff / a_rcnt tell the consumer where the rows of each unit are.
// PACKED (with !ORDERED): the same, but a row is ONE 16-byte record (brb_batch_out::packed_rows, include/bramble_b200.h): a one-word
// CIGAR below 2^27 travels inside it, any other CIGAR goes to the pool behind a length word -- the pool then only
// holds the rare multi-word CIGARs and a short-read row costs 16 bytes on the way back to the host instead of 22.
// LONG (long-read mode): small batches, rows with long CIGARs merged and copied by the whole warp.  A separate
// instantiation, so that the short-read kernel keeps its registers (48 at 5 CTAs per SM) for the code it actually runs.
template <bool ORDERED, bool PACKED, bool LONG>
__global__ void __launch_bounds__(256, LONG ? 3 : VAR_BACK_CTAS) k_back(DevBatch b, Ctl* __restrict__ ctl, BackIn v, uint4* __restrict__ gstage, uint32_t* __restrict__ tpool,
                                                 BatchTotals tl, ProvRows pv_, Rows o, Ctl* __restrict__ host_result) {
  __shared__ uint4 s_stage[8][BACK_STAGE];                 // per match of the batch: tid (NONE: not emitted), CIGAR words, partner, scratch offset
  __shared__ uint16_t s_queue[8][BACK_STAGE];              // matches that need the general merge
  __shared__ uint32_t s_own[8][9][32];                     // per alignment of the batch: off, excl, ro, nrc, mro, mnrc, words of merged rows, rsum, mrsum
  const bool skip = front_overflow(ctl);                   // uniform for the whole grid (only the front kernels move these counters)
  if (ORDERED && skip) return;
  const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  uint32_t* s_off = s_own[wid][0]; uint32_t* s_excl = s_own[wid][1]; uint32_t* s_ro = s_own[wid][2]; uint32_t* s_nrc = s_own[wid][3];
  uint32_t* s_mro = s_own[wid][4]; uint32_t* s_mnrc = s_own[wid][5]; uint32_t* s_words = s_own[wid][6];
  uint32_t* s_rsum = s_own[wid][7]; uint32_t* s_mrsum = s_own[wid][8];
  uint16_t* queue = s_queue[wid];
  // alignments per batch: 32 (short reads: lane = alignment), fewer for long reads, whose rows are 10^2..10^4 CIGAR words each --
  // with 32 of them per ticket a bundle has fewer batches than the GPU has warps and the slowest batch is the kernel time
  const uint32_t BS = LONG ? b.back_shift : 5u, BA = 1u << BS;
  const uint32_t n_batches = (b.n_aln + BA - 1) >> BS;
  const uint32_t cap_tmpw = ctl->cap_tmpw, cap_rows = ctl->cap_rows, cap_outw = ctl->cap_outw;
  WarpPool wp; wp.cur = 0; wp.end = 0;
  uint32_t n_pooled = 0;                                    // PACKED: rows of this lane whose CIGAR went to the pool (behind a length word)
  uint32_t wb_next = 0, wb_end = 0;
  for (; !skip;) {
    if (wb_next == wb_end) {                                // batch tickets, kTicketBack batches each
      uint32_t t = 0;
      if (lane == 0) t = atomicAdd(&ctl->ticket_a, kTicketBack);
      wb_next = __shfl_sync(0xffffffffu, t, 0); wb_end = wb_next + kTicketBack;
    }
    const uint32_t wb = wb_next++;
    if (wb >= n_batches) break;
    const uint32_t a = wb * BA + lane;
    const bool in_batch = lane < BA && a < b.n_aln;
    uint32_t off = 0, n = 0, m_off = 0, m_n = 0, ro = 0, nrc = 0, mro = 0, mnrc = 0, rsum = 0, mrsum = 0, pflag = 0, nm = 0;
    int32_t mate = -1;
    bool owner = false;
    if (in_batch) {
      const uint4 ac = ld_stream<2>(&v.a_cand[a]);
      mate = ld_stream<2>(&b.mate_idx[a]);
      ro = ld_stream<2>(&b.cigar_off[a]); nrc = ld_stream<2>(&b.cigar_off[a + 1]) - ro;
      if (ac.z) rsum = real_msum(b.cigar + ro, nrc);          // every lane sums its OWN real CIGAR ...
      owner = (mate < 0 || (uint32_t)mate > a) && ac.z;       // unit owner with at least one match
      if (owner) {
        off = ac.x; n = ac.y;
        if (mate >= 0) {
          const uint4 mc = ld_stream<2>(&v.a_cand[mate]);
          m_off = mc.x; m_n = mc.y; nm = mc.z;
          mro = ld_stream<2>(&b.cigar_off[mate]); mnrc = ld_stream<2>(&b.cigar_off[mate + 1]) - mro;
        }
        pflag = (nm == 0 ? 1u : 0u) | ((ac.z == 1 && nm == 1) ? 2u : 0u);
      }
    }
    {                                                       // ... and an owner takes its mate's from the mate's lane when that is in the batch
      const uint32_t ml = (uint32_t)mate - wb * BA;
      const uint32_t got = __shfl_sync(0xffffffffu, rsum, ml & 31u);
      if (owner && nm) mrsum = ml < BA ? got : real_msum(b.cigar + mro, mnrc);
    }
    const uint32_t incl = warp_incl_scan(n, lane), excl = incl - n, total = __shfl_sync(0xffffffffu, incl, 31);
    const bool staged = total <= BACK_STAGE;
    uint4* stage = staged ? s_stage[wid] : gstage;          // staged: indexed by item; otherwise by slot
    s_off[lane] = off; s_excl[lane] = excl; s_ro[lane] = ro; s_nrc[lane] = nrc; s_mro[lane] = mro; s_mnrc[lane] = mnrc; s_words[lane] = 0; s_rsum[lane] = rsum; s_mrsum[lane] = mrsum;
    __syncwarp();
    uint32_t rows = 0, words = 0, qn = 0;
    // general merge of one emitted row (pair): all lanes call, `go` lanes work
    auto merge_rows = [&](bool go, uint32_t L, uint32_t sidx, uint32_t tid, uint32_t part, const uint4& mr, const uint4& pr) {
      const bool paired = part != SELF;
      const uint32_t l_nrc = go ? s_nrc[L] : 0, l_mnrc = go ? s_mnrc[L] : 0;
      const uint32_t cap_r = l_nrc + (mr.y >> 1) + 2, want = go ? 1 + cap_r + (paired ? l_mnrc + (pr.y >> 1) + 2 : 0) : 0;
      bool fit;
      const uint32_t toff = warp_pool_alloc(wp, &ctl->n_tmpw, cap_tmpw, want, lane, fit);
      if (!go) return;
      uint32_t clr = 0, clm = 0;
      if (fit) {
        // a row of the pair may still be the trivial one-word case
        if ((mr.y >> 1) == 1 && merge_trivial(mr.z, s_rsum[L])) { tpool[toff + 1] = mr.z; clr = 1; }
        else clr = merge_cigars(b.cigar + s_ro[L], l_nrc, v.ipool + mr.z, mr.y >> 1, mr.z, tpool + toff + 1);
        if (paired) {
          if ((pr.y >> 1) == 1 && merge_trivial(pr.z, s_mrsum[L])) { tpool[toff + 1 + cap_r] = pr.z; clm = 1; }
          else clm = merge_cigars(b.cigar + s_mro[L], l_mnrc, v.ipool + pr.z, pr.y >> 1, pr.z, tpool + toff + 1 + cap_r);
        }
        tpool[toff] = clr;
      }
      const uint32_t hdr = PACKED ? (paired ? 2u : 1u) : 0u;     // PACKED: every pooled CIGAR sits behind its length word
      n_pooled += hdr;
      stage[sidx] = make_uint4(tid, clr + clm + hdr, part, fit ? toff : NOFIT);
      atomicAdd(&s_words[L], clr + clm + hdr);
    };
    // ---- phase 1
    for (uint32_t base = 0; base < total; base += 32) {
      const uint32_t item = base + lane;
      const bool actv = item < total;
      const unsigned L = warp_owner(incl, actv ? item : total - 1);
      const uint32_t o_off_ = __shfl_sync(0xffffffffu, off, L);
      const uint32_t c = o_off_ + (item - __shfl_sync(0xffffffffu, excl, L));
      const uint32_t o_pf = __shfl_sync(0xffffffffu, pflag, L);
      const uint32_t o_moff = __shfl_sync(0xffffffffu, m_off, L), o_mn = __shfl_sync(0xffffffffu, m_n, L);
      const uint32_t o_rsum = __shfl_sync(0xffffffffu, rsum, L), o_mrsum = __shfl_sync(0xffffffffu, mrsum, L);
      uint4 mr = make_uint4(0, 0, 0, NONE);
      if (actv) mr = ld_stream<2>(&v.m_rec[c]);
      const uint32_t tid = mr.w;
      uint32_t part = NONE;
      if (tid != NONE) {
        if (o_pf & 1u) part = SELF;
        else {
          // (the matches of an alignment have distinct tids: one guide-exon chain per transcript)
          uint32_t only = NONE;
#if VAR_PARTNER_GUESS
          // both ends usually hit the same transcripts and both lists are in index order: the partner of my k-th slot
          // is most often the mate's k-th slot (tids are distinct inside a list, so a hit is THE partner)
          { const uint32_t k = c - o_off_; if (k < o_mn && __ldg(&v.m_rec[o_moff + k].w) == tid) part = o_moff + k; }
          if (part == NONE)
#endif
#pragma unroll kUnroll
          for (uint32_t d = o_moff, e = o_moff + o_mn; d < e; d++) { const uint32_t t = __ldg(&v.m_rec[d].w); if (t == tid) part = d; if (t != NONE) only = d; }
          if (part == NONE && (o_pf & 2u)) part = only;           // different transcripts
        }
      }
      const bool emitted = part != NONE, paired = emitted && part != SELF;
      uint4 pr = make_uint4(0, 0, 0, NONE);
      if (paired) pr = __ldg(&v.m_rec[part]);
      bool inl = emitted && (mr.y >> 1) == 1 && merge_trivial(mr.z, o_rsum) && (!paired || ((pr.y >> 1) == 1 && merge_trivial(pr.z, o_mrsum)));
      if (PACKED) inl = inl && ((mr.z | (paired ? pr.z : 0u)) >> 27) == 0;      // the word has to fit the record
      const bool slow = emitted && !inl;
      const uint32_t B_e = __ballot_sync(0xffffffffu, emitted), B_p = __ballot_sync(0xffffffffu, paired), B_s = __ballot_sync(0xffffffffu, slow);
      if (n && incl > base && excl < base + 32) {           // rows / one-word CIGARs of MY alignment in this chunk
        const uint32_t mine = bit_range(excl > base ? excl - base : 0, incl - base < 32 ? incl - base : 32);
        rows += __popc(B_e & mine) + __popc(B_p & mine);
        if (!PACKED) words += __popc(B_e & ~B_s & mine) + __popc(B_p & ~B_s & mine);
      }
      const uint32_t sidx = staged ? item : c;
      if (LONG) {                                             // the merges are done (k_merge_long): a slow row only needs their sizes
        uint32_t wds = PACKED ? 0u : (paired ? 2u : 1u), where = INLINE;
        if (slow) {
          const uint2 gr = __ldg(&v.mg[c]), gm = paired ? __ldg(&v.mg[part]) : make_uint2(NONE, 0u);
          const bool fit = gr.x != NOFIT && gm.x != NOFIT;
          const uint32_t hdr = PACKED ? (paired ? 2u : 1u) : 0u;
          wds = fit ? gr.y + gm.y + hdr : 0u; where = fit ? PREMERGED : NOFIT;
          n_pooled += hdr;
          atomicAdd(&s_words[L], wds);
        }
        if (actv) stage[sidx] = make_uint4(emitted ? tid : NONE, wds, part, where);
        continue;
      }
      if (actv) stage[sidx] = make_uint4(emitted ? tid : NONE, slow ? L : (PACKED ? 0u : (paired ? 2u : 1u)), part, slow ? PENDING : INLINE);
      if (B_s) {
        if (staged) { if (slow) queue[qn + __popc(B_s & lt_mask)] = (uint16_t)item; }
        qn += __popc(B_s);
      }
    }
    __syncwarp();
    // ---- phase 1b (short reads): the general merges, 32 at a time (staged: from the queue; otherwise a second walk over the items)
    if (!LONG) for (uint32_t qi = 0, qe = qn ? (staged ? qn : total) : 0; qi < qe; qi += 32) {
      bool go = qi + lane < qe;
      uint32_t sidx = 0, L = 0, c = 0;
      if (staged) { if (go) { sidx = queue[qi + lane]; L = stage[sidx].y; c = s_off[L] + (sidx - s_excl[L]); } }
      else {
        L = warp_owner(incl, go ? qi + lane : total - 1);
        c = s_off[L] + (qi + lane - s_excl[L]); sidx = c;
        go = go && stage[sidx].w == PENDING;
      }
      uint32_t tid = NONE, part = SELF; uint4 mr = make_uint4(0, 0, 0, NONE), pr = mr;
      if (go) {
        const uint4 sg = stage[sidx];
        tid = sg.x; part = sg.z;
        mr = __ldg(&v.m_rec[c]);
        if (part != SELF) pr = __ldg(&v.m_rec[part]);
      }
      merge_rows(go, L, sidx, tid, part, mr, pr);
    }
    __syncwarp();
    words += s_words[lane];
    // ---- row / word room of the batch: claimed in completion order (no batch ever waits for another one);
    //      k_place moves the rows to their place in bundle order once every batch total is known
    const uint32_t irows = warp_incl_scan(rows, lane), iwords = warp_incl_scan(words, lane);
    const uint32_t R_b = __shfl_sync(0xffffffffu, irows, 31), W_b = __shfl_sync(0xffffffffu, iwords, 31);
    const unsigned long long agg = (unsigned long long)R_b | ((unsigned long long)W_b << 32);
    unsigned long long claim = 0;
    if (lane == 0) { claim = atomicAdd(&ctl->claimed, agg); if (ORDERED) { tl.agg[wb] = agg; tl.claim[wb] = claim; } }
    claim = __shfl_sync(0xffffffffu, claim, 0);
    const uint32_t prow0 = (uint32_t)claim, pword0 = (uint32_t)(claim >> 32);
    const uint32_t row0 = prow0 + irows - rows, word0 = pword0 + iwords - words;
    if (PACKED && !LONG && b.compact_rows) {                  // one row offset per batch + 16-bit unit sizes (units follow each other in alignment order)
      if (in_batch) reinterpret_cast<uint16_t*>(o.a_rcnt)[a] = (uint16_t)rows;
      if (lane == 0) o.a_roff[wb] = prow0;
      if (__any_sync(0xffffffffu, rows > 0xffffu)) { if (lane == 0) atomicMax(&ctl->bad_input, 6u); __syncwarp(); continue; }
    } else
    if (in_batch) { o.a_roff[a] = ORDERED ? irows - rows : row0; o.a_rcnt[a] = rows; }     // ORDERED: batch-relative, k_place adds the first row of the batch
    if (ORDERED && W_b >= (1u << 28)) { if (lane == 0) atomicMax(&ctl->bad_input, 3u); __syncwarp(); continue; }
    if (PACKED && (unsigned long long)pword0 + W_b >= (1u << 27)) { if (lane == 0) atomicMax(&ctl->bad_input, 4u); __syncwarp(); continue; }
    if ((unsigned long long)prow0 + R_b > cap_rows || (unsigned long long)pword0 + W_b > cap_outw) { __syncwarp(); continue; }   // the host grows the row pools and replays
    // ---- phase 2: rank inside the unit, rows, CIGAR words
    for (uint32_t base = 0; base < total; base += 32) {
      const uint32_t item = base + lane;
      const bool actv = item < total;
      const unsigned L = warp_owner(incl, actv ? item : total - 1);
      const uint32_t o_off = __shfl_sync(0xffffffffu, off, L), o_n = __shfl_sync(0xffffffffu, n, L), o_ex = __shfl_sync(0xffffffffu, excl, L);
      const uint32_t o_row0 = __shfl_sync(0xffffffffu, row0, L), o_word0 = __shfl_sync(0xffffffffu, word0, L);
      const int32_t o_mate = __shfl_sync(0xffffffffu, mate, L);
      if (!LONG) {                                            // short reads: every lane writes its own row(s) and CIGAR words
        if (!actv) continue;
        const uint32_t c = o_off + (item - o_ex);
        const uint32_t s0 = staged ? o_ex : o_off, sidx = staged ? item : c;
        const uint4 sg = stage[sidx];
        const uint32_t tid = sg.x;
        if (tid == NONE || sg.w == NOFIT) continue;
        // rank among the emitted matches of the read: ascending tid, slot order among equals; CIGAR words before me
        uint32_t rank = 0, wpre = 0;
#pragma unroll kUnroll
        for (uint32_t k = 0; k < o_n; k++) {
          const uint2 e = *reinterpret_cast<const uint2*>(&stage[s0 + k]);
          if (e.x < tid || (e.x == tid && s0 + k < sidx)) { rank++; wpre += e.y; }        // (NONE is never below a tid)
        }
        const uint32_t part = sg.z;
        const bool paired = part != SELF, inl = sg.w == INLINE;
        const uint4 mr = __ldg(&v.m_rec[c]);
        const uint32_t row = o_row0 + (paired ? 2 * rank : rank), w0 = o_word0 + wpre;
        const uint32_t clr = inl ? 1u : tpool[sg.w];
        uint4 m = mr; uint32_t cc = c, aln = wb * BA + L, fl = BRB_REC_FIRST, wpos = w0, len = clr;
        const uint32_t* src = tpool + sg.w + 1;
        uint4 pm = make_uint4(0, 0, 0, NONE);
        if (paired) { pm = __ldg(&v.m_rec[part]); fl |= BRB_REC_PAIRED | (pm.w == tid ? BRB_REC_SAMETX : 0); }
        for (uint32_t e = 0, ne = paired ? 2 : 1; e < ne; e++) {
          if (e) {                                            // the mate row follows the read row
            m = pm; cc = part; aln = (uint32_t)o_mate; fl &= ~BRB_REC_FIRST; wpos = w0 + clr;
            src = tpool + sg.w + 1 + (s_nrc[L] + (mr.y >> 1) + 2); len = inl ? 1u : sg.y - clr - (PACKED ? 2u : 0u);
            if (PACKED) wpos = w0 + 1 + clr;
          }
          const uint32_t minus = m.y & 1u, r = row + e;
          uint32_t* pool = ORDERED ? pv_.pool : o.pool;
          if (ORDERED) {
            pv_.row[r] = make_uint4(aln, m.w, m.x, (wpos - pword0) | (minus << 28) | (fl << 29));
            if (v.m_sim) { pv_.sim[r] = v.m_sim[cc]; pv_.clip[r] = v.m_clip[cc]; }
          } else if (PACKED) {
            const uint32_t w = (inl ? (m.z | BRB_ROW_INLINE) : wpos) | (minus << 28) | (fl << 29);
            if (b.compact_rows == 2) {                        // 8 bytes: tid 22 | pos 24 | flags 3 | '-' 1 | inline 1 | 13 bits: M length (inline) or words of the pool entry
              const uint32_t cl = inl ? (m.z >> 4) : len;
              if ((m.w >> 22) | (m.x >> 24) | (cl >> 13)) atomicMax(&ctl->bad_input, 7u);
              st_stream(reinterpret_cast<unsigned long long*>(pv_.row) + r,
                        (unsigned long long)m.w | ((unsigned long long)m.x << 22) | ((unsigned long long)fl << 46) | ((unsigned long long)minus << 49) |
                        ((unsigned long long)(inl ? 1u : 0u) << 50) | ((unsigned long long)cl << 51));
            } else
            if (b.compact_rows) { uint32_t* q = reinterpret_cast<uint32_t*>(pv_.row) + 3 * (size_t)r; st_stream(q, m.w); st_stream(q + 1, m.x); st_stream(q + 2, w); }
            else st_stream(&pv_.row[r], make_uint4(aln, m.w, m.x, w));
            if (v.m_sim) { o.sim[r] = v.m_sim[cc]; o.clip[r] = v.m_clip[cc]; }
            if (!inl) pool[wpos++] = len;                     // length word, then the CIGAR
          } else {
            st_stream(&o.src[r], aln); st_stream(&o.tid[r], m.w); st_stream(&o.pos[r], m.x); st_stream(&o.strand[r], (int8_t)(minus ? '-' : '+'));
            st_stream(&o.flags[r], (uint8_t)fl); st_stream(&o.coff[r], wpos);
            if (v.m_sim) { o.sim[r] = v.m_sim[cc]; o.clip[r] = v.m_clip[cc]; }
          }
          if (inl) { if (!PACKED) st_stream(&pool[wpos], m.z); }
          else if (minus) for (uint32_t i = 0; i < len; i++) pool[wpos + i] = src[len - 1 - i];      // reverse_complement_bam, src/bam.cpp:636-702
          else for (uint32_t i = 0; i < len; i++) pool[wpos + i] = src[i];
        }
        continue;
      }
      // long reads: the same, with the CIGAR copies of long rows done by the whole warp
      const uint32_t c = o_off + (item - o_ex);
      const uint32_t s0 = staged ? o_ex : o_off, sidx = staged ? item : c;
      uint4 sg = make_uint4(NONE, 0, 0, NOFIT);
      if (actv) sg = stage[sidx];
      const uint32_t tid = sg.x;
      // CIGAR copies of this lane's row(s), done after the row columns: short ones by the lane, long ones by the whole warp
      const uint32_t* j_src[2] = {nullptr, nullptr}; uint32_t j_dst[2] = {0, 0}, j_len[2] = {0, 0}, j_minus = 0;
      uint32_t* pool = ORDERED ? pv_.pool : o.pool;
      if (tid != NONE && sg.w != NOFIT) {
        // rank among the emitted matches of the read: ascending tid, slot order among equals; CIGAR words before me
        uint32_t rank = 0, wpre = 0;
#pragma unroll kUnroll
        for (uint32_t k = 0; k < o_n; k++) {
          const uint2 e = *reinterpret_cast<const uint2*>(&stage[s0 + k]);
          if (e.x < tid || (e.x == tid && s0 + k < sidx)) { rank++; wpre += e.y; }        // (NONE is never below a tid)
        }
        const uint32_t part = sg.z;
        const bool paired = part != SELF, inl = sg.w == INLINE;
        const uint4 mr = __ldg(&v.m_rec[c]);
        const uint32_t row = o_row0 + (paired ? 2 * rank : rank), w0 = o_word0 + wpre;
        // a merged CIGAR made by k_merge_long sits in the scratch pool; the one-word case is the match record's own word
        const uint2 gr = inl ? make_uint2(NONE, 1u) : __ldg(&v.mg[c]);
        const uint2 gm = (inl || !paired) ? make_uint2(NONE, 1u) : __ldg(&v.mg[part]);
        const uint32_t clr = gr.y;
        uint4 m = mr; uint32_t cc = c, aln = wb * BA + L, fl = BRB_REC_FIRST, wpos = w0, len = clr;
        const uint32_t* src = gr.x == NONE ? &v.m_rec[c].z : tpool + gr.x;
        uint4 pm = make_uint4(0, 0, 0, NONE);
        if (paired) { pm = __ldg(&v.m_rec[part]); fl |= BRB_REC_PAIRED | (pm.w == tid ? BRB_REC_SAMETX : 0); }
        for (uint32_t e = 0, ne = paired ? 2 : 1; e < ne; e++) {
          if (e) {                                            // the mate row follows the read row
            m = pm; cc = part; aln = (uint32_t)o_mate; fl &= ~BRB_REC_FIRST; wpos = w0 + clr;
            src = gm.x == NONE ? &v.m_rec[part].z : tpool + gm.x; len = gm.y;
            if (PACKED) wpos = w0 + 1 + clr;
          }
          const uint32_t minus = m.y & 1u, r = row + e;
          if (ORDERED) {
            pv_.row[r] = make_uint4(aln, m.w, m.x, (wpos - pword0) | (minus << 28) | (fl << 29));
            if (v.m_sim) { pv_.sim[r] = v.m_sim[cc]; pv_.clip[r] = v.m_clip[cc]; }
          } else if (PACKED) {
            st_stream(&pv_.row[r], make_uint4(aln, m.w, m.x, (inl ? (m.z | BRB_ROW_INLINE) : wpos) | (minus << 28) | (fl << 29)));
            if (v.m_sim) { o.sim[r] = v.m_sim[cc]; o.clip[r] = v.m_clip[cc]; }
            if (!inl) pool[wpos++] = len;                     // length word, then the CIGAR
          } else {
            st_stream(&o.src[r], aln); st_stream(&o.tid[r], m.w); st_stream(&o.pos[r], m.x); st_stream(&o.strand[r], (int8_t)(minus ? '-' : '+'));
            st_stream(&o.flags[r], (uint8_t)fl); st_stream(&o.coff[r], wpos);
            if (v.m_sim) { o.sim[r] = v.m_sim[cc]; o.clip[r] = v.m_clip[cc]; }
          }
          if (inl) { if (!PACKED) st_stream(&pool[wpos], m.z); }
          else { j_src[e] = src; j_dst[e] = wpos; j_len[e] = len; j_minus |= minus << e; }
        }
      }
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const uint32_t len = j_len[e]; const bool minus = (j_minus >> e) & 1u;
        const bool wide = LONG && len >= 32;                  // (reversed for '-' transcripts: reverse_complement_bam, src/bam.cpp:636-702)
        if (len && !wide) { if (minus) for (uint32_t i = 0; i < len; i++) pool[j_dst[e] + i] = j_src[e][len - 1 - i]; else for (uint32_t i = 0; i < len; i++) pool[j_dst[e] + i] = j_src[e][i]; }
        if (LONG) for (uint32_t Bw = __ballot_sync(0xffffffffu, wide); Bw; Bw &= Bw - 1) {
          const int sl = __ffs(Bw) - 1;
          const uint32_t* bs = reinterpret_cast<const uint32_t*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(j_src[e]), sl));
          const uint32_t bd = __shfl_sync(0xffffffffu, j_dst[e], sl), bl = __shfl_sync(0xffffffffu, len, sl);
          const bool bm = __shfl_sync(0xffffffffu, (uint32_t)minus, sl);
          if (bm) for (uint32_t i = lane; i < bl; i += 32) pool[bd + i] = bs[bl - 1 - i]; else for (uint32_t i = lane; i < bl; i += 32) pool[bd + i] = bs[i];
        }
      }
        }
    __syncwarp();                                           // the staging area is reused by the next batch
  }
  if (!ORDERED) {                                           // last kernel of the sequence: the last CTA to leave publishes the result
    if (PACKED) {                                           // (out_words is free until the publisher turns it into the CIGAR word count)
      const uint32_t np = __reduce_add_sync(0xffffffffu, n_pooled);
      if (lane == 0 && np) atomicAdd(&ctl->out_words, (unsigned long long)np);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(&ctl->ticket_b, 1u) == gridDim.x - 1) {
        __threadfence();
        if (!skip) {                                        // the totals claimed are the result sizes
          const unsigned long long tot = atomicAdd(&ctl->claimed, 0ull);
          const uint32_t R = (uint32_t)tot, W = (uint32_t)(tot >> 32);
          ctl->n_rows = R; ctl->n_outw = W;
          // the work counter is the CIGAR word count of the rows whatever the layout: a pooled row has one length word
          // too many in W, a row with its word inside the record has none at all
          ctl->out_words = PACKED ? (unsigned long long)W + R - 2ull * atomicAdd(&ctl->out_words, 0ull) : (unsigned long long)W;
          if (!PACKED && R <= cap_rows && W <= cap_outw) o.coff[R] = W;
        }
        ctl_publish_and_reset(ctl, host_result);
      }
    }
  }
}

// Clip-rescue path only: the front kernel cannot know which candidates survive k_ksw / k_build, so the match
// count of each alignment (a_cand.z) is filled in afterwards.
__global__ void __launch_bounds__(256) k_count_valid(DevBatch b, const Ctl* __restrict__ ctl, uint4* __restrict__ a_cand, const uint4* __restrict__ m_rec) {
  if (front_overflow(ctl)) return;
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= b.n_aln) return;
  uint4 ac = a_cand[a]; uint32_t nv = 0;
  for (uint32_t c = ac.x, e = ac.x + ac.y; c < e; c++) nv += m_rec[c].w != NONE;
  ac.z = nv; a_cand[a] = ac;
}

// Long reads: the merged CIGAR of every match, BEFORE pairing.  A row's CIGAR depends on its alignment (real CIGAR) and its transcript
// (ideal CIGAR) only, and a long-read bundle emits practically every match, so the 10^2..10^4-word merges do not have to sit inside the
// pairing kernel: here they run one warp per match (merge_cigars_warp, k_merge.cuh; short ones by lane 0) with their own occupancy, and
// k_back<.., LONG> only copies.  The one-word case (ideal = one M of the real CIGAR's M/=/X sum) is recognised and left alone, as in k_back.
__global__ void __launch_bounds__(256, 4) k_merge_long(DevBatch b, Ctl* __restrict__ ctl, const uint32_t* __restrict__ c_task, BackIn v,
                                                       uint32_t* __restrict__ tpool, uint2* __restrict__ mg) {
  if (front_overflow(ctl)) return;
  const unsigned lane = threadIdx.x & 31;
  const uint32_t n_cand = ctl->n_cand, cap_tmpw = ctl->cap_tmpw;
  WarpPool wp; wp.cur = 0; wp.end = 0;
  for (;;) {
    uint32_t c0 = 0;
    if (lane == 0) c0 = atomicAdd(&ctl->ticket_m, 32u);
    c0 = __shfl_sync(0xffffffffu, c0, 0);
    if (c0 >= n_cand) break;
    const uint32_t c = c0 + lane;
    uint4 mr = make_uint4(0, 0, 0, NONE);
    if (c < n_cand) mr = __ldg(&v.m_rec[c]);
    uint32_t ro = 0, nR = 0, rep = c;
    if (mr.w != NONE) {
      const uint32_t a = __ldg(&c_task[c]) >> 1; ro = __ldg(&b.cigar_off[a]); nR = __ldg(&b.cigar_off[a + 1]) - ro;
      // An earlier match of the SAME alignment with the same ideal CIGAR (isoforms that share the exons under the read) has the same
      // merged CIGAR: the merge runs once (for the earliest such match among this warp's 32 slots) and the others point at its result.
      const uint32_t nI = mr.y >> 1, first = __ldg(&v.a_cand[a].x);
      const uint32_t lo = first > c0 ? first : c0;             // (the SAME window start for every slot of the alignment: the earliest equal slot is then nobody's duplicate)
      for (uint32_t c2 = lo; c2 < c; c2++) {
        const uint4 m2 = __ldg(&v.m_rec[c2]);
        if (m2.w == NONE || (m2.y >> 1) != nI || (__ldg(&c_task[c2]) >> 1) != a) continue;
        bool same = nI == 1 ? m2.z == mr.z : true;
        if (nI != 1) for (uint32_t i = 0; i < nI && same; i++) same = __ldg(&v.ipool[m2.z + i]) == __ldg(&v.ipool[mr.z + i]);
        if (same) { rep = c2; break; }
      }
    }
    const bool dup = rep != c;
    // short CIGARs: every lane merges its own match
    const bool small = mr.w != NONE && !dup && nR + (mr.y >> 1) < MERGE_COOP_MIN;
    {
      bool triv = false;
      if (small && (mr.y >> 1) == 1) triv = merge_trivial(mr.z, real_msum(b.cigar + ro, nR));
      bool fit;
      const uint32_t toff = warp_pool_alloc(wp, &ctl->n_tmpw, cap_tmpw, (small && !triv) ? nR + (mr.y >> 1) + 2 : 0u, lane, fit);
      if (small) {
        if (triv) mg[c] = make_uint2(NONE, 1u);
        else if (!fit) mg[c] = make_uint2(NOFIT, 0u);
        else mg[c] = make_uint2(toff, merge_cigars(b.cigar + ro, nR, v.ipool + mr.z, mr.y >> 1, mr.z, tpool + toff));
      }
    }
    for (uint32_t Bv = __ballot_sync(0xffffffffu, mr.w != NONE && !dup && !small); Bv; Bv &= Bv - 1) {        // long ones: one match at a time, by the whole warp
      const int src = __ffs(Bv) - 1;
      const uint32_t cc = c0 + (uint32_t)src, nI = __shfl_sync(0xffffffffu, mr.y, src) >> 1, z = __shfl_sync(0xffffffffu, mr.z, src);
      const uint32_t r0 = __shfl_sync(0xffffffffu, ro, src), nr = __shfl_sync(0xffffffffu, nR, src);
      const uint32_t* real = b.cigar + r0;
      if (nI == 1 && (z & 0xfu) == CM) {                       // merge_trivial: M/=/X sum of an M/N-only real CIGAR == the ideal M
        uint32_t sum = 0; bool other = false;
        for (uint32_t i = lane; i < nr; i += 32) { const uint32_t rw = __ldg(&real[i]), op = rw & 0xfu; if (op == CM || op == CEQ || op == CX) sum += rw >> 4; else if (op != CN) other = true; }
        sum = __reduce_add_sync(0xffffffffu, sum);
        if (!__any_sync(0xffffffffu, other) && sum && sum == (z >> 4)) { if (lane == 0) mg[cc] = make_uint2(NONE, 1u); continue; }
      }
      const uint32_t cap = nr + nI + 2;
      bool fit;
      uint32_t toff = warp_pool_alloc(wp, &ctl->n_tmpw, cap_tmpw, lane == 0 ? cap + merge_warp_scratch(nr, nI) : 0u, lane, fit);
      toff = __shfl_sync(0xffffffffu, toff, 0);
      if (!fit) { if (lane == 0) mg[cc] = make_uint2(NOFIT, 0u); continue; }
      const uint32_t len = merge_cigars_any_warp(real, nr, v.ipool + z, nI, z, tpool + toff, tpool + toff + cap, lane);
      if (lane == 0) mg[cc] = make_uint2(toff, len);
      __syncwarp();
    }
    __syncwarp();                                             // the representatives' results are in mg
    if (dup) mg[c] = mg[rep];
  }
}

// BRB_OPT_ORDERED_ROWS: the sequence ends with k_place / k_group_rows; one thread publishes after them.
__global__ void k_publish(Ctl* __restrict__ ctl, Ctl* __restrict__ host_result) { ctl_publish_and_reset(ctl, host_result); }

// Rows of batch wb: from where k_back claimed room (tl.claim) to their place in bundle order (t_off = exclusive
// scan of tl.agg; t_off[n_batches] = totals).  One warp per batch.
__global__ void __launch_bounds__(256, VAR_PLACE_CTAS) k_place(DevBatch b, Ctl* __restrict__ ctl, BatchTotals tl, const unsigned long long* __restrict__ t_off,
                                               ProvRows pv, Rows o) {
  if (front_overflow(ctl)) return;
  const uint32_t BS = b.back_shift, BA = 1u << BS;
  const uint32_t n_batches = (b.n_aln + BA - 1) >> BS;
  const unsigned long long tot = t_off[n_batches];
  const uint32_t R = (uint32_t)tot, W = (uint32_t)(tot >> 32);
  if (blockIdx.x == 0 && threadIdx.x == 0) { ctl->n_rows = R; ctl->n_outw = W; ctl->out_words = W; }
  if (R > ctl->cap_rows || W > ctl->cap_outw) return;       // k_back skipped what did not fit; the host grows the pools and replays
  if (blockIdx.x == 0 && threadIdx.x == 0) { o.coff[R] = W; o.a_roff[b.n_aln] = R; }
  const unsigned lane = threadIdx.x & 31;
  const uint32_t wstride = gridDim.x * (blockDim.x >> 5);
  for (uint32_t wb = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); wb < n_batches; wb += wstride) {
    const unsigned long long agg = tl.agg[wb], src = tl.claim[wb], dst = t_off[wb];
    const uint32_t R_b = (uint32_t)agg, W_b = (uint32_t)(agg >> 32);
    const uint32_t srow = (uint32_t)src, sword = (uint32_t)(src >> 32), drow = (uint32_t)dst, dword = (uint32_t)(dst >> 32);
    const uint32_t a = wb * BA + lane;
    if (lane < BA && a < b.n_aln) o.a_roff[a] += drow;
    for (uint32_t r0 = 0; r0 < R_b; r0 += 32 * VAR_PLACE_ROWS) {            // several rows per lane in flight
      uint4 x[VAR_PLACE_ROWS];
#pragma unroll
      for (int k = 0; k < VAR_PLACE_ROWS; k++) { const uint32_t r = r0 + k * 32 + lane; if (r < R_b) x[k] = __ldcs(&pv.row[srow + r]); }
#pragma unroll
      for (int k = 0; k < VAR_PLACE_ROWS; k++) {
        const uint32_t r = r0 + k * 32 + lane;
        if (r < R_b) {
          o.src[drow + r] = x[k].x; o.tid[drow + r] = x[k].y; o.pos[drow + r] = x[k].z; o.coff[drow + r] = dword + (x[k].w & 0x0fffffffu);
          o.strand[drow + r] = (x[k].w >> 28) & 1u ? '-' : '+'; o.flags[drow + r] = (uint8_t)(x[k].w >> 29);
          if (pv.sim) { o.sim[drow + r] = pv.sim[srow + r]; o.clip[drow + r] = pv.clip[srow + r]; }
        }
      }
    }
    for (uint32_t w0 = 0; w0 < W_b; w0 += 128) {
      uint32_t y[4];
#pragma unroll
      for (int k = 0; k < 4; k++) { const uint32_t w = w0 + k * 32 + lane; if (w < W_b) y[k] = __ldcs(&pv.pool[sword + w]); }
#pragma unroll
      for (int k = 0; k < 4; k++) { const uint32_t w = w0 + k * 32 + lane; if (w < W_b) o.pool[dword + w] = y[k]; }
    }
  }
}

__global__ void k_group_rows(DevBatch b, const Ctl* __restrict__ ctl, const uint32_t* __restrict__ a_roff, uint32_t* __restrict__ g_roff) {
  if (front_overflow(ctl)) return;
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g <= b.n_groups) g_roff[g] = a_roff[g < b.n_groups ? b.group_off[g] : b.n_aln];
}

}  // namespace brb
