// Kernel group 5: the output BAM records, assembled on the device from the rows of the resident result and the bundle's
// input records (reference write_to_bam src/core.cpp:96-212; update_cigar / set_mate_info / reverse_complement_bam / set_*_tag
// src/bam.cpp:475-702).  The uncompressed record stream never exists on the host: k_bgzf_deflate (k_deflate.cuh) takes it from
// HBM and only BGZF blocks go back.
//
//   k_emit_plan    thread per input alignment: record geometry (name, CIGAR, sequence, aux area) checked against block_size, and
//                  ONE walk over the aux tags for what the reference edits per output record: the first NH, the first XS (short
//                  reads) or ts (long reads), the first AS (long reads: its value feeds the new AS) and the first HI are dropped,
//                  NH [AS] HI are appended as 'i' tags.  The reference edits tag by tag (bam_aux_get + bam_aux_del + append); on
//                  a well-formed or a malformed aux area that is the same as one walk, because a walk stops at the first tag it
//                  cannot skip, whichever tag it is looking for.
//   k_emit_sizes   thread per output unit (one row, or a read row + its mate row): bytes of its records
//   k_emit_write   warp per output unit: block_size, the 32 fixed bytes with tid / pos / mapq / n_cigar / flag / mate fields replaced,
//                  name, the merged CIGAR of the row, sequence + qualities (reverse-complemented for '-' transcripts), the kept
//                  aux bytes and the three appended tags
#pragma once
#include "dev_common.cuh"

namespace brb {

struct EmitPlan {                 // per input alignment (16 words)
  uint32_t l_name, n_cigar_in; int32_t l_seq; uint32_t aux_off;        // aux_off: from the start of the record body
  uint32_t aux_len, n_del; int32_t as_val; uint32_t ok;
  uint32_t del_s[4], del_e[4];                                         // dropped [s, e) ranges of the aux area, ascending
};
static_assert(sizeof(EmitPlan) == 64, "EmitPlan is 16 words");

struct EmitIn {
  const uint8_t* raw; uint64_t raw_bytes; const uint64_t* rec_off; const uint32_t* rec_len; uint32_t n_aln;
  uint32_t n_units; const uint4* units;                                // .x first row  .y HI of its first record  .z NH  .w mapq | primary << 8
  uint32_t n_rows; const uint32_t* src; const uint32_t* tid; const uint32_t* pos; const int8_t* strand; const uint8_t* flags;
  const double* sim; const int32_t* clip; const uint32_t* coff; const uint32_t* pool;
  uint32_t long_reads;
};

__device__ __forceinline__ uint32_t emit_type_size(uint32_t t) {
  switch (t) { case 'A': case 'c': case 'C': return 1; case 's': case 'S': return 2; case 'i': case 'I': case 'f': return 4; case 'd': return 8; default: return 0; }
}
// end of the value whose TYPE byte is at s (offsets inside [0, end)), or NONE when it cannot be skipped
__device__ __forceinline__ uint32_t emit_aux_skip(const uint8_t* a, uint32_t s, uint32_t end) {
  if (s >= end) return NONE;
  const uint32_t t = a[s++];
  if (t == 'Z' || t == 'H') { while (s < end && a[s]) s++; return s < end ? s + 1 : NONE; }
  if (t == 'B') {
    if (end - s < 5) return NONE;
    const uint32_t sz = emit_type_size(a[s]);
    const uint32_t n = (uint32_t)a[s + 1] | ((uint32_t)a[s + 2] << 8) | ((uint32_t)a[s + 3] << 16) | ((uint32_t)a[s + 4] << 24);
    s += 5;
    if (!sz || (unsigned long long)(end - s) < (unsigned long long)sz * n) return NONE;
    return s + sz * n;
  }
  const uint32_t sz = emit_type_size(t);
  if (!sz || end - s < sz) return NONE;
  return s + sz;
}

__global__ void __launch_bounds__(128) k_emit_plan(EmitIn in, EmitPlan* __restrict__ plan, uint32_t* __restrict__ err) {
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= in.n_aln) return;
  EmitPlan p; memset(&p, 0, sizeof p);
  const uint64_t off = in.rec_off[a]; const uint32_t len = in.rec_len[a];
  if (len < 32 || off > in.raw_bytes || in.raw_bytes - off < len) { atomicMax(err, 1u); plan[a] = p; return; }
  const uint8_t* b = in.raw + off;
  p.l_name = b[8]; p.n_cigar_in = (uint32_t)b[12] | ((uint32_t)b[13] << 8);
  p.l_seq = (int32_t)((uint32_t)b[16] | ((uint32_t)b[17] << 8) | ((uint32_t)b[18] << 16) | ((uint32_t)b[19] << 24));
  const unsigned long long fixed = 32ull + p.l_name + 4ull * p.n_cigar_in + (p.l_seq < 0 ? 0ull : (unsigned long long)((p.l_seq + 1ll) / 2) + (unsigned long long)p.l_seq);
  if (p.l_seq < 0 || fixed > len || p.l_name == 0) { atomicMax(err, 1u); plan[a] = p; return; }
  p.aux_off = (uint32_t)fixed; p.aux_len = len - p.aux_off; p.ok = 1;
  const uint8_t* x = b + p.aux_off; const uint32_t end = p.aux_len;
  const uint8_t t3a = in.long_reads ? 't' : 'X', t3b = in.long_reads ? 's' : 'S';
  uint32_t seen = 0;                                             // bit 0 NH, 1 XS|ts, 2 AS, 3 HI: first occurrence handled
  for (uint32_t s = 0; end - s >= 3 && s < end;) {
    const uint32_t e = emit_aux_skip(x, s + 2, end);
    int which = -1;
    if (x[s] == 'N' && x[s + 1] == 'H') which = 0; else if (x[s] == t3a && x[s + 1] == t3b) which = 1;
    else if (in.long_reads && x[s] == 'A' && x[s + 1] == 'S') which = 2; else if (x[s] == 'H' && x[s + 1] == 'I') which = 3;
    if (which >= 0 && !((seen >> which) & 1u)) {
      seen |= 1u << which;
      if (which == 2) {                                          // bam_aux2i of the first AS, readable or not
        const uint32_t t = x[s + 2]; const uint32_t v = s + 3;
        int32_t val = 0;
        const uint32_t need = emit_type_size(t);
        if (v + need <= end) {
          if (t == 'c') val = (int8_t)x[v]; else if (t == 'C') val = x[v];
          else if (t == 's') val = (int16_t)((uint32_t)x[v] | ((uint32_t)x[v + 1] << 8)); else if (t == 'S') val = (int32_t)((uint32_t)x[v] | ((uint32_t)x[v + 1] << 8));
          else if (t == 'i' || t == 'I') val = (int32_t)((uint32_t)x[v] | ((uint32_t)x[v + 1] << 8) | ((uint32_t)x[v + 2] << 16) | ((uint32_t)x[v + 3] << 24));
        }
        p.as_val = val;
      }
      if (e != NONE) { p.del_s[p.n_del] = s; p.del_e[p.n_del] = e; p.n_del++; }
    }
    if (e == NONE) break;
    s = e;
  }
  plan[a] = p;
}

__device__ __forceinline__ uint32_t emit_aux_out_len(const EmitPlan& p, bool lng) {
  uint32_t n = p.aux_len;
  for (uint32_t k = 0; k < p.n_del; k++) n -= p.del_e[k] - p.del_s[k];
  return n + (lng ? 21u : 14u);
}
__device__ __forceinline__ uint32_t emit_record_bytes(const EmitIn& in, const EmitPlan& p, uint32_t row) {
  const uint32_t nc = in.coff[row + 1] - in.coff[row];
  return 4u + 32u + p.l_name + 4u * nc + (uint32_t)((p.l_seq + 1) / 2) + (uint32_t)p.l_seq + emit_aux_out_len(p, in.long_reads != 0);
}

__global__ void __launch_bounds__(256) k_emit_sizes(EmitIn in, const EmitPlan* __restrict__ plan, unsigned long long* __restrict__ sizes, uint32_t* __restrict__ err) {
  const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= in.n_units) return;
  const uint32_t row = in.units[u].x;
  unsigned long long n = 0;
  if (row >= in.n_rows) { atomicMax(err, 2u); sizes[u] = 0; return; }
  const bool paired = in.flags[row] & BRB_REC_PAIRED;
  if (paired && row + 1 >= in.n_rows) { atomicMax(err, 2u); sizes[u] = 0; return; }
  for (uint32_t e = 0; e < (paired ? 2u : 1u); e++) {
    const uint32_t a = in.src[row + e];
    if (a >= in.n_aln || !plan[a].ok) { atomicMax(err, 2u); sizes[u] = 0; return; }
    n += emit_record_bytes(in, plan[a], row + e);
  }
  sizes[u] = n;
}

__device__ __forceinline__ void emit_put16(uint8_t* d, uint32_t v) { d[0] = (uint8_t)v; d[1] = (uint8_t)(v >> 8); }
__device__ __forceinline__ void emit_put32(uint8_t* d, uint32_t v) { d[0] = (uint8_t)v; d[1] = (uint8_t)(v >> 8); d[2] = (uint8_t)(v >> 16); d[3] = (uint8_t)(v >> 24); }

__global__ void __launch_bounds__(256) k_emit_write(EmitIn in, const EmitPlan* __restrict__ plan, const unsigned long long* __restrict__ offs, uint8_t* __restrict__ out) {
  const unsigned lane = threadIdx.x & 31;
  const uint32_t wstride = gridDim.x * (blockDim.x >> 5);
  for (uint32_t u = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); u < in.n_units; u += wstride) {
    const uint4 un = in.units[u];
    const uint32_t row0 = un.x;
    const bool paired = in.flags[row0] & BRB_REC_PAIRED, same_tx = in.flags[row0] & BRB_REC_SAMETX;
    const bool primary = (un.w >> 8) & 1u, lng = in.long_reads != 0;
    uint8_t* d = out + offs[u];
    for (uint32_t e = 0; e < (paired ? 2u : 1u); e++) {
      const uint32_t r = row0 + e, other = row0 + (1 - e);
      const uint32_t a = in.src[r];
      const EmitPlan p = plan[a];
      const uint8_t* b = in.raw + in.rec_off[a];
      const uint32_t c0 = in.coff[r], nc = in.coff[r + 1] - c0;
      const uint32_t lq = (uint32_t)p.l_seq, seqb = (lq + 1) >> 1;
      const uint32_t aux_out = emit_aux_out_len(p, lng);
      const uint32_t body = 32u + p.l_name + 4u * nc + seqb + lq + aux_out;
      const bool minus = in.strand[r] == '-';
      // ---- the 32 fixed bytes + name: copied, then the replaced fields (set_mate_info src/bam.cpp:531-588)
      for (uint32_t i = lane; i < 32 + p.l_name; i += 32) d[4 + i] = b[i];
      __syncwarp();
      if (lane == 0) {
        emit_put32(d, body);
        uint8_t* h = d + 4;
        uint32_t flag = (uint32_t)b[14] | ((uint32_t)b[15] << 8);
        flag = primary ? (flag & ~0x100u) : (flag | 0x100u);
        if (minus) flag ^= 16u;
        const uint32_t my = in.pos[r];
        int32_t mtid, mpos, isize;
        if (!paired) { flag &= ~(1u | 2u | 32u); mtid = -1; mpos = -1; isize = 0; }
        else {
          flag |= 1u;
          if (minus) flag |= 32u;
          const uint32_t mate = in.pos[other];
          if (same_tx) {
            mtid = (int32_t)in.tid[r]; mpos = (int32_t)mate; flag |= 2u;
            isize = (int32_t)my <= (int32_t)mate ? ((int32_t)mate + (int32_t)lq) - (int32_t)my : -(((int32_t)my + (int32_t)lq) - (int32_t)mate);
          } else { mtid = (int32_t)in.tid[other]; mpos = (int32_t)mate; isize = 0; flag &= ~2u; }
        }
        emit_put32(h, in.tid[r]); emit_put32(h + 4, my); h[9] = (uint8_t)(un.w & 0xffu); emit_put16(h + 12, nc);
        emit_put16(h + 14, flag); emit_put32(h + 20, (uint32_t)mtid); emit_put32(h + 24, (uint32_t)mpos); emit_put32(h + 28, (uint32_t)isize);
      }
      uint8_t* q = d + 4 + 32 + p.l_name;
      // ---- merged CIGAR of the row (already reversed for '-' transcripts by k_back)
      for (uint32_t i = lane; i < nc; i += 32) emit_put32(q + 4 * i, in.pool[c0 + i]);
      q += 4 * nc;
      // ---- sequence and qualities (reverse_complement_bam, src/bam.cpp:636-702)
      const uint8_t* s = b + 32 + p.l_name + 4 * p.n_cigar_in; const uint8_t* ql = s + seqb;
      if (minus && lq > 0) {
        for (uint32_t k = lane; k < seqb; k += 32) {
          // 4-bit complement: A(1)<->T(8), C(2)<->G(4), everything else N(15)
          const uint32_t j0 = lq - 1 - 2 * k;
          const uint32_t n0 = (s[j0 >> 1] >> ((~j0 & 1u) << 2)) & 15u;
          uint32_t n1 = 0; bool have1 = 2 * k + 1 < lq;
          if (have1) { const uint32_t j1 = j0 - 1; n1 = (s[j1 >> 1] >> ((~j1 & 1u) << 2)) & 15u; }
          const uint32_t c_hi = n0 == 1 ? 8u : n0 == 8 ? 1u : n0 == 2 ? 4u : n0 == 4 ? 2u : 15u;
          const uint32_t c_lo = !have1 ? 0u : (n1 == 1 ? 8u : n1 == 8 ? 1u : n1 == 2 ? 4u : n1 == 4 ? 2u : 15u);
          q[k] = (uint8_t)((c_hi << 4) | c_lo);
        }
        const bool rev = ql[0] != 0xff;
        for (uint32_t i = lane; i < lq; i += 32) q[seqb + i] = rev ? ql[lq - 1 - i] : ql[i];
      } else for (uint32_t i = lane; i < seqb + lq; i += 32) q[i] = s[i];
      q += seqb + lq;
      // ---- aux: the kept bytes, then NH [AS] HI
      const uint8_t* x = b + p.aux_off;
      uint32_t from = 0, w = 0;
      for (uint32_t k = 0; k <= p.n_del; k++) {
        const uint32_t to = k < p.n_del ? p.del_s[k] : p.aux_len;
        for (uint32_t i = from + lane; i < to; i += 32) q[w + (i - from)] = x[i];
        w += to - from; from = k < p.n_del ? p.del_e[k] : to;
      }
      if (lane == 0) {
        uint8_t* t = q + w;
        t[0] = 'N'; t[1] = 'H'; t[2] = 'i'; emit_put32(t + 3, un.z); t += 7;
        if (lng) {                                             // set_as_tag, src/bam.cpp:622-634
          const double score = ((double)p.as_val + (double)(in.clip ? in.clip[r] : 0)) * (in.sim ? in.sim[r] : 0.0);
          t[0] = 'A'; t[1] = 'S'; t[2] = 'i'; emit_put32(t + 3, (uint32_t)(int32_t)score); t += 7;
        }
        t[0] = 'H'; t[1] = 'I'; t[2] = 'i'; emit_put32(t + 3, un.y + e);
      }
      d += 4 + body;
      __syncwarp();
    }
  }
}

}  // namespace brb
