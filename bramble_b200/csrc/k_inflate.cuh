// Kernel group 7: BGZF inflate on the device -- the input side of what htslib does behind the reference's BAM reader
// (bgzf_read -> inflate, gclib/GSam.h:489-499 via sam_read1).  One warp per BGZF block (blocks are independent, a file has
// thousands of them): lane 0 decodes the Huffman symbols of a batch of up to 32 tokens (literal | length, distance) into shared
// memory, then all 32 lanes place them -- literals in one step, matches one after the other, each copied by the whole warp.
// RFC 1951 in full: stored, fixed and dynamic blocks, several deflate blocks per BGZF block.  The CRC-32 and ISIZE of every
// block are verified here as well (the footer of the block says what they must be).
//
// Tables per warp (shared memory): a 10-bit first-level lookup for literal/length codes and an 8-bit one for distance codes
// (entry = symbol << 4 | code length; 0 = longer code), plus the canonical count / symbol arrays for the codes the lookup does
// not cover (decoded bit by bit, as zlib's puff.c does).
#pragma once
#include "dev_common.cuh"
#include "k_deflate.cuh"

namespace brb {

constexpr int INFLATE_WARPS = 4;
constexpr uint32_t INF_LBITS = 10, INF_DBITS = 8;

struct InflateJob {
  const uint8_t* in;                 // the BGZF blocks, back to back (device copy of a piece of the file); readable 8 bytes past the end
  const uint64_t* in_off;            // per block: offset of its deflate payload
  const uint32_t* clen;              // per block: payload bytes (the 8-byte footer CRC32 | ISIZE follows the payload)
  const uint64_t* out_off;           // per block: where its output goes (prefix sums of ISIZE)
  uint8_t* out;
  uint32_t n_blocks;
  uint32_t* ticket;
  uint32_t* err;                     // first failing block + 1 (0: none), and err[1] = what
};
// what: 1 bad block type, 2 bad stored length, 3 bad code lengths, 4 bad symbol / distance too far, 5 output does not match ISIZE,
//       6 CRC-32 mismatch, 7 input ran out

struct HuffTab { uint16_t count[16]; uint16_t symbol[288]; };

struct BitIn {
  const uint8_t* p; uint32_t pos, end;      // byte cursor inside the payload
  unsigned long long buf; uint32_t cnt;     // bit buffer (LSB first), valid bits
  bool over;                                // read past the end of the payload
};
__device__ __forceinline__ void bi_refill(BitIn& b) {
  while (b.cnt <= 56) {
    unsigned long long v = 0;
    if (b.pos < b.end) v = b.p[b.pos]; else if (b.pos >= b.end + 8) b.over = true;      // (a few zero bytes past the end are fine: the last code may be peeked wide)
    b.buf |= v << b.cnt; b.cnt += 8; b.pos++;
  }
}
__device__ __forceinline__ uint32_t bi_peek(BitIn& b, uint32_t n) { return (uint32_t)(b.buf & ((1ull << n) - 1ull)); }
__device__ __forceinline__ void bi_drop(BitIn& b, uint32_t n) { b.buf >>= n; b.cnt -= n; }
__device__ __forceinline__ uint32_t bi_get(BitIn& b, uint32_t n) { if (b.cnt < n) bi_refill(b); const uint32_t v = bi_peek(b, n); bi_drop(b, n); return v; }

// Canonical Huffman code from code lengths (lane 0): count[], symbol[] (puff.c construct) and the first-level lookup.
// Returns false for an over-subscribed or (beyond the one-code case deflate allows) incomplete set.
__device__ bool huff_build(const uint8_t* len, uint32_t n, HuffTab& h, uint16_t* lut, uint32_t lbits) {
  for (int i = 0; i < 16; i++) h.count[i] = 0;
  for (uint32_t s = 0; s < n; s++) h.count[len[s]]++;
  for (uint32_t i = 0; i < (1u << lbits); i++) lut[i] = 0;
  if (h.count[0] == n) return true;                          // no codes at all: legal for the distance code of a literal-only block
  int left = 1;
  for (int l = 1; l < 16; l++) { left <<= 1; left -= h.count[l]; if (left < 0) return false; }
  uint16_t offs[16]; offs[1] = 0;
  for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + h.count[l];
  for (uint32_t s = 0; s < n; s++) if (len[s]) h.symbol[offs[len[s]]++] = (uint16_t)s;
  // first-level lookup: canonical codes in (length, symbol) order; the stream carries a code MSB first, i.e. bit-reversed in our LSB-first buffer
  uint32_t code = 0, idx = 0;
  for (uint32_t l = 1; l <= lbits; l++) {
    for (uint32_t k = 0; k < h.count[l]; k++, idx++, code++) {
      const uint32_t rev = __brev(code) >> (32 - l);
      const uint16_t e = (uint16_t)((h.symbol[idx] << 4) | l);
      for (uint32_t f = rev; f < (1u << lbits); f += 1u << l) lut[f] = e;
    }
    code <<= 1;
  }
  return left == 0 || (n - h.count[0]) == 1;                 // complete, or a single code (incomplete sets are otherwise an error, as in zlib)
}

// One symbol: lookup, else bit by bit (puff.c decode).  -1: invalid code.
__device__ __forceinline__ int huff_decode(BitIn& b, const HuffTab& h, const uint16_t* lut, uint32_t lbits) {
  if (b.cnt < 15) bi_refill(b);
  const uint16_t e = lut[bi_peek(b, lbits)];
  if (e) { bi_drop(b, e & 15u); return e >> 4; }
  int code = 0, first = 0, index = 0;
  unsigned long long bits = b.buf;
  for (int l = 1; l <= 15; l++) {
    code |= (int)(bits & 1ull); bits >>= 1;
    const int count = h.count[l];
    if (code - count < first) { bi_drop(b, (uint32_t)l); return h.symbol[index + (code - first)]; }
    index += count; first += count; first <<= 1; code <<= 1;
  }
  return -1;
}

__constant__ uint16_t c_len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
__constant__ uint8_t c_len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
__constant__ uint16_t c_dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
__constant__ uint8_t c_dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
__constant__ uint8_t c_clen_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

struct InflateShared {
  uint16_t lut_l[1u << INF_LBITS]; uint16_t lut_d[1u << INF_DBITS];
  HuffTab hl, hd;
  uint8_t lens[352];                 // 19 code-length-code lengths, then (from 32) up to 286 + 30 lengths of the two alphabets
  uint16_t tlen[32], tval[32];
};

__device__ __forceinline__ uint32_t inf_le32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

// Header of a dynamic block (lane 0): code lengths of both alphabets into S.lens, tables built.  0 = fine, else the error code.
__device__ int inflate_dynamic_header(BitIn& b, InflateShared& S) {
  const uint32_t nlen = bi_get(b, 5) + 257, ndist = bi_get(b, 5) + 1, ncode = bi_get(b, 4) + 4;
  if (nlen > 286 || ndist > 30) return 3;
  for (int i = 0; i < 19; i++) S.lens[i] = 0;
  for (uint32_t i = 0; i < ncode; i++) S.lens[c_clen_order[i]] = (uint8_t)bi_get(b, 3);
  if (!huff_build(S.lens, 19, S.hl, S.lut_d, 7)) return 3;          // (the code-length code borrows the distance lookup, 7 bits of it)
  uint8_t* L = S.lens + 32;                                          // lengths of the two alphabets, behind the 19 above
  uint32_t idx = 0;
  while (idx < nlen + ndist) {
    const int sym = huff_decode(b, S.hl, S.lut_d, 7);
    if (sym < 0) return 3;
    if (sym < 16) { L[idx++] = (uint8_t)sym; continue; }
    uint32_t rep, val = 0;
    if (sym == 16) { if (idx == 0) return 3; val = L[idx - 1]; rep = 3 + bi_get(b, 2); }
    else if (sym == 17) rep = 3 + bi_get(b, 3);
    else rep = 11 + bi_get(b, 7);
    if (idx + rep > nlen + ndist) return 3;
    while (rep--) L[idx++] = (uint8_t)val;
  }
  if (L[256] == 0) return 3;                                         // no end-of-block code
  if (!huff_build(L, nlen, S.hl, S.lut_l, INF_LBITS)) return 3;
  // (the distance lengths follow the literal/length ones; huff_build reads them before it clears the lookup they were decoded with)
  if (!huff_build(L + nlen, ndist, S.hd, S.lut_d, INF_DBITS)) return 3;
  return 0;
}
__device__ void inflate_fixed_tables(InflateShared& S) {
  uint8_t* L = S.lens;
  for (int s = 0; s < 144; s++) L[s] = 8;
  for (int s = 144; s < 256; s++) L[s] = 9;
  for (int s = 256; s < 280; s++) L[s] = 7;
  for (int s = 280; s < 288; s++) L[s] = 8;
  huff_build(L, 288, S.hl, S.lut_l, INF_LBITS);
  for (int s = 0; s < 30; s++) L[s] = 5;
  huff_build(L, 30, S.hd, S.lut_d, INF_DBITS);
}

__global__ void __launch_bounds__(32 * INFLATE_WARPS) k_bgzf_inflate(InflateJob J) {
  __shared__ InflateShared s_w[INFLATE_WARPS];
  __shared__ uint32_t s_crc[256], s_x2n[32];
  const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  {                                                           // CRC table and the x^(2^k) powers, once per CTA
    for (uint32_t v = threadIdx.x; v < 256; v += blockDim.x) { uint32_t c = v; for (int k = 0; k < 8; k++) c = (c & 1) ? (c >> 1) ^ 0xedb88320u : c >> 1; s_crc[v] = c; }
    if (threadIdx.x == 0) { uint32_t p = 1u << 30; s_x2n[0] = p; for (int n = 1; n < 32; n++) { p = crc_multmodp(p, p); s_x2n[n] = p; } }
  }
  __syncthreads();
  InflateShared& S = s_w[wib];
  const uint32_t lt_mask = (1u << lane) - 1u;
  for (;;) {
    uint32_t blk = 0;
    if (lane == 0) blk = atomicAdd(J.ticket, 1u);
    blk = __shfl_sync(0xffffffffu, blk, 0);
    if (blk >= J.n_blocks) break;
    const uint8_t* in = J.in + J.in_off[blk];
    const uint32_t clen = J.clen[blk];
    const uint32_t want_crc = inf_le32(in + clen), isize = inf_le32(in + clen + 4);
    uint8_t* out = J.out + J.out_off[blk];
    BitIn b; b.p = in; b.pos = 0; b.end = clen; b.buf = 0; b.cnt = 0; b.over = false;
    uint32_t opos = 0; int err = 0;
    bool last = isize == 0 && clen == 0;                      // (nothing at all: not a deflate stream, tolerated as an empty block)
    while (!last && !err) {
      // ---- block header (lane 0)
      uint32_t type = 0, stored_src = 0, stored_len = 0;
      if (lane == 0) {
        last = bi_get(b, 1) != 0; type = bi_get(b, 2);
        if (type == 0) {
          bi_drop(b, b.cnt & 7u);
          const uint32_t len = bi_get(b, 16), nlen = bi_get(b, 16);
          stored_src = b.pos - (b.cnt >> 3); stored_len = len;
          if ((len ^ nlen) != 0xffffu || stored_src + len > clen || opos + len > isize) err = 2;
          else { b.pos = stored_src + len; b.buf = 0; b.cnt = 0; }
        } else if (type == 1) inflate_fixed_tables(S);
        else if (type == 2) err = inflate_dynamic_header(b, S);
        else err = 1;
      }
      last = __shfl_sync(0xffffffffu, (uint32_t)last, 0) != 0; type = __shfl_sync(0xffffffffu, type, 0); err = __shfl_sync(0xffffffffu, err, 0);
      __syncwarp();
      if (err) break;
      if (type == 0) {
        stored_src = __shfl_sync(0xffffffffu, stored_src, 0); stored_len = __shfl_sync(0xffffffffu, stored_len, 0);
        for (uint32_t i = lane; i < stored_len; i += 32) out[opos + i] = in[stored_src + i];
        opos += stored_len;
        __syncwarp();
        continue;
      }
      // ---- compressed data: batches of up to 32 tokens
      bool eob = false;
      while (!eob && !err) {
        uint32_t n_tok = 0;
        if (lane == 0) {
          uint32_t o = opos;
          while (n_tok < 32) {
            const int sym = huff_decode(b, S.hl, S.lut_l, INF_LBITS);
            if (sym < 0) { err = 4; break; }
            if (sym < 256) { S.tlen[n_tok] = 1; S.tval[n_tok] = (uint16_t)sym; n_tok++; o++; continue; }
            if (sym == 256) { eob = true; break; }
            const uint32_t ls = (uint32_t)sym - 257u;
            if (ls >= 29) { err = 4; break; }
            const uint32_t len = c_len_base[ls] + bi_get(b, c_len_extra[ls]);
            const int ds = huff_decode(b, S.hd, S.lut_d, INF_DBITS);
            if (ds < 0 || ds >= 30) { err = 4; break; }
            const uint32_t dist = c_dist_base[ds] + bi_get(b, c_dist_extra[ds]);
            if (dist > o) { err = 4; break; }                 // distance too far back
            S.tlen[n_tok] = (uint16_t)len; S.tval[n_tok] = (uint16_t)(dist - 1); n_tok++; o += len;
          }
          if (!err && o > isize) err = 5;
          if (b.over) err = 7;
        }
        n_tok = __shfl_sync(0xffffffffu, n_tok, 0); eob = __shfl_sync(0xffffffffu, (uint32_t)eob, 0) != 0; err = __shfl_sync(0xffffffffu, err, 0);
        __syncwarp();
        if (err) break;
        // ---- placement: literals at once, matches in order, each by the whole warp
        const uint32_t len = lane < n_tok ? S.tlen[lane] : 0u, val = lane < n_tok ? S.tval[lane] : 0u;
        const uint32_t incl = warp_incl_scan(len, lane), at = opos + incl - len;
        if (len == 1) out[at] = (uint8_t)val;
        __syncwarp();
        for (uint32_t M = __ballot_sync(0xffffffffu, len > 1); M; M &= M - 1) {
          const int src = __ffs(M) - 1;
          const uint32_t l = __shfl_sync(0xffffffffu, len, src), d = __shfl_sync(0xffffffffu, val, src) + 1, p = __shfl_sync(0xffffffffu, at, src);
          const uint8_t* from = out + p - d;
          for (uint32_t i = lane; i < l; i += 32) out[p + i] = from[i < d ? i : i % d];
          __syncwarp();
        }
        opos += __shfl_sync(0xffffffffu, incl, 31);
        (void)lt_mask;
        __syncwarp();                                         // the token arrays are rewritten by the next batch
      }
    }
    if (!err && opos != isize) err = 5;
    // ---- CRC-32 of the output: 2 KiB per lane, combined left to right (as k_bgzf_deflate computes it for the blocks it writes)
    if (!err && isize) {
      const uint32_t n = isize, c0 = lane * 2048u, c1 = c0 + 2048u < n ? c0 + 2048u : n;
      uint32_t crc = 0xffffffffu;
      for (uint32_t i = c0; i < c1; i++) crc = s_crc[(crc ^ out[i]) & 0xffu] ^ (crc >> 8);
      crc = c0 < n ? ~crc : 0u;
      const uint32_t op_full = crc_x2nmodp(2048u, 3, s_x2n);
      uint32_t acc = 0;
      for (uint32_t k = 0; k * 2048u < n; k++) {
        const uint32_t ck = __shfl_sync(0xffffffffu, crc, k), ln = (k + 1) * 2048u <= n ? 2048u : n - k * 2048u;
        if (lane == 0) acc = k == 0 ? ck : (crc_multmodp(ln == 2048u ? op_full : crc_x2nmodp(ln, 3, s_x2n), acc) ^ ck);
      }
      acc = __shfl_sync(0xffffffffu, acc, 0);
      if (acc != want_crc) err = 6;
    }
    if (err && lane == 0) { if (atomicCAS(&J.err[0], 0u, blk + 1u) == 0u) J.err[1] = (uint32_t)err; }
    __syncwarp();
  }
}

}  // namespace brb
