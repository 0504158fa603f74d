// Kernel group 7: BGZF block encoder -- the output side of the reference's BAM writer (htslib bgzf_write / deflate behind
// sam_write1, gclib/GSam.h:553,642-645: four zlib threads there).  One WARP compresses one BGZF block (<= 0xff00 input bytes):
//   * LZ77: 32 consecutive positions per step (lane = position); 4-byte hash into a per-warp table of previous positions in
//     shared memory, match verification word by word, greedy parse of the 32 candidates (a match jumps over the positions it
//     covers).  The window is the block itself (BGZF blocks are independent), matches up to 258 bytes: the 3-4 output records
//     of a read that repeat its name, sequence and qualities collapse to a few tokens.
//   * fixed Huffman codes (BTYPE=01): no tree building; token bits are placed with a warp prefix sum and OR-ed into the
//     zeroed output words.  A block that would not fit BGZF's 64 KiB this way is emitted as a stored block.
//   * CRC-32 of the input: every lane runs the byte-wise table over its 2 KiB share, the shares are combined with zlib's
//     crc32_combine arithmetic (x^(8n) mod P by repeated squaring).
// tests: round trip through zlib inflate on the host, CRC and ISIZE fields, incompressible / repetitive / BAM-shaped / empty input.
#pragma once
#include "dev_common.cuh"

namespace brb {

constexpr uint32_t BGZF_IN = 0xff00;                       // input bytes per block (htslib BGZF_BLOCK_SIZE)
constexpr uint32_t BGZF_SLOT = 80 * 1024;                  // per-block scratch for the deflate payload (worst case 9 bits per byte)
constexpr uint32_t BGZF_MAX_PAYLOAD = 65536 - 26;          // header 18 + payload + CRC32 + ISIZE must stay within 64 KiB
constexpr int DEFLATE_WARPS = 4;
constexpr uint32_t DEFLATE_HBITS = 12;

__device__ __forceinline__ uint32_t crc_multmodp(uint32_t a, uint32_t b) {     // zlib crc32.c multmodp (reflected CRC-32 polynomial)
  uint32_t m = 1u << 31, p = 0;
  for (;;) {
    if (a & m) { p ^= b; if ((a & (m - 1)) == 0) break; }
    m >>= 1;
    b = (b & 1) ? (b >> 1) ^ 0xedb88320u : b >> 1;
  }
  return p;
}
__device__ __forceinline__ uint32_t crc_x2nmodp(uint32_t n, uint32_t k, const uint32_t* x2n) {   // x^(n * 2^k) mod P
  uint32_t p = 1u << 31;
  while (n) { if (n & 1) p = crc_multmodp(x2n[k & 31], p); n >>= 1; k++; }
  return p;
}

struct DeflateJob { const uint8_t* in; uint64_t n; uint8_t* slots; uint32_t* sizes; uint32_t* crcs; uint32_t n_blocks; uint32_t* ticket; };

// the 32-bit word at byte offset p of a buffer whose base is 4-byte aligned (2 aligned loads + funnel shift)
__device__ __forceinline__ uint32_t load_u32_at(const uint32_t* w, uint32_t p) {
  const uint32_t i = p >> 2, sh = (p & 3u) * 8u;
  const uint32_t lo = __ldg(&w[i]);
  return sh ? __funnelshift_r(lo, __ldg(&w[i + 1]), sh) : lo;
}

__global__ void __launch_bounds__(32 * DEFLATE_WARPS) k_bgzf_deflate(DeflateJob J) {
  __shared__ uint16_t s_tab[DEFLATE_WARPS][1u << DEFLATE_HBITS];
  __shared__ uint32_t s_crc[256], s_x2n[32];
  const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  {                                                           // CRC table and the x^(2^k) powers, once per CTA
    uint32_t c = threadIdx.x;
    for (uint32_t v = threadIdx.x; v < 256; v += blockDim.x) { c = v; for (int k = 0; k < 8; k++) c = (c & 1) ? (c >> 1) ^ 0xedb88320u : c >> 1; s_crc[v] = c; }
    if (threadIdx.x == 0) { uint32_t p = 1u << 30; s_x2n[0] = p; for (int n = 1; n < 32; n++) { p = crc_multmodp(p, p); s_x2n[n] = p; } }
  }
  __syncthreads();
  uint16_t* tab = s_tab[wib];
  const uint32_t lt_mask = (1u << lane) - 1u;
  for (;;) {
    uint32_t blk = 0;
    if (lane == 0) blk = atomicAdd(J.ticket, 1u);
    blk = __shfl_sync(0xffffffffu, blk, 0);
    if (blk >= J.n_blocks) break;
    const uint64_t off = (uint64_t)blk * BGZF_IN;
    const uint32_t n = (uint32_t)(J.n - off < BGZF_IN ? J.n - off : BGZF_IN);
    const uint8_t* in = J.in + off;                           // (off is a multiple of 4: aligned word view)
    const uint32_t* inw = reinterpret_cast<const uint32_t*>(in);
    uint32_t* out = reinterpret_cast<uint32_t*>(J.slots + (size_t)blk * BGZF_SLOT);
    for (uint32_t i = lane; i < BGZF_SLOT / 4; i += 32) out[i] = 0;
    for (uint32_t i = lane; i < (1u << DEFLATE_HBITS); i += 32) tab[i] = 0xffff;
    __syncwarp();
    // ---- CRC-32 of the block: 2 KiB per lane, then crc32_combine left to right
    {
      const uint32_t c0 = lane * 2048u, c1 = c0 + 2048u < n ? c0 + 2048u : n;
      uint32_t crc = 0xffffffffu;
      for (uint32_t i = c0; i < c1; i++) crc = s_crc[(crc ^ in[i]) & 0xffu] ^ (crc >> 8);
      crc = c0 < n ? ~crc : 0u;
      const uint32_t op_full = crc_x2nmodp(2048u, 3, s_x2n);       // shift over one full share
      uint32_t acc = 0;
      for (uint32_t k = 0; k * 2048u < n; k++) {
        const uint32_t ck = __shfl_sync(0xffffffffu, crc, k), len = (k + 1) * 2048u <= n ? 2048u : n - k * 2048u;
        if (lane == 0) acc = k == 0 ? ck : (crc_multmodp(len == 2048u ? op_full : crc_x2nmodp(len, 3, s_x2n), acc) ^ ck);
      }
      if (lane == 0) J.crcs[blk] = acc;
    }
    // ---- LZ77 + fixed Huffman
    uint32_t bitpos = 3;                                       // BFINAL = 1, BTYPE = 01 (fixed): bits 1,1,0
    if (lane == 0) out[0] = 0x3u;
    __syncwarp();
    uint32_t pos = 0;
    while (pos < n) {
      const uint32_t p = pos + lane;
      uint32_t mlen = 0, dist = 0, lit = 0;
      if (p < n) lit = in[p];
      const bool can = p + 3 < n;
      uint32_t h = 0, cand = 0xffff;
      if (can) { const uint32_t v = load_u32_at(inw, p); h = (v * 2654435761u) >> (32 - DEFLATE_HBITS); cand = tab[h]; }
      __syncwarp();
      if (can) tab[h] = (uint16_t)p;
      __syncwarp();
      if (cand != 0xffff && cand < p && p - cand <= 32768u) {          // (deflate distances end at 32 KiB; the block is up to 0xff00 bytes)
        uint32_t l = 0; const uint32_t maxl = n - p < 258u ? n - p : 258u;
        while (l + 4 <= maxl) { const uint32_t x = load_u32_at(inw, cand + l) ^ load_u32_at(inw, p + l); if (x) { l += (__ffs(x) - 1) >> 3; break; } l += 4; }
        if (l + 4 > maxl) while (l < maxl && in[cand + l] == in[p + l]) l++;     // (tail, or the word loop stopped short of a difference)
        if (l > maxl) l = maxl;
        if (l >= 4) { mlen = l; dist = p - cand; }
      }
      // greedy parse of the step: a match at the current position is taken and skips what it covers
      uint32_t starts = 0, cur = 0;
      const uint32_t lim = n - pos < 32u ? n - pos : 32u;
      while (cur < lim) { const uint32_t l = __shfl_sync(0xffffffffu, mlen, cur); starts |= 1u << cur; cur += l ? l : 1u; }
      // token bits
      uint32_t bits = 0, nb = 0;
      if ((starts >> lane) & 1u) {
        if (mlen == 0) {
          if (lit < 144) { bits = __brev(0x30u + lit) >> 24; nb = 8; } else { bits = __brev(0x190u + (lit - 144)) >> 23; nb = 9; }
        } else {
          const uint32_t l = mlen - 3; uint32_t code, e = 0, ev = 0;
          if (l < 8) code = 257 + l; else if (mlen == 258) code = 285;
          else { const uint32_t lg = 31 - __clz(l); e = lg - 2; code = 265 + 4 * (e - 1) + ((l >> e) & 3u); ev = l & ((1u << e) - 1u); }
          if (code < 280) { bits = __brev(code - 256) >> 25; nb = 7; } else { bits = __brev(0xC0u + (code - 280)) >> 24; nb = 8; }
          bits |= ev << nb; nb += e;
          const uint32_t d = dist - 1; uint32_t dc, de = 0, dv = 0;
          if (d < 4) dc = d; else { const uint32_t lg = 31 - __clz(d); dc = 2 * lg + ((d >> (lg - 1)) & 1u); de = lg - 1; dv = d & ((1u << de) - 1u); }
          bits |= (__brev(dc) >> 27) << nb; nb += 5;
          bits |= dv << nb; nb += de;
        }
      }
      const uint32_t incl = warp_incl_scan(nb, lane), at = bitpos + incl - nb;
      if (nb) {
        const uint32_t w = at >> 5, sh = at & 31u;
        atomicOr(&out[w], bits << sh);
        if (sh + nb > 32) atomicOr(&out[w + 1], bits >> (32 - sh));
      }
      bitpos += __shfl_sync(0xffffffffu, incl, 31);
      pos += cur;
      (void)lt_mask;
    }
    bitpos += 7;                                               // end of block: symbol 256 = seven 0 bits (already there)
    uint32_t bytes = (bitpos + 7) >> 3;
    __syncwarp();
    if (bytes > BGZF_MAX_PAYLOAD || bytes >= n + 5) {          // incompressible: a stored block (BFINAL=1, BTYPE=00, LEN, ~LEN, bytes)
      uint8_t* ob = reinterpret_cast<uint8_t*>(out);
      if (lane == 0) { ob[0] = 1; ob[1] = (uint8_t)n; ob[2] = (uint8_t)(n >> 8); ob[3] = (uint8_t)~n; ob[4] = (uint8_t)(~n >> 8); }
      for (uint32_t i = lane; i < n; i += 32) ob[5 + i] = in[i];
      bytes = n + 5;
    }
    if (lane == 0) J.sizes[blk] = bytes;
    __syncwarp();
  }
}

// The finished blocks, densely: 18-byte BGZF header, payload, CRC32, ISIZE -- at the offsets an exclusive scan of (size + 26) gave.
__global__ void __launch_bounds__(256) k_bgzf_pack(DeflateJob J, const unsigned long long* __restrict__ off, uint8_t* __restrict__ dst) {
  const uint32_t blk = blockIdx.x;
  if (blk >= J.n_blocks) return;
  const uint32_t sz = J.sizes[blk];
  const uint64_t o = off[blk];
  const uint64_t in_off = (uint64_t)blk * BGZF_IN;
  const uint32_t n = (uint32_t)(J.n - in_off < BGZF_IN ? J.n - in_off : BGZF_IN);
  uint8_t* d = dst + o;
  const uint8_t* s = J.slots + (size_t)blk * BGZF_SLOT;
  if (threadIdx.x == 0) {
    const uint8_t head[16] = {0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 'B', 'C', 2, 0};
    for (int i = 0; i < 16; i++) d[i] = head[i];
    const uint32_t bs = sz + 25; d[16] = (uint8_t)bs; d[17] = (uint8_t)(bs >> 8);
    const uint32_t crc = J.crcs[blk];
    for (int i = 0; i < 4; i++) { d[18 + sz + i] = (uint8_t)(crc >> (8 * i)); d[22 + sz + i] = (uint8_t)(n >> (8 * i)); }
  }
  for (uint32_t i = threadIdx.x; i < sz; i += blockDim.x) d[18 + i] = s[i];
}
__global__ void k_bgzf_sizes(DeflateJob J, unsigned long long* __restrict__ tot) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < J.n_blocks) tot[i] = (unsigned long long)J.sizes[i] + 26ull;
}

}  // namespace brb
