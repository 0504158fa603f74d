// Kernel group 0: the index build on the device (reference build_g2t_tree src/bramble.cpp:132-211 + cgranges IITree::index).
// The host keeps exactly one step: the ORDER of the records inside a table, because cgranges sorts with libstdc++'s unstable
// std::sort and the order of equal starts is that implementation's (brb_sort_by_start, host_util.cpp).  Everything derived
// from the sorted columns is computed here:
//   k_index_pmax    running maximum of `en` inside each table (cgranges' max augmentation, flattened): one warp per table
//   k_index_links   for every record the genomically previous / next exon of its transcript, found by binary search of the
//                   neighbour's start among the table's starts (the neighbour columns of the reference are in transcript
//                   order: on '-' tables prev/next swap); validates what the reference's GTF loader guarantees (exons of a
//                   transcript disjoint and ascending, gclib/gff.cpp:988-1075) and the neighbour columns themselves
//   k_index_bins    the coarse bin directory of every table
//   k_index_recs    the 32-byte device records
//   k_pack_genome   one code byte per base -> two 4-bit codes per byte (the resident genome is always packed)
#pragma once
#include "dev_common.cuh"

namespace brb {

struct IndexCols {                    // device copies of brb_index_desc's columns (table_off as 32-bit)
  uint32_t n, n_tables;
  const uint32_t* table_off; const int32_t* st; const int32_t* en; const uint32_t* tid; const uint8_t* exon_id; const uint32_t* pos_start;
  const uint32_t* prev_st; const uint32_t* prev_en; const uint32_t* next_st; const uint32_t* next_en; const uint8_t* hpn;
};
struct IndexErr { uint32_t code, rec, aux, max_tid; unsigned long long span; uint32_t max_chain, pad_; };
// code: 1 table not sorted by start, 2 empty or negative interval, 3 overlapping exons of a transcript, 4 neighbour columns inconsistent

__device__ __forceinline__ void index_fail(IndexErr* e, uint32_t code, uint32_t rec, uint32_t aux) {
  if (atomicCAS(&e->code, 0u, code) == 0u) { e->rec = rec; e->aux = aux; }
}

__global__ void __launch_bounds__(32) k_index_pmax(IndexCols c, int32_t* __restrict__ pmax, IndexErr* err) {
  const uint32_t t = blockIdx.x; const unsigned lane = threadIdx.x;
  const uint32_t b = c.table_off[t], e = c.table_off[t + 1];
  int32_t carry = INT32_MIN, prev_st = INT32_MIN;
  for (uint32_t base = b; base < e; base += 32) {
    const uint32_t i = base + lane;
    int32_t v = INT32_MIN, s = INT32_MAX;
    if (i < e) { v = c.en[i]; s = c.st[i]; if (s < 0 || v <= s) index_fail(err, 2, i, 0); }
    int32_t ps = __shfl_up_sync(0xffffffffu, s, 1); if (lane == 0) ps = prev_st;
    if (i < e && s < ps) index_fail(err, 1, i, t);
    for (int d = 1; d < 32; d <<= 1) { const int32_t y = __shfl_up_sync(0xffffffffu, v, d); if (lane >= (unsigned)d) v = v > y ? v : y; }
    v = v > carry ? v : carry;
    if (i < e) pmax[i] = v;
    carry = __shfl_sync(0xffffffffu, v, 31); prev_st = __shfl_sync(0xffffffffu, s, 31);
  }
  if (lane == 0 && e > b) atomicAdd(&err->span, (unsigned long long)carry);
}

// index of the record of `tid` that starts at `start` in table [b, e), or NONE
__device__ __forceinline__ uint32_t index_find(const IndexCols& c, uint32_t b, uint32_t e, int32_t start, uint32_t tid) {
  uint32_t lo = b, n = e - b;
  while (n > 0) { const uint32_t h = n >> 1; if (c.st[lo + h] < start) { lo += h + 1; n -= h + 1; } else n = h; }
  for (; lo < e && c.st[lo] == start; lo++) if (c.tid[lo] == tid) return lo;
  return NONE;
}

__global__ void __launch_bounds__(256) k_index_links(IndexCols c, uint32_t* __restrict__ gprev, uint32_t* __restrict__ gnext, IndexErr* err) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.n) return;
  uint32_t t = 0;                                             // table of record i
  { uint32_t lo = 0, n = c.n_tables; while (n > 0) { const uint32_t h = n >> 1; if (c.table_off[lo + h + 1] <= i) { lo += h + 1; n -= h + 1; } else n = h; } t = lo; }
  const uint32_t b = c.table_off[t], e = c.table_off[t + 1];
  const bool minus = t & 1;
  const uint32_t tid = c.tid[i];
  atomicMax(&err->max_tid, tid);
  // neighbour columns are in TRANSCRIPT order: '+' prev = genomic prev, '-' prev = genomic next
  const bool hp = c.hpn[i] & 1, hn = (c.hpn[i] >> 1) & 1;
  const bool g_has_p = minus ? hn : hp, g_has_n = minus ? hp : hn;
  const uint32_t gp_st = minus ? c.next_st[i] : c.prev_st[i], gp_en = minus ? c.next_en[i] : c.prev_en[i];
  const uint32_t gn_st = minus ? c.prev_st[i] : c.next_st[i], gn_en = minus ? c.prev_en[i] : c.next_en[i];
  uint32_t p = NONE, nx = NONE;
  if (g_has_p) {
    p = index_find(c, b, e, (int32_t)gp_st, tid);
    if (p == NONE || (uint32_t)c.en[p] != gp_en) index_fail(err, 4, i, tid);
    else if (c.en[p] > c.st[i]) index_fail(err, 3, i, p);
  } else if (gp_st != 0 || gp_en != 0) index_fail(err, 4, i, tid);
  if (g_has_n) {
    nx = index_find(c, b, e, (int32_t)gn_st, tid);
    if (nx == NONE || (uint32_t)c.en[nx] != gn_en || nx == i) index_fail(err, 4, i, tid);
    else if (c.st[nx] < c.en[i]) index_fail(err, 3, nx, i);
  } else if (gn_st != 0 || gn_en != 0) index_fail(err, 4, i, tid);
  // ... and the neighbours must point back (their own columns name this record)
  if (p != NONE && p < c.n) { const bool back = minus ? (c.hpn[p] & 1) : ((c.hpn[p] >> 1) & 1); const uint32_t bs = minus ? c.prev_st[p] : c.next_st[p]; if (!back || bs != (uint32_t)c.st[i]) index_fail(err, 4, i, tid); }
  if (nx != NONE && nx < c.n) { const bool back = minus ? ((c.hpn[nx] >> 1) & 1) : (c.hpn[nx] & 1); const uint32_t bs = minus ? c.next_st[nx] : c.prev_st[nx]; if (!back || bs != (uint32_t)c.st[i]) index_fail(err, 4, i, tid); }
  gprev[i] = p; gnext[i] = nx;
}

// exons per transcript (a chain of more than 256 wraps the reference's uint8 exon_id): count, then the maximum
// cnt[tid] = exons | chain heads (records without a genomic predecessor) << 20: a transcript is ONE chain
__global__ void __launch_bounds__(256) k_index_chain_count(IndexCols c, const uint32_t* __restrict__ gprev, uint32_t* __restrict__ cnt) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < c.n) atomicAdd(&cnt[c.tid[i]], 1u + (gprev[i] == NONE ? (1u << 20) : 0u));
}
__global__ void __launch_bounds__(256) k_index_chain_max(uint32_t n_tid, const uint32_t* __restrict__ cnt, IndexErr* err) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t w = i < n_tid ? cnt[i] : 0;
  if (w && (w >> 20) != 1) index_fail(err, 4, NONE, i);               // exons of the transcript that the neighbour columns do not connect
  const uint32_t v = __reduce_max_sync(0xffffffffu, w & 0xfffffu);
  if ((threadIdx.x & 31) == 0 && v) atomicMax(&err->max_chain, v);
}

// bins[bin_off[t] + k] = { first i with pmax[i] > k << shift, first i with st[i] >= k << shift } (absolute record indices)
__global__ void __launch_bounds__(256) k_index_bins(IndexCols c, const int32_t* __restrict__ pmax, const uint32_t* __restrict__ bin_off, uint32_t shift, uint2* __restrict__ bins) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= bin_off[c.n_tables]) return;
  uint32_t t = 0;
  { uint32_t lo = 0, n = c.n_tables; while (n > 0) { const uint32_t h = n >> 1; if (bin_off[lo + h + 1] <= g) { lo += h + 1; n -= h + 1; } else n = h; } t = lo; }
  const uint32_t b = c.table_off[t], e = c.table_off[t + 1];
  const long long x = (long long)(g - bin_off[t]) << shift;
  uint32_t il = b, n1 = e - b, ih = b, n2 = e - b;
  while (n1 > 0) { const uint32_t h = n1 >> 1; if ((long long)pmax[il + h] <= x) { il += h + 1; n1 -= h + 1; } else n1 = h; }
  while (n2 > 0) { const uint32_t h = n2 >> 1; if ((long long)c.st[ih + h] < x) { ih += h + 1; n2 -= h + 1; } else n2 = h; }
  bins[g] = make_uint2(il, ih);
}

__global__ void __launch_bounds__(256) k_index_recs(IndexCols c, const uint32_t* __restrict__ gprev, const uint32_t* __restrict__ gnext, uint4* __restrict__ rec, bool split) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c.n) return;
  const uint32_t p = gprev[i], nx = gnext[i];
  const uint4 key = make_uint4((uint32_t)c.st[i], (uint32_t)c.en[i], c.tid[i], c.pos_start[i]);
  const uint4 link = make_uint4(p != NONE ? (uint32_t)c.en[p] : 0u, nx != NONE ? (uint32_t)c.st[nx] : (uint32_t)INT32_MAX, nx,
                                (uint32_t)c.exon_id[i] | ((uint32_t)(c.hpn[i] & 1) << 8) | ((uint32_t)((c.hpn[i] >> 1) & 1) << 9));
  if (split) { rec[i] = key; rec[(size_t)c.n + i] = link; } else { rec[2 * (size_t)i] = key; rec[2 * (size_t)i + 1] = link; }
}

// one code (0..4) per byte -> two per byte, low nibble = even base
__global__ void __launch_bounds__(256) k_pack_genome(const uint8_t* __restrict__ src, unsigned long long n, uint8_t* __restrict__ dst) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, nb = (n + 1) / 2;
  if (i >= nb) return;
  const uint8_t lo = src[2 * i], hi = 2 * i + 1 < n ? src[2 * i + 1] : 4;
  dst[i] = (uint8_t)((lo & 15u) | (hi << 4));
}

}  // namespace brb
