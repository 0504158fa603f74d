// Exclusive prefix sum over uint32 (hand-written; tile scan + recursive scan of tile sums + add).
// out has n+1 entries: out[n] is the total.  Used for candidate / record / CIGAR-pool offsets.
#pragma once
#include "dev_common.cuh"

namespace brb {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;   // 2048

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total, uint32_t* sh /* >= 33 */) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= (unsigned)d) inc += y; }
  if (lane == 31) sh[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < nwarp ? sh[lane] : 0, winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, winc, d); if (lane >= (unsigned)d) winc += y; }
    if (lane < nwarp) sh[lane] = winc - w;
    if (lane == 31) sh[32] = winc;
  }
  __syncthreads();
  uint32_t r = sh[warp] + inc - v;
  *total = sh[32];
  __syncthreads();
  return r;
}

// tile-local exclusive scan; tile total -> sums[blockIdx.x]
// n_dev (optional): the live element count in device memory (<= n); tiles beyond it only report a zero sum,
// so a scan launched for a pool's CAPACITY costs what its live part costs.
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_tiles(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                             uint32_t* __restrict__ sums, uint32_t n, const uint32_t* __restrict__ n_dev) {
  __shared__ uint32_t sh[33];
  if (n_dev) { const uint32_t live = *n_dev; if (live < n) n = live; }
  if (blockIdx.x * SCAN_TILE >= n) { if (threadIdx.x == 0) sums[blockIdx.x] = 0; return; }
  const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS], s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { v[k] = (base + k < n) ? in[base + k] : 0; s += v[k]; }
  uint32_t total, ex = block_exclusive_scan(s, &total, sh);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { if (base + k < n) out[base + k] = ex; ex += v[k]; }
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_add(uint32_t* __restrict__ out, const uint32_t* __restrict__ sums_ex,
                                                           uint32_t n, uint32_t n_tiles, const uint32_t* __restrict__ n_dev) {
  if (n_dev) { const uint32_t live = *n_dev; if (live < n) n = live; }
  if (blockIdx.x * SCAN_TILE >= n) { if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = sums_ex[n_tiles]; return; }
  const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  const uint32_t add = sums_ex[blockIdx.x];
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) if (base + k < n) out[base + k] += add;
  if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = sums_ex[n_tiles];
}

// host driver; scratch must hold at least scan_scratch_words(n) uint32
inline size_t scan_scratch_words(size_t n) {
  size_t w = 0;
  while (true) { size_t t = (n + SCAN_TILE - 1) / SCAN_TILE; if (t == 0) t = 1; w += 2 * (t + 1); if (t <= 1) break; n = t; }
  return w + 8;
}

__global__ void k_scan_total1(uint32_t* __restrict__ out, const uint32_t* __restrict__ sums, uint32_t n, const uint32_t* __restrict__ n_dev) {
  if (n_dev) { const uint32_t live = *n_dev; if (live < n) n = live; }
  out[n] = sums[0];
}

// out[live n] receives the total.  n = launch size (capacity); n_dev = optional live count on the device.
inline int exclusive_scan(const uint32_t* in, uint32_t* out, uint32_t n, uint32_t* scratch, cudaStream_t st, uint64_t* launches,
                          const uint32_t* n_dev = nullptr) {
  uint32_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE; if (tiles == 0) tiles = 1;
  uint32_t* sums = scratch; uint32_t* sums_ex = scratch + (tiles + 1);
  k_scan_tiles<<<tiles, SCAN_THREADS, 0, st>>>(in, out, sums, n, n_dev); (*launches)++;
  if (tiles == 1) {
    k_scan_total1<<<1, 1, 0, st>>>(out, sums, n, n_dev); (*launches)++;   // single tile: total is sums[0]
    return 0;
  }
  exclusive_scan(sums, sums_ex, tiles, scratch + 2 * (tiles + 1), st, launches);
  k_scan_add<<<tiles, SCAN_THREADS, 0, st>>>(out, sums_ex, n, tiles, n_dev); (*launches)++;
  return 0;
}

}  // namespace brb
