// Exclusive prefix sum (hand-written; tile scan + recursive scan of tile sums + add) over uint32 or uint64.
// out has n+1 entries: out[n] is the total.  Used for record / CIGAR-pool offsets; the 64-bit form carries two
// 32-bit counters at once (rows in the low half, CIGAR words in the high half; neither total reaches 2^32).
#pragma once
#include "dev_common.cuh"

namespace brb {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;   // 2048

template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T* total, T* sh /* >= 33 */) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  T inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { T y = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= (unsigned)d) inc += y; }
  if (lane == 31) sh[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    T w = lane < nwarp ? sh[lane] : 0, winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { T y = __shfl_up_sync(0xffffffffu, winc, d); if (lane >= (unsigned)d) winc += y; }
    if (lane < nwarp) sh[lane] = winc - w;
    if (lane == 31) sh[32] = winc;
  }
  __syncthreads();
  T r = sh[warp] + inc - v;
  *total = sh[32];
  __syncthreads();
  return r;
}

// tile-local exclusive scan; tile total -> sums[blockIdx.x]
// n_dev (optional): the live element count in device memory (<= n); tiles beyond it only report a zero sum,
// so a scan launched for a pool's CAPACITY costs what its live part costs.
template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_tiles(const T* __restrict__ in, T* __restrict__ out,
                                                             T* __restrict__ sums, uint32_t n, const uint32_t* __restrict__ n_dev) {
  __shared__ T sh[33];
  if (n_dev) { const uint32_t live = *n_dev; if (live < n) n = live; }
  if (blockIdx.x * SCAN_TILE >= n) { if (threadIdx.x == 0) sums[blockIdx.x] = 0; return; }
  const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  T v[SCAN_ITEMS], s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { v[k] = (base + k < n) ? in[base + k] : 0; s += v[k]; }
  T total, ex = block_exclusive_scan<T>(s, &total, sh);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { if (base + k < n) out[base + k] = ex; ex += v[k]; }
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_add(T* __restrict__ out, const T* __restrict__ sums_ex,
                                                           uint32_t n, uint32_t n_tiles, const uint32_t* __restrict__ n_dev) {
  if (n_dev) { const uint32_t live = *n_dev; if (live < n) n = live; }
  if (blockIdx.x * SCAN_TILE >= n) { if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = sums_ex[n_tiles]; return; }
  const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  const T add = sums_ex[blockIdx.x];
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) if (base + k < n) out[base + k] += add;
  if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = sums_ex[n_tiles];
}

template <typename T>
__global__ void k_scan_total1(T* __restrict__ out, const T* __restrict__ sums, uint32_t n, const uint32_t* __restrict__ n_dev) {
  if (n_dev) { const uint32_t live = *n_dev; if (live < n) n = live; }
  out[n] = sums[0];
}

// host driver; scratch must hold at least scan_scratch_words(n) elements of T
inline size_t scan_scratch_words(size_t n) {
  size_t w = 0;
  while (true) { size_t t = (n + SCAN_TILE - 1) / SCAN_TILE; if (t == 0) t = 1; w += 2 * (t + 1); if (t <= 1) break; n = t; }
  return w + 8;
}

// out[live n] receives the total.  n = launch size (capacity); n_dev = optional live count on the device.
template <typename T>
inline int exclusive_scan(const T* in, T* out, uint32_t n, T* scratch, cudaStream_t st, uint64_t* launches,
                          const uint32_t* n_dev = nullptr) {
  uint32_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE; if (tiles == 0) tiles = 1;
  T* sums = scratch; T* sums_ex = scratch + (tiles + 1);
  k_scan_tiles<T><<<tiles, SCAN_THREADS, 0, st>>>(in, out, sums, n, n_dev); (*launches)++;
  if (tiles == 1) {
    k_scan_total1<T><<<1, 1, 0, st>>>(out, sums, n, n_dev); (*launches)++;   // single tile: total is sums[0]
    return 0;
  }
  exclusive_scan<T>(sums, sums_ex, tiles, scratch + 2 * (tiles + 1), st, launches);
  k_scan_add<T><<<tiles, SCAN_THREADS, 0, st>>>(out, sums_ex, n, tiles, n_dev); (*launches)++;
  return 0;
}

}  // namespace brb
