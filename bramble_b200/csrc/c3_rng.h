// The three libstdc++ (GCC 13) pieces that decide the reference's primary tie-break (src/core.cpp:214-218, 298-299):
//     seed = std::hash<std::string>{}(read_name);  std::mt19937_64 gen(seed);
//     idx  = std::uniform_int_distribution<uint32_t>(0, k - 1)(gen);
// restated so that they compile for the host (tools/c3_proto.cpp checks them against the real library) and for the device
// (k_flush.cuh).  Not a general RNG: exactly the first draw of a freshly seeded engine.
#pragma once
#include <stdint.h>
#ifdef __CUDACC__
#define BRB_HD __host__ __device__ __forceinline__
#else
#define BRB_HD inline
#endif

namespace brb {

// std::hash<std::string>: _Hash_bytes(ptr, len, 0xc70f6907), libstdc++-v3/libsupc++/hash_bytes.cc (64-bit size_t: a
// Murmur-style mix, 8 bytes at a time, little endian)
BRB_HD uint64_t c3_hash_bytes(const uint8_t* p, uint64_t len) {
  const uint64_t mul = (0xc6a4a793ull << 32) + 0x5bd1e995ull;
  uint64_t hash = 0xc70f6907ull ^ (len * mul);
  const uint64_t n8 = len & ~7ull;
  for (uint64_t i = 0; i < n8; i += 8) {
    uint64_t d = 0;
    for (int k = 7; k >= 0; k--) d = (d << 8) | p[i + k];
    d *= mul; d ^= d >> 47; d *= mul;
    hash ^= d; hash *= mul;
  }
  if (len & 7) {
    uint64_t d = 0;
    for (int k = (int)(len & 7) - 1; k >= 0; k--) d = (d << 8) + p[n8 + k];
    hash ^= d; hash *= mul;
  }
  hash ^= hash >> 47; hash *= mul; hash ^= hash >> 47;
  return hash;
}

// First output of std::mt19937_64 seeded with `seed`: the seeding recurrence up to word 156 (the first twist only reads
// words 0, 1 and 156), one twist, the tempering.
BRB_HD uint64_t c3_mt19937_64_first(uint64_t seed) {
  uint64_t x = seed, m0 = seed, m1 = 0, m156 = 0;
  for (uint32_t i = 1; i <= 156; i++) {
    x = 6364136223846793005ull * (x ^ (x >> 62)) + i;
    if (i == 1) m1 = x;
    if (i == 156) m156 = x;
  }
  const uint64_t y = (m0 & 0xffffffff80000000ull) | (m1 & 0x7fffffffull);
  uint64_t z = m156 ^ (y >> 1) ^ ((y & 1) ? 0xb5026f5aa96619e9ull : 0ull);
  z ^= (z >> 29) & 0x5555555555555555ull;
  z ^= (z << 17) & 0x71d67fffeda60000ull;
  z ^= (z << 37) & 0xfff7eee000000000ull;
  z ^= z >> 43;
  return z;
}

// std::uniform_int_distribution<uint32_t>(0, k - 1) on a 64-bit engine: Lemire's multiply-shift on 128 bits
// (bits/uniform_int_dist.h, _S_nd).  *reject is set when the library would draw again (low 64 bits below 2^64 mod k --
// probability k / 2^64): the caller then leaves that name to the host.
BRB_HD uint32_t c3_uniform_pick(uint64_t g, uint32_t k, bool* reject) {
#ifdef __CUDA_ARCH__
  const uint64_t hi = __umul64hi(g, (uint64_t)k), lo = g * (uint64_t)k;
#else
  const unsigned __int128 prod = (unsigned __int128)g * k;
  const uint64_t hi = (uint64_t)(prod >> 64), lo = (uint64_t)prod;
#endif
  *reject = lo < k && lo < (0 - (uint64_t)k) % k;
  return (uint32_t)hi;
}

// get_mapq, src/core.cpp:46-58
BRB_HD uint32_t c3_mapq(uint32_t nh, bool long_reads) {
  if (!long_reads) return nh == 1 ? 255u : nh == 2 ? 3u : (nh == 3 || nh == 4) ? 1u : 0u;
  return nh > 1 ? 0u : 3u;
}

}  // namespace brb
