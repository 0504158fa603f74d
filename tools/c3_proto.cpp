// Development harness: bramble_b200/csrc/c3_rng.h (the restated std::hash<std::string> / std::mt19937_64 first draw /
// std::uniform_int_distribution multiply-shift) against the real libstdc++ of this toolchain, on random read names.
//   g++ -O2 -std=c++17 -o /tmp/c3_proto tools/c3_proto.cpp && /tmp/c3_proto
#include <cstdio>
#include <random>
#include <string>
#include "../bramble_b200/csrc/c3_rng.h"

int main() {
  std::mt19937_64 rng(99);
  long bad = 0, rejects = 0;
  for (long t = 0; t < 2000000; t++) {
    std::string name;
    const int len = (int)(rng() % 40);
    for (int i = 0; i < len; i++) name.push_back((char)(33 + rng() % 94));
    if (t % 3 == 0) { char buf[32]; snprintf(buf, sizeof buf, "r%09ld", t); name = buf; }
    const uint64_t h = std::hash<std::string>{}(name);
    const uint64_t mine = brb::c3_hash_bytes((const uint8_t*)name.data(), name.size());
    const uint32_t k = 2 + (uint32_t)(rng() % (t % 5 == 0 ? 4000000000u : 60));
    std::mt19937_64 gen(h);
    std::uniform_int_distribution<uint32_t> dis(0, k - 1);
    const uint32_t want = dis(gen);
    bool rej = false;
    const uint32_t got = brb::c3_uniform_pick(brb::c3_mt19937_64_first(mine), k, &rej);
    rejects += rej;
    if (mine != h || (!rej && got != want)) { if (bad < 5) printf("MISMATCH '%s' k=%u hash %llx/%llx pick %u/%u\n", name.c_str(), k, (unsigned long long)mine, (unsigned long long)h, got, want); bad++; }
  }
  printf("2000000 names, %ld mismatches, %ld rejections\n", bad, rejects);
  return bad != 0;
}
