// Kernel group 6: the per-read-name bookkeeping of the reference's flush() (src/core.cpp:237-332) that does not depend on the
// order in which names are visited: HI (running 1-based counter over read row, then mate row, of every unit in emit
// order), NH (records of the name), MAPQ (get_mapq, :46-58) and the primary unit -- the unit with the strictly greatest
// pair score max(r.sim, m.sim), or, on a tie, the one picked by
//     std::uniform_int_distribution<uint32_t>(0, k - 1)(std::mt19937_64(std::hash<std::string>{}(read_name)))
// among the tied units in emit order (:214-218, 283-299).  c3_rng.h restates those three library pieces (checked against
// libstdc++ by tools/c3_proto.cpp and, on the device, by the parity test).  With short reads every multi-record name ties
// (similarity scores stay 0), so this moves one 312-word Mersenne-Twister seeding per such name off the host.
// What stays on the host is the ORDER of the names inside a flush chunk (std::unordered_map iteration order).
// One thread per read-name group; rows are found through the unit blocks (a_roff / a_rcnt), whichever way they are laid out.
#pragma once
#include "dev_common.cuh"
#include "c3_rng.h"

namespace brb {

struct FlushIn {
  const uint32_t* a_roff; const uint32_t* a_rcnt; const uint8_t* flags; const double* sim;     // sim == nullptr: every score is 0
  const uint64_t* name_off; const uint8_t* names;
  uint32_t cap_rows;
};
struct FlushOut { int32_t* hi; int32_t* nh; uint8_t* mapq; uint8_t* primary; };

__global__ void __launch_bounds__(128) k_flush_tags(DevBatch b, FlushIn in, FlushOut out, bool long_reads) {
  const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= b.n_groups) return;
  const uint32_t a0 = b.group_off[g], a1 = b.group_off[g + 1];
  // pass 1: HI, record count, best pair score and how many units share it
  int32_t hit = 0; uint32_t n_best = 0;
  double best = -INFINITY;
  for (uint32_t a = a0; a < a1; a++) {
    const uint32_t r0 = in.a_roff[a], cnt = in.a_rcnt[a];
    if ((unsigned long long)r0 + cnt > in.cap_rows) return;          // (the sequence is being replayed with larger pools)
    for (uint32_t r = r0; r < r0 + cnt;) {
      const bool paired = in.flags[r] & BRB_REC_PAIRED;
      double s = in.sim ? in.sim[r] : 0.0;
      out.hi[r] = ++hit;
      if (paired) { out.hi[r + 1] = ++hit; if (in.sim) { const double m = in.sim[r + 1]; s = s < m ? m : s; } }
      if (s > best) { best = s; n_best = 1; } else if (s == best) n_best++;
      r += paired ? 2 : 1;
    }
  }
  if (hit == 0) return;
  uint32_t pick = 0; bool reject = false;
  if (n_best > 1) {
    const uint64_t n0 = in.name_off[g];
    const uint64_t h = c3_hash_bytes(in.names + n0, in.name_off[g + 1] - n0);
    pick = c3_uniform_pick(c3_mt19937_64_first(h), n_best, &reject);
  }
  const uint8_t mq = (uint8_t)c3_mapq((uint32_t)hit, long_reads);
  // pass 2: NH, MAPQ, primary flag (2 = the library would have drawn again: the host decides this name)
  uint32_t k = 0;
  for (uint32_t a = a0; a < a1; a++) {
    const uint32_t r0 = in.a_roff[a], cnt = in.a_rcnt[a];
    for (uint32_t r = r0; r < r0 + cnt;) {
      const bool paired = in.flags[r] & BRB_REC_PAIRED;
      double s = in.sim ? in.sim[r] : 0.0;
      if (paired && in.sim) { const double m = in.sim[r + 1]; s = s < m ? m : s; }
      uint8_t prim = 0;
      if (s == best) { prim = reject ? 2 : (k == pick ? 1 : 0); k++; }
      for (uint32_t e = 0; e < (paired ? 2u : 1u); e++) { out.nh[r + e] = hit; out.mapq[r + e] = mq; out.primary[r + e] = prim; }
      r += paired ? 2 : 1;
    }
  }
}

}  // namespace brb
