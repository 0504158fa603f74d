// TEST INFRASTRUCTURE.  Calls the reference's OWN ksw2 translation unit
// (/root/reference/subprojects/packagefiles/ksw2/ksw2_extz2_sse.cpp, compiled unmodified into oracle/_ref/) with the
// arguments of the reference's align() (src/evaluate.cpp:285-317: 5x5 matrix match 1 / mismatch -4 / N row 0 -> sc_N = -e,
// q=4, e=1, w=-1, zdrop=40, end_bonus=0, EXTZ_ONLY|APPROX_MAX|APPROX_DROP), on problems given as base codes 0..4.
// Same signature as brb_port_ksw_extend: a CPU test pins the restatement (oracle/brb_port.cpp) on it problem by problem.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include "ksw2.h"

extern "C" int brb_ref_ksw_extend(uint32_t n, const uint64_t* q_off, const uint8_t* q, const uint64_t* t_off, const uint8_t* t,
                                  int32_t* max, int32_t* score, int32_t* max_q, int32_t* max_t, uint8_t* zdropped,
                                  uint32_t cigar_cap, uint32_t* n_cigar, uint32_t* cigar_out) {
  const int a = 1, b = -4;
  int8_t mat[25] = {(int8_t)a, (int8_t)b, (int8_t)b, (int8_t)b, 0, (int8_t)b, (int8_t)a, (int8_t)b, (int8_t)b, 0, (int8_t)b, (int8_t)b, (int8_t)a, (int8_t)b, 0,
                    (int8_t)b, (int8_t)b, (int8_t)b, (int8_t)a, 0, 0, 0, 0, 0, 0};
  const int flag = KSW_EZ_EXTZ_ONLY | KSW_EZ_APPROX_MAX | KSW_EZ_APPROX_DROP;
  for (uint32_t i = 0; i < n; i++) {
    const int ql = (int)(q_off[i + 1] - q_off[i]), tl = (int)(t_off[i + 1] - t_off[i]);
    ksw_extz_t ez; memset(&ez, 0, sizeof ez);
    if (ql > 0 && tl > 0) ksw_extz2_sse(0, ql, q + q_off[i], tl, t + t_off[i], 5, mat, 4, 1, -1, 40, 0, flag, &ez);
    else ksw_reset_extz(&ez);
    max[i] = (int32_t)ez.max; score[i] = ez.score; max_q[i] = ez.max_q; max_t[i] = ez.max_t; zdropped[i] = (uint8_t)ez.zdropped;
    n_cigar[i] = (uint32_t)ez.n_cigar;
    for (int k = 0; k < ez.n_cigar && (uint32_t)k < cigar_cap && cigar_out; k++) cigar_out[(size_t)i * cigar_cap + k] = ez.cigar[k];
    free(ez.cigar);
  }
  return 0;
}
