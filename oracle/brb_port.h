/* TEST INFRASTRUCTURE -- CPU restatement ("port") of the reference's projection hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may call this; the product
 * (bramble_b200/) never links or loads it.  Same data contract as include/bramble_b200.h. */
#ifndef BRB_PORT_H
#define BRB_PORT_H
#include "../include/bramble_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct brb_port_out {           /* malloc'ed by brb_port_project, released by brb_port_free */
  uint32_t n_rec; uint64_t n_cigar_words;
  uint32_t* group_rec_off; uint32_t* src_aln; uint32_t* tid; uint32_t* pos; int8_t* tx_strand; uint8_t* flags;
  double* sim_score; int32_t* clip_score; uint32_t* cigar_off; uint32_t* cigar_len; uint32_t* cigar;
  uint64_t n_matches, n_rescue, n_hits;  /* work counters (for the algorithmic-bytes formula) */
  double seconds_project;                /* wall time of the projection itself (index build excluded), one thread: the CPU figure at
                                          * the boundary of the engine's resident step (columns in, rows out) */
} brb_port_out;

int brb_port_project(const brb_index_desc* idx, const brb_params* p, const brb_batch_in* in, brb_port_out* out);
void brb_port_free(brb_port_out* out);

int brb_port_lookup(const brb_index_desc* idx, const brb_params* p, uint32_t n_query, const int32_t* refid,
                    const int8_t* strand, const uint32_t* qs, const uint32_t* qe, const uint8_t* status,
                    uint32_t cap, uint32_t* hits, uint32_t* n_hits);
int brb_port_merge_cigars(uint32_t n, const uint32_t* real_off, const uint32_t* real, const uint32_t* ideal_off,
                          const uint32_t* ideal, uint32_t* out_off, uint32_t* out_len, uint32_t* out);
/* flush() of the reference (src/core.cpp:237-332) per read-name group, with the REAL libstdc++ std::hash<std::string>,
 * std::mt19937_64 and std::uniform_int_distribution (src/core.cpp:214-218): HI, NH, MAPQ and the primary flag of every row.
 * Rows are in emit order; group g owns rows [group_rec_off[g], group_rec_off[g + 1]). */
int brb_port_flush_tags(uint32_t n_groups, const uint32_t* group_rec_off, const uint8_t* flags, const double* sim,
                        const uint64_t* name_off, const uint8_t* names, int long_reads,
                        int32_t* hi, int32_t* nh, uint8_t* mapq, uint8_t* primary);
int brb_port_ksw_extend(uint32_t n, const uint64_t* q_off, const uint8_t* q, const uint64_t* t_off, const uint8_t* t,
                        int32_t* max, int32_t* score, int32_t* max_q, int32_t* max_t, uint8_t* zdropped,
                        uint32_t cigar_cap, uint32_t* n_cigar, uint32_t* cigar_out);
#ifdef __cplusplus
}
#endif
#endif
