/* bramble_b200 -- C ABI of the B200 projection engine.
 *
 * The reference (zrudnick/bramble, C++) has no plugin/FFI surface; the seam this library replaces
 * is the per-bundle call
 *     void convert_reads(std::vector<CReadAln>&, std::shared_ptr<g2tTree>, std::shared_ptr<ReadEvaluator>,
 *                        quill::Logger*, BamIO*)                  reference include/evaluate.h:377-381, src/core.cpp:220
 * plus the one-time index build
 *     std::shared_ptr<g2tTree> build_g2t_tree(GVec<GRefData>, BamIO*)   reference src/bramble.cpp:132-211
 * A maintainer of the reference would call brb_index_create() where build_g2t_tree() is called
 * (src/bramble.cpp:639) and brb_project() inside process_bundle() (src/threads.cpp:100-104); see
 * INTEGRATION.md for the binding.
 *
 * Everything is plain pointers + sizes.  All arrays are host memory owned by the caller unless stated
 * otherwise; output arrays are owned by the brb_ctx and stay valid until the next call on that ctx.
 * Every function returns 0 on success or a negative brb_status; nothing exits or throws.  There is no
 * CPU implementation behind this ABI: without a CUDA device every compute entry point returns
 * BRB_E_NO_DEVICE.
 */
#ifndef BRAMBLE_B200_H
#define BRAMBLE_B200_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum brb_status {
  BRB_OK = 0,
  BRB_E_NO_DEVICE = -1,   /* no CUDA device / driver */
  BRB_E_CUDA = -2,        /* a CUDA call failed; brb_last_error() has the text */
  BRB_E_ARG = -3,         /* malformed input (NULL pointer, inconsistent CSR, ...) */
  BRB_E_NOMEM = -4,
  BRB_E_UNSUPPORTED = -5  /* e.g. use_fasta requested but the index holds no genome */
} brb_status;

typedef struct brb_index brb_index;   /* device-resident flattened g2t table; immutable after create */
typedef struct brb_ctx brb_ctx;       /* one CUDA stream + scratch + pinned staging; one per host thread */

/* Resolved thresholds: reference ShortReadEvaluator/LongReadEvaluator::evaluate,
 * src/evaluate.cpp:1136-1221, struct ReadEvaluationConfig include/evaluate.h:275-285. */
typedef struct brb_params {
  uint32_t max_clip, max_junc_ins, max_junc_gap, max_error_exon;
  float    similarity_threshold;        /* compared as (double)float, src/evaluate.cpp:854-856 */
  uint8_t  long_reads;                  /* LONG_READS global, src/bramble.cpp:504 */
  uint8_t  use_fasta;                   /* USE_FASTA global, src/bramble.cpp:499-502 */
  uint8_t  ignore_small_exons;          /* max_error_exon > 0 */
  uint8_t  filter_by_similarity;        /* similarity_threshold < 1.0 */
} brb_params;

/* Flattened interval table = the reference's per-(refid,strand) IITree<int,IITData> arrays
 * (include/g2t.h:27-43, src/g2t.cpp:34-62) concatenated; table t = 2*refid + (strand=='-').
 * Records of one table are in the reference's own order (std::sort by start on the reference's
 * insertion sequence, cgranges IITree::index; brb_sort_by_start yields it), so "hit order" == ascending record index.
 * That order is the host's only share of the index build: brb_index_create uploads the columns and derives the
 * max-augmentation, the exon chains, the bin directory and the device records on the GPU, validating on the way
 * (sorted tables, disjoint ascending exons per transcript, consistent neighbour columns). */
typedef struct brb_index_desc {
  uint32_t n_tables;                    /* 2 * number of annotated chromosomes */
  const uint64_t* table_off;            /* n_tables+1 offsets into the record arrays */
  const int32_t*  st;                   /* interval start, 1-based */
  const int32_t*  en;                   /* interval end, exclusive */
  const uint32_t* tid;                  /* output-header index of the transcript */
  const uint8_t*  exon_id;              /* (uint8_t) genomic exon index -- wraps at 256 like the reference */
  const uint32_t* pos_start;            /* cumulative transcript offset of the exon (5'->3') */
  const uint32_t* prev_st; const uint32_t* prev_en;   /* neighbour exons in transcript order, 0,0 if none */
  const uint32_t* next_st; const uint32_t* next_en;
  const uint8_t*  has_prev_next;        /* bit0 has_prev, bit1 has_next */
  const uint32_t* transcript_len;
  /* genome, only for use_fasta: per-refid base codes 0..3 = A,C,G,T, 4 = anything else (the ksw2
   * encoding of reference src/evaluate.cpp:299-305), 0-based genomic offset.  genome_format 0: one code per byte;
   * 1: two codes per byte (base o in byte o / 2, low nibble = even o) -- half the upload.  Resident on the device it is
   * always the packed form. */
  uint32_t n_chrom;
  uint32_t genome_format;
  const uint64_t* chrom_off;            /* n_chrom+1 offsets (in bases) into the genome, or NULL */
  const uint8_t*  genome_codes;         /* or NULL */
} brb_index_desc;

/* One bundle (reference BundleData::reads, include/bramble.h:153-183) as structure-of-arrays. */
typedef struct brb_batch_in {
  uint32_t n_aln, n_groups;
  const uint32_t* group_off;            /* n_groups+1: read-name groups (src/core.cpp:347-380) */
  const int32_t*  refid;                /* n_aln: annotated chromosome index, <0 or >=n_tables/2: none */
  const int8_t*   strand;               /* n_aln: '+', '-', '.' (src/bramble.cpp:213-244,382) */
  const uint32_t* exon_off;             /* n_aln+1 */
  const uint32_t* exon_st;              /* read exon segments, 1-based half-open (src/bramble.cpp:246-255) */
  const uint32_t* exon_en;
  const uint32_t* cigar_off;            /* n_aln+1 */
  const uint32_t* cigar;                /* raw BAM CIGAR words */
  const int32_t*  mate_idx;             /* n_aln: batch-local index of the linked mate or -1 (process_pairs) */
  const uint64_t* group_seq_off;        /* n_groups+1 base offsets into seq_codes; NULL unless use_fasta */
  const uint8_t*  seq_codes;            /* group-shared read sequence (src/core.cpp:353-378), codes 0..4 */
  const uint64_t* group_name_off;       /* n_groups+1 byte offsets into name_bytes; NULL unless BRB_OPT_FLUSH_TAGS is used */
  const uint8_t*  name_bytes;           /* the read names (QNAME without the terminator), group after group */
  /* Compact alternatives (each optional, NULL = not used): what bounds a caller that streams bundles through brb_project() is the
   * bytes that cross the bus, and four of the columns above can be said in fewer.  When one of these is given, the column it
   * replaces may be NULL; the engine expands it on the device (k_expand.cuh). */
  const int32_t*  pos0;                 /* n_aln: 0-based leftmost coordinate (BAM pos).  Replaces exon_off / exon_st / exon_en: the read
                                         * exons are derived from pos0 + CIGAR on the device exactly as the reference's ingest derives them
                                         * (GSamRecord::setupCoordinates + process_exons, src/bramble.cpp:246-255; an alignment whose CIGAR
                                         * yields no exon is the reference's "invalid CIGAR" error: BRB_E_ARG) */
  const uint16_t* cigar_cnt16;          /* n_aln: CIGAR words per alignment (BAM's n_cigar_op is 16 bits too).  Replaces cigar_off;
                                         * n_cigar_total must hold their sum */
  const int16_t*  mate_delta16;         /* n_aln: 0 = no mate, else mate index - own index (mates sit in the same read-name group).  Replaces mate_idx */
  const int16_t*  refid16;              /* n_aln: as refid.  Replaces refid */
  uint64_t        n_cigar_total;        /* only read with cigar_cnt16 */
} brb_batch_in;

/* One row per output BAM record.  Rows come in UNIT blocks: the unit owned by alignment a (an unpaired alignment, or
 * the lower-indexed end of a linked pair -- `this_read` of src/core.cpp:384-415) has its rows at
 * [aln_rec_off[a], aln_rec_off[a] + aln_rec_cnt[a]), ascending by tid exactly as add_mate_info emits them
 * (src/mates.cpp:28-119); for a paired BamInfo the read1 row is immediately followed by its read2 row.  The reference's
 * emit order of a bundle is the concatenation of these blocks in alignment order (group by group).  By default the
 * blocks of different 32-alignment batches sit in the row arrays in the order the GPU finished them; with
 * BRB_OPT_ORDERED_ROWS the row arrays themselves are in emit order and group_rec_off is filled in as well. */
#define BRB_REC_FIRST  1u   /* this row is BamInfo::read1 */
#define BRB_REC_PAIRED 2u   /* BamInfo::is_paired */
#define BRB_REC_SAMETX 4u   /* BamInfo::same_transcript */
typedef struct brb_batch_out {
  uint32_t n_rec;
  uint64_t n_cigar_words;               /* size of the cigar pool == cigar_off[n_rec] (dense, row order) */
  const uint32_t* group_rec_off;        /* n_groups+1: rows of group g -- only with BRB_OPT_ORDERED_ROWS, else NULL */
  const uint32_t* src_aln;              /* input alignment the record is a projection of */
  const uint32_t* tid;
  const uint32_t* pos;                  /* fwpos or rcpos by transcript strand (src/core.cpp:153-158) */
  const int8_t*   tx_strand;            /* '+' or '-' */
  const uint8_t*  flags;                /* BRB_REC_* */
  const double*   sim_score;            /* AlignInfo::similarity_score; NULL when filter_by_similarity == 0 and no clip rescue
                                         * ran: every score is then 0.0, as in the reference (src/evaluate.cpp:843-865) */
  const int32_t*  clip_score;           /* AlignInfo::clip_score; NULL under the same condition (all 0) */
  const uint32_t* cigar_off;            /* n_rec+1: merged CIGAR of row r = cigar[cigar_off[r] .. cigar_off[r+1]),
                                         * ALREADY reversed for '-' transcripts */
  const uint32_t* cigar;
  const uint32_t* aln_rec_off;          /* n_aln: first row of the unit owned by alignment a */
  const uint32_t* aln_rec_cnt;          /* n_aln: its rows (0 for the higher-indexed end of a pair and for units that emit nothing) */
  const uint32_t* packed_rows;          /* BRB_OPT_PACKED_ROWS only (else NULL): 4 words per row, see below; src_aln, tid, pos,
                                         * tx_strand, flags and cigar_off are then NULL */
  /* BRB_OPT_FLUSH_TAGS only (else NULL): the per-name results of the reference's flush() (src/core.cpp:237-332), per row */
  const int32_t*  hi;                   /* HI tag: 1-based counter over the records of the read name in emit order */
  const int32_t*  nh;                   /* NH tag: records of the read name */
  const uint8_t*  mapq;                 /* get_mapq(nh) (src/core.cpp:46-58) */
  const uint8_t*  primary;              /* 1: the record belongs to the name's primary unit, 0: secondary, 2: tie whose library draw would
                                         * have been rejected and repeated (probability k / 2^64): the caller decides that name itself */
  /* BRB_OPT_PACKED_ROWS = 2 only (else NULL; aln_rec_off / aln_rec_cnt are then NULL instead): */
  const uint32_t* batch_rec_off;        /* (n_aln + 31) / 32: first row of the units owned by alignments 32k .. 32k + 31; their blocks follow
                                         * each other in alignment order */
  const uint16_t* aln_rec_cnt16;        /* n_aln: rows of the unit owned by alignment a */
  uint32_t        packed_row_words;     /* 32-bit words per row of packed_rows: 4 (option value 1), 3 (2), 2 (3); 0 when packed_rows is NULL */
} brb_batch_out;

/* BRB_OPT_PACKED_ROWS: the row columns as ONE 16-byte record per row -- a third fewer bytes device->host, which is what
 * bounds the end-to-end rate.  Row r = packed_rows[4r .. 4r+3] = { src_aln, tid, pos, w }; the unit blocks are found
 * through aln_rec_off / aln_rec_cnt as usual.  A merged CIGAR of one word below 2^27 (every plain short-read row:
 * one M) sits in w itself; any other CIGAR is in `cigar` behind a length word: cigar[i] = n, cigar[i+1 .. i+n] =
 * the words, i = BRB_ROW_CIGAR(w).  n_cigar_words counts the pool with its length words. */
/* BRB_OPT_PACKED_ROWS = 2 ("compact", short-read modes): 12 bytes per row, packed_rows[3r .. 3r+2] = { tid, pos, w } -- the source
 * alignment is implied by the unit block the row sits in (BRB_REC_FIRST rows belong to the owner a, the others to mate_idx[a]) --
 * and the unit blocks are found through batch_rec_off / aln_rec_cnt16 (2.1 bytes per alignment instead of 8).  Another quarter
 * off the device->host bytes. */
/* BRB_OPT_PACKED_ROWS = 3 (short-read modes): as 2 with EIGHT bytes per row, one little-endian 64-bit word
 *   tid : 22 | pos : 24 | BRB_REC_* flags : 3 | '-' : 1 | inline : 1 | n : 13
 * inline = 1: the merged CIGAR is the single word (n << 4 | M) -- every plain short-read row.  inline = 0: the CIGAR has n words and sits in
 * `cigar` behind a length word, and the pool entries come in ROW order: the k-th row with inline = 0 owns the k-th entry (a consumer that
 * walks the rows keeps one cursor: entry = cigar[cur + 1 .. cur + 1 + n], cur += n + 1).  A bundle with a tid >= 2^22, a pos >= 2^24 or an
 * n >= 2^13 is refused with BRB_E_UNSUPPORTED (project it with the option at 2). */
#define BRB_ROW8_TID(v)    ((uint32_t)((v) & 0x3fffffu))
#define BRB_ROW8_POS(v)    ((uint32_t)(((v) >> 22) & 0xffffffu))
#define BRB_ROW8_FLAGS(v)  ((uint32_t)(((v) >> 46) & 7u))
#define BRB_ROW8_MINUS(v)  ((uint32_t)(((v) >> 49) & 1u))
#define BRB_ROW8_INLINE(v) ((uint32_t)(((v) >> 50) & 1u))
#define BRB_ROW8_N(v)      ((uint32_t)((v) >> 51))
#define BRB_ROW_FLAGS(w)   ((uint32_t)(w) >> 29)                 /* BRB_REC_* */
#define BRB_ROW_MINUS(w)   (((uint32_t)(w) >> 28) & 1u)          /* tx_strand == '-' */
#define BRB_ROW_INLINE     (1u << 27)                            /* set: BRB_ROW_CIGAR(w) is the CIGAR word itself */
#define BRB_ROW_CIGAR(w)   ((uint32_t)(w) & 0x07ffffffu)

typedef struct brb_stats {
  uint64_t n_launches;                  /* kernels launched by the last brb_project* call */
  float    kernel_ms;                   /* CUDA-event time of those kernels on the ctx stream */
  float    ms_lookup, ms_chain, ms_rescue, ms_finalize, ms_pair, ms_merge;   /* per phase: front kernel (lookup + evaluation; on the
                                         * clip-rescue path lookup | chain | rescue | build), k_back, and -- BRB_OPT_ORDERED_ROWS only -- scan + k_place + k_group_rows */
  uint64_t n_candidates, n_matches, n_rescue, n_rec;
  uint64_t alg_bytes;                   /* algorithmic bytes of the batch (DESIGN.md formula) */
  uint64_t h2d_bytes, d2h_bytes;
  uint64_t alg_bytes_front;             /* the part of alg_bytes that belongs to the front kernel (alignment columns + index hits) */
  uint64_t ksw_cells;                   /* clip rescue: DP cells (query length x target length) of all problems of the call */
  uint32_t host_syncs;                  /* host round trips of the call: 1, plus one per pool that had to grow (the sequence replays) */
  uint32_t reserved_;
  uint64_t ksw_cells_run;               /* ... of the problems whose DP ran: identical problems of one read (same clip, same neighbour exons) share one */
} brb_stats;

int brb_device_count(void);
const char* brb_last_error(void);
const char* brb_version(void);

int brb_index_create(int device, const brb_index_desc* desc, brb_index** out);
int brb_index_destroy(brb_index* idx);

int brb_ctx_create(brb_index* idx, brb_ctx** out);
int brb_ctx_destroy(brb_ctx* ctx);

/* End to end: H2D of the batch, all kernels, D2H of the result, stream synchronised on return. */
int brb_project(brb_ctx* ctx, const brb_params* p, const brb_batch_in* in, brb_batch_out* out);

/* The same split in three so the kernels can be timed with the batch resident in HBM.  brb_batch_upload returns after the
 * copies have completed: the caller's columns may be reused at once. */
int brb_batch_upload(brb_ctx* ctx, const brb_batch_in* in);
int brb_project_resident(brb_ctx* ctx, const brb_params* p);
/* The read exons of the resident batch as the engine holds them (exon_off: n_aln+1 entries; cap = room in exon_st / exon_en; any of the
 * three may be NULL): what a batch uploaded with pos0 instead of exon columns was expanded to on the device. */
int brb_batch_read_exons(brb_ctx* ctx, uint32_t* exon_off, uint32_t* exon_st, uint32_t* exon_en, uint64_t cap, uint64_t* n_exon);
int brb_batch_download(brb_ctx* ctx, brb_batch_out* out);

int brb_ctx_stats(const brb_ctx* ctx, brb_stats* out);
/* Options of a context.  BRB_OPT_GENERIC_FRONT != 0: use the generic lookup/evaluate kernel (every read exon is
 * looked up, as the reference does) even where the short-read kernel (exon 0 only, later exons verified along
 * the transcript's exon chain) applies -- and, in the long-read modes, count EVERY hit of every read exon (by default a
 * later exon only has to have a surviving hit and the scan of its range stops there).  Results are identical;
 * brb_stats.alg_bytes is the reference-defined figure only with this option (it counts the index hits of ALL exons). */
#define BRB_OPT_GENERIC_FRONT 1
/* BRB_OPT_NO_FASTPATH != 0: short-read preset without the fast path for perfectly spliced candidates (every surviving
 * candidate goes through the general evaluation).  Results are identical; for tests and A/B timing. */
#define BRB_OPT_NO_FASTPATH 2
/* BRB_OPT_ORDERED_ROWS != 0: one more pass on the device moves the unit blocks into the reference's emit order
 * (group_rec_off is then valid and aln_rec_off is ascending).  Costs one streaming copy of the rows. */
#define BRB_OPT_ORDERED_ROWS 3
#define BRB_OPT_PACKED_ROWS 4      /* rows as 16-byte records (brb_batch_out::packed_rows); ignored while BRB_OPT_ORDERED_ROWS is set */
/* BRB_OPT_MERGE_WARP != 0: brb_merge_cigars() runs the warp-cooperative merge (the code long-read rows go through inside
 * brb_project) instead of the one-thread-per-pair code.  Results are identical; for the parity tests. */
#define BRB_OPT_MERGE_WARP 5
/* BRB_OPT_FLUSH_TAGS != 0: HI / NH / MAPQ and the primary choice -- incl. the reference's tie-break
 * uniform_int_distribution(mt19937_64(std::hash(read_name))) -- are computed on the device per read-name group
 * (brb_batch_out::hi, nh, mapq, primary).  Needs brb_batch_in::group_name_off / name_bytes; not available together with
 * BRB_OPT_PACKED_ROWS.  A read name is assumed to form ONE group of the bundle (name-collated input). */
#define BRB_OPT_FLUSH_TAGS 6
int brb_ctx_set_option(brb_ctx* ctx, int option, int value);
/* CUDA-event stopwatch on the ctx stream (what bench.py brackets its K timed steps with). */
int brb_ctx_timer_start(brb_ctx* ctx);
int brb_ctx_timer_stop(brb_ctx* ctx, float* ms);   /* records, synchronises, returns elapsed ms */

/* Standalone kernel entry points used by the parity tests (device work, host buffers). */
/* G1/G2: hits of one query per row, filtered like IntervalTree::findOverlapping (src/g2t.cpp:103-257).
 * status: 0 FIRST, 1 MIDDLE, 2 LAST, 3 ONLY.  Writes up to cap hit record indices per query into
 * hits[q*cap..], the full count into n_hits[q]. */
int brb_lookup(brb_ctx* ctx, const brb_params* p, uint32_t n_query, const int32_t* refid, const int8_t* strand,
               const uint32_t* qs, const uint32_t* qe, const uint8_t* status, uint32_t cap,
               uint32_t* hits, uint32_t* n_hits);
/* B1: merge_cigars (src/bam.cpp:113-315,443-472) for n pairs of (real, ideal) CIGARs given as CSR. */
int brb_merge_cigars(brb_ctx* ctx, uint32_t n, const uint32_t* real_off, const uint32_t* real,
                     const uint32_t* ideal_off, const uint32_t* ideal, uint32_t* out_off /* n+1, in: capacity layout */,
                     uint32_t* out_len, uint32_t* out);
/* The same merge WITHOUT any allocation.  No counterpart in the reference: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch of the
 * warp-cooperative merge lives in a fixed __device__ array of the library: 4 MiB (2^20 words) + 256 status bytes, held by every GPU
 * that loads the library until the process exits.
 *   - ALL array arguments are DEVICE pointers owned by the caller (layouts as brb_merge_cigars: CSR offsets of n+1 entries; pair i
 *     writes at d_out + d_out_off[i], room for real + ideal + 2 words, and its length to d_out_len[i]).
 *   - max_real_ops / max_ideal_ops: upper bounds on the op counts of any pair.  A pair needs max_real_ops + 4 * max_ideal_ops + 16
 *     scratch words; beyond 2^20 the call is rejected with BRB_E_ARG (use brb_merge_cigars).  A pair that exceeds the bounds the
 *     caller gave is still merged correctly, serially by one lane; *d_n_serial (device, optional) receives how many did.
 *   - `stream` is a cudaStream_t (NULL: the default stream).  The call only enqueues (one memset of the status words, which must
 *     start at zero, one or two kernels) and returns; it calls no cudaMalloc / cudaFree and synchronises nothing.
 *   - Every call on a device shares the one array, which keeps the last call's contents: two calls that overlap on the GPU (different
 *     streams, or different host threads) race.  Serialise them on one stream or with events. */
int brb_merge_cigars_static(int device, void* stream, uint32_t n, uint32_t max_real_ops, uint32_t max_ideal_ops,
                            const uint32_t* d_real_off, const uint32_t* d_real, const uint32_t* d_ideal_off, const uint32_t* d_ideal,
                            const uint32_t* d_out_off, uint32_t* d_out_len, uint32_t* d_out, uint32_t* d_n_serial);

/* K3/K4: ksw_extz2 with bramble's fixed parameters (src/evaluate.cpp:285-317) for n problems.
 * Sequences are codes 0..4.  cigar_out row i holds up to cigar_cap words (len<<4|op, op 0=M,1=I,2=D). */
int brb_ksw_extend(brb_ctx* ctx, uint32_t n, const uint64_t* q_off, const uint8_t* q, const uint64_t* t_off,
                   const uint8_t* t, int32_t* max, int32_t* score, int32_t* max_q, int32_t* max_t,
                   uint8_t* zdropped, uint32_t cigar_cap, uint32_t* n_cigar, uint32_t* cigar_out);

/* BGZF encoder on the device: `n` bytes in, complete BGZF blocks out (every block covers 0xff00 input bytes, the last one
 * what is left; no EOF marker block) -- what htslib's bgzf_write + deflate do behind the reference's sam_write1
 * (gclib/GSam.h:553,642-645).  LZ77 + fixed Huffman codes per block, CRC-32 and ISIZE filled in; any inflater reads it.
 * `*out` points into pinned memory owned by the ctx, valid until the next call on it. */
int brb_bgzf_compress(brb_ctx* ctx, const uint8_t* data, uint64_t n, const uint8_t** out, uint64_t* out_bytes);

/* BGZF decoder on the device: `n` bytes holding complete BGZF blocks back to back (any piece of a BAM file cut at block boundaries)
 * in, their inflated bytes out -- what htslib's bgzf_read + inflate do behind the reference's sam_read1 (gclib/GSam.h:489-499).  Full
 * RFC 1951 (stored, fixed and dynamic Huffman blocks); CRC-32 and ISIZE of every block are verified.  Malformed input: BRB_E_ARG with
 * the block number in brb_last_error().  `*out` points into pinned memory owned by the ctx, valid until the next call on it. */
int brb_bgzf_inflate(brb_ctx* ctx, const uint8_t* bgzf, uint64_t n, const uint8_t** out, uint64_t* out_bytes);

/* Output records on the device: what the reference's write_to_bam builds per row (src/core.cpp:96-212 with update_cigar,
 * set_mate_info, reverse_complement_bam, set_nh_tag / set_hi_tag / set_as_tag of src/bam.cpp:475-702) from the rows of the
 * ctx's resident result (the last brb_project / brb_project_resident call, plain columns) and the bundle's INPUT records,
 * followed by the BGZF encoder above.  The caller keeps what the reference's flush() decides on the host -- the ORDER of the
 * units and, where it wants, their tag values -- and passes one unit per emitted BamInfo, in output order:
 *   units[4u + 0]  first row of the unit (a paired unit emits rows r and r + 1: read, then mate)
 *   units[4u + 1]  HI of its first record (the mate record gets HI + 1, as in the reference's counter)
 *   units[4u + 2]  NH
 *   units[4u + 3]  MAPQ | primary << 8
 * raw / rec_off / rec_len: body of input alignment a (the bytes behind its block_size word) = raw[rec_off[a] .. + rec_len[a]).
 * The uncompressed record stream (raw_out_bytes bytes, optional) stays in device memory; `*out` = its BGZF blocks in pinned
 * memory owned by the ctx, valid until the next call on it. */
/* Hands the resident result rows of `from` over to `to` (another ctx of the same index; a device-to-device copy of a few MB), so
 * that `from` can project the next bundle while the rows wait for brb_emit_bgzf.  Afterwards `from` holds no result, and the
 * result in `to` can only be read by brb_emit_bgzf (its batch did not come along). */
int brb_ctx_move_result(brb_ctx* from, brb_ctx* to);

typedef struct brb_emit_in {
  const uint8_t*  raw;
  uint64_t        raw_bytes;
  const uint64_t* rec_off;              /* n_aln of the resident batch */
  const uint32_t* rec_len;
  uint32_t        n_units;
  const uint32_t* units;                /* 4 * n_units words */
} brb_emit_in;
int brb_emit_bgzf(brb_ctx* ctx, const brb_params* params, const brb_emit_in* in, const uint8_t** out, uint64_t* out_bytes,
                  uint64_t* raw_out_bytes);

/* ---- host-side helpers (no device work) ---- */
/* Permutation cgranges' IITree::index() applies to one tree: libstdc++ std::sort (unstable) by start
 * only, on the reference's insertion sequence (src/g2t.cpp:57-62).  perm[k] = input index of sorted record k. */
int brb_sort_by_start(uint64_t n, const int32_t* st, uint32_t* perm);
/* Pinned host memory for batch columns (makes the H2D copies of brb_project asynchronous DMA). */
int brb_host_alloc(size_t bytes, void** out);
int brb_host_free(void* p);

#ifdef __cplusplus
}
#endif
#endif
