// TEST INFRASTRUCTURE (oracle shim) -- not product code.
// Functional mini stand-in for the parts of htslib 1.22 (subprojects/htslib.wrap in the
// reference; un-vendored, no network) that the reference's own sources call:
//   gclib/GSam.h:5-7,355-645   gclib/GSam.cpp   src/bam.cpp   src/core.cpp:96-212
//   src/bramble.cpp:552-555    include/evaluate.h:108-126
// Layouts follow the public htslib ABI (bam1_core_t / bam1_t); BGZF is done over zlib.
// Differences that do not reach record content: qnames are not NUL-padded to 4 bytes in
// memory (l_extranul is always 0; the BAM file never holds the padding), CRAM and SAM-text
// records are unsupported (SAM text is parsed for its header only).
#ifndef BRB_SHIM_SAM_H
#define BRB_SHIM_SAM_H
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include "htslib/kstring.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef int64_t hts_pos_t;

typedef struct sam_hdr_t {
  int32_t n_targets;
  uint32_t *target_len;
  char **target_name;
  char *text;
  size_t l_text;
  void *priv;   // name -> tid map
} sam_hdr_t;
typedef sam_hdr_t bam_hdr_t;

typedef struct htsFile {
  uint32_t is_bin:1, is_write:1, is_be:1, is_cram:1, is_bgzf:1, dummy:27;
  void *priv;
} htsFile;
typedef htsFile samFile;

typedef struct bam1_core_t {
  hts_pos_t pos;
  int32_t tid;
  uint16_t bin;
  uint8_t qual;
  uint8_t l_extranul;
  uint16_t flag;
  uint16_t l_qname;
  uint32_t n_cigar;
  int32_t l_qseq;
  int32_t mtid;
  hts_pos_t mpos;
  hts_pos_t isize;
} bam1_core_t;

typedef struct bam1_t {
  bam1_core_t core;
  uint64_t id;
  uint8_t *data;
  int l_data;
  uint32_t m_data;
  uint32_t mempolicy:2, :30;
} bam1_t;

#define BAM_CMATCH      0
#define BAM_CINS        1
#define BAM_CDEL        2
#define BAM_CREF_SKIP   3
#define BAM_CSOFT_CLIP  4
#define BAM_CHARD_CLIP  5
#define BAM_CPAD        6
#define BAM_CEQUAL      7
#define BAM_CDIFF       8
#define BAM_CBACK       9
#define BAM_CIGAR_STR   "MIDNSHP=XB"
#define BAM_CIGAR_SHIFT 4
#define BAM_CIGAR_MASK  0xf
#define BAM_CIGAR_TYPE  0x3C1A7
#define bam_cigar_op(c) ((c)&BAM_CIGAR_MASK)
#define bam_cigar_oplen(c) ((c)>>BAM_CIGAR_SHIFT)
#define bam_cigar_opchr(c) (BAM_CIGAR_STR "??????" [bam_cigar_op(c)])
#define bam_cigar_gen(l, o) ((l)<<BAM_CIGAR_SHIFT|(o))
#define bam_cigar_type(o) (BAM_CIGAR_TYPE>>((o)<<1)&3)

#define BAM_FPAIRED        1
#define BAM_FPROPER_PAIR   2
#define BAM_FUNMAP         4
#define BAM_FMUNMAP        8
#define BAM_FREVERSE      16
#define BAM_FMREVERSE     32
#define BAM_FREAD1        64
#define BAM_FREAD2       128
#define BAM_FSECONDARY   256
#define BAM_FQCFAIL      512
#define BAM_FDUP        1024
#define BAM_FSUPPLEMENTARY 2048

#define bam_is_rev(b) (((b)->core.flag&BAM_FREVERSE) != 0)
#define bam_get_qname(b) ((char*)(b)->data)
#define bam_get_cigar(b) ((uint32_t*)((b)->data + (b)->core.l_qname))
#define bam_get_seq(b)   ((b)->data + ((b)->core.n_cigar<<2) + (b)->core.l_qname)
#define bam_get_qual(b)  ((b)->data + ((b)->core.n_cigar<<2) + (b)->core.l_qname + (((b)->core.l_qseq + 1)>>1))
#define bam_get_aux(b)   ((b)->data + ((b)->core.n_cigar<<2) + (b)->core.l_qname + (((b)->core.l_qseq + 1)>>1) + (b)->core.l_qseq)
#define bam_get_l_aux(b) ((b)->l_data - ((b)->core.n_cigar<<2) - (b)->core.l_qname - (b)->core.l_qseq - (((b)->core.l_qseq + 1)>>1))
#define bam_seqi(s, i) ((s)[(i)>>1] >> ((~(i)&1)<<2) & 0xf)
#define bam_set_seqi(s,i,b) ((s)[(i)>>1] = ((s)[(i)>>1] & (0xf0 >> ((~(i)&1)<<2))) | ((b)<<((~(i)&1)<<2)))

extern const char seq_nt16_str[];
extern const unsigned char seq_nt16_table[256];

enum hts_fmt_option {
  CRAM_OPT_DECODE_MD, CRAM_OPT_PREFIX, CRAM_OPT_VERBOSITY, CRAM_OPT_SEQS_PER_SLICE,
  CRAM_OPT_SLICES_PER_CONTAINER, CRAM_OPT_RANGE, CRAM_OPT_VERSION, CRAM_OPT_EMBED_REF,
  CRAM_OPT_IGNORE_MD5, CRAM_OPT_REFERENCE, CRAM_OPT_MULTI_SEQ_PER_SLICE, CRAM_OPT_NO_REF,
  CRAM_OPT_USE_BZIP2, CRAM_OPT_SHARED_REF, CRAM_OPT_NTHREADS, CRAM_OPT_THREAD_POOL,
  CRAM_OPT_USE_LZMA, CRAM_OPT_USE_RANS, CRAM_OPT_REQUIRED_FIELDS
};
#define SAM_QNAME 0x00000001
#define SAM_FLAG  0x00000002
#define SAM_RNAME 0x00000004
#define SAM_POS   0x00000008
#define SAM_MAPQ  0x00000010
#define SAM_CIGAR 0x00000020
#define SAM_RNEXT 0x00000040
#define SAM_PNEXT 0x00000080
#define SAM_TLEN  0x00000100
#define SAM_SEQ   0x00000200
#define SAM_QUAL  0x00000400
#define SAM_AUX   0x00000800
#define SAM_RGAUX 0x00001000

htsFile *hts_open(const char *fn, const char *mode);
int hts_close(htsFile *fp);
int hts_set_threads(htsFile *fp, int n);
int hts_set_opt(htsFile *fp, enum hts_fmt_option opt, ...);

sam_hdr_t *sam_hdr_read(htsFile *fp);
int sam_hdr_write(htsFile *fp, const sam_hdr_t *h);
sam_hdr_t *sam_hdr_dup(const sam_hdr_t *h);
void sam_hdr_destroy(sam_hdr_t *h);
const char *sam_hdr_str(sam_hdr_t *h);
int sam_hdr_name2tid(sam_hdr_t *h, const char *ref);
int sam_hdr_add_pg(sam_hdr_t *h, const char *name, ...);

bam1_t *bam_init1(void);
void bam_destroy1(bam1_t *b);
bam1_t *bam_dup1(const bam1_t *b);
int bam_set1(bam1_t *bam, size_t l_qname, const char *qname, uint16_t flag, int32_t tid,
             hts_pos_t pos, uint8_t mapq, size_t n_cigar, const uint32_t *cigar,
             int32_t mtid, hts_pos_t mpos, hts_pos_t isize, size_t l_seq, const char *seq,
             const char *qual, size_t l_aux);
int sam_read1(htsFile *fp, sam_hdr_t *h, bam1_t *b);
int sam_write1(htsFile *fp, const sam_hdr_t *h, const bam1_t *b);

uint8_t *bam_aux_get(const bam1_t *b, const char tag[2]);
int bam_aux_del(bam1_t *b, uint8_t *s);
int bam_aux_append(bam1_t *b, const char tag[2], char type, int len, const uint8_t *data);
int bam_aux_update_int(bam1_t *b, const char tag[2], int64_t val);
int64_t bam_aux2i(const uint8_t *s);
double bam_aux2f(const uint8_t *s);
char bam_aux2A(const uint8_t *s);
char *bam_aux2Z(const uint8_t *s);

#ifdef __cplusplus
}
#endif
#endif
