// TEST INFRASTRUCTURE (oracle shim) -- not product code.
// Implementation of the mini htslib declared in oracle/shim/htslib/sam.h: BGZF over zlib,
// BAM record and header I/O, aux-tag editing, @PG chaining.  Behaviour follows htslib 1.22
// as far as record/header CONTENT goes (bam_read1/bam_write1, sam_hdr_add_pg, bam_aux_*).
// Env knobs (shim only): BRB_SHIM_NULL_WRITE=1 makes sam_write1 a no-op (projection-only
// timing); BRB_SHIM_LEVEL=<0-9> sets the deflate level (default 6, like htslib/zlib).
#include "htslib/sam.h"
#include <zlib.h>
#include <stdio.h>
#include <stdarg.h>
#include <string>
#include <vector>
#include <unordered_map>

const char seq_nt16_str[] = "=ACMGRSVTWYHKDBN";
const unsigned char seq_nt16_table[256] = {
  15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15,
  15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 1,2,4,8, 15,15,15,15, 15,15,15,15, 15, 0,15,15,
  15, 1,14, 2, 13,15,15, 4, 11,15,15,12, 15, 3,15,15, 15,15, 5, 6,  8,15, 7, 9, 15,10,15,15, 15,15,15,15,
  15, 1,14, 2, 13,15,15, 4, 11,15,15,12, 15, 3,15,15, 15,15, 5, 6,  8,15, 7, 9, 15,10,15,15, 15,15,15,15,
  15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15,
  15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15,
  15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15,
  15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15, 15,15,15,15 };

namespace {

const size_t BGZF_BLOCK = 0xff00;

struct Stream {
  FILE *f = nullptr;
  bool write = false, bgzf = false;
  // reader state
  std::vector<uint8_t> in;  size_t in_pos = 0;
  // writer state
  std::vector<uint8_t> out;
  int level = 6;
  bool null_write = false;

  bool fill() {  // read one BGZF block into `in`
    uint8_t h[12];
    for (;;) {
      if (fread(h, 1, 12, f) != 12) return false;
      if (h[0] != 0x1f || h[1] != 0x8b) return false;
      unsigned xlen = h[10] | (h[11] << 8);
      std::vector<uint8_t> extra(xlen);
      if (xlen && fread(extra.data(), 1, xlen, f) != xlen) return false;
      unsigned bsize = 0; bool found = false;
      for (size_t p = 0; p + 4 <= xlen;) {
        unsigned slen = extra[p + 2] | (extra[p + 3] << 8);
        if (extra[p] == 'B' && extra[p + 1] == 'C' && slen == 2) { bsize = extra[p + 4] | (extra[p + 5] << 8); found = true; }
        p += 4 + slen;
      }
      if (!found) return false;
      size_t clen = (size_t)bsize + 1 - xlen - 12 - 8;
      std::vector<uint8_t> c(clen + 8);
      if (fread(c.data(), 1, clen + 8, f) != clen + 8) return false;
      uint32_t isize = c[clen + 4] | (c[clen + 5] << 8) | (c[clen + 6] << 16) | ((uint32_t)c[clen + 7] << 24);
      in.resize(isize); in_pos = 0;
      if (isize == 0) continue;  // empty block (e.g. EOF marker): try the next one
      z_stream zs; memset(&zs, 0, sizeof zs);
      if (inflateInit2(&zs, -15) != Z_OK) return false;
      zs.next_in = c.data(); zs.avail_in = (uInt)clen;
      zs.next_out = in.data(); zs.avail_out = isize;
      int rc = inflate(&zs, Z_FINISH);
      inflateEnd(&zs);
      if (rc != Z_STREAM_END) return false;
      return true;
    }
  }
  size_t read(void *dst, size_t n) {
    uint8_t *d = (uint8_t *)dst; size_t got = 0;
    while (got < n) {
      if (in_pos == in.size() && !fill()) break;
      size_t k = in.size() - in_pos; if (k > n - got) k = n - got;
      memcpy(d + got, in.data() + in_pos, k); in_pos += k; got += k;
    }
    return got;
  }
  void flush_block() {
    size_t n = out.size() < BGZF_BLOCK ? out.size() : BGZF_BLOCK;
    uLong bound = compressBound((uLong)n) + 64;
    std::vector<uint8_t> c(bound);
    z_stream zs; memset(&zs, 0, sizeof zs);
    deflateInit2(&zs, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY);
    zs.next_in = out.data(); zs.avail_in = (uInt)n;
    zs.next_out = c.data(); zs.avail_out = (uInt)bound;
    deflate(&zs, Z_FINISH);
    size_t clen = zs.total_out;
    deflateEnd(&zs);
    uint32_t crc = (uint32_t)crc32(crc32(0L, Z_NULL, 0), out.data(), (uInt)n);
    unsigned bsize = (unsigned)(clen + 25);
    uint8_t h[18] = {0x1f, 0x8b, 8, 4, 0, 0, 0, 0, 0, 0xff, 6, 0, 'B', 'C', 2, 0,
                     (uint8_t)(bsize & 0xff), (uint8_t)(bsize >> 8)};
    uint8_t t[8] = {(uint8_t)crc, (uint8_t)(crc >> 8), (uint8_t)(crc >> 16), (uint8_t)(crc >> 24),
                    (uint8_t)n, (uint8_t)(n >> 8), (uint8_t)(n >> 16), (uint8_t)(n >> 24)};
    fwrite(h, 1, 18, f); fwrite(c.data(), 1, clen, f); fwrite(t, 1, 8, f);
    out.erase(out.begin(), out.begin() + n);
  }
  void put(const void *src, size_t n) {
    const uint8_t *s = (const uint8_t *)src;
    out.insert(out.end(), s, s + n);
    while (out.size() >= BGZF_BLOCK) flush_block();
  }
  void flush_try(size_t n) { if (out.size() + n > BGZF_BLOCK && !out.empty()) flush_block(); }
};

typedef std::unordered_map<std::string, int> NameMap;

void hdr_build_map(sam_hdr_t *h) {
  NameMap *m = new NameMap();
  for (int i = 0; i < h->n_targets; i++) m->emplace(h->target_name[i], i);  // first wins
  delete (NameMap *)h->priv;
  h->priv = m;
}

sam_hdr_t *hdr_new() { return (sam_hdr_t *)calloc(1, sizeof(sam_hdr_t)); }

void hdr_set_text(sam_hdr_t *h, const std::string &t) {
  free(h->text);
  h->text = (char *)malloc(t.size() + 1);
  memcpy(h->text, t.data(), t.size()); h->text[t.size()] = 0; h->l_text = t.size();
}

std::string field_of(const std::string &line, const char *key) {  // key like "ID:"
  size_t p = 0;
  while ((p = line.find('\t', p)) != std::string::npos) {
    ++p;
    if (line.compare(p, strlen(key), key) == 0) {
      size_t e = line.find('\t', p);
      return line.substr(p + strlen(key), (e == std::string::npos ? line.size() : e) - p - strlen(key));
    }
  }
  return std::string();
}

int aux_type_size(int t) {
  switch (t) { case 'A': case 'c': case 'C': return 1; case 's': case 'S': return 2;
    case 'i': case 'I': case 'f': return 4; case 'd': return 8; default: return 0; }
}
// s points at the type byte; returns pointer past the value, or NULL if malformed
const uint8_t *aux_skip(const uint8_t *s, const uint8_t *end) {
  if (s >= end) return NULL;
  int t = *s++;
  int sz = aux_type_size(t);
  if (sz) return s + sz <= end ? s + sz : NULL;
  if (t == 'Z' || t == 'H') { while (s < end && *s) ++s; return s < end ? s + 1 : NULL; }
  if (t == 'B') {
    if (s + 5 > end) return NULL;
    int es = aux_type_size(*s); uint32_t n; memcpy(&n, s + 1, 4);
    s += 5 + (size_t)es * n; return (es && s <= end) ? s : NULL;
  }
  return NULL;
}
void ensure(bam1_t *b, size_t need) {
  if (b->m_data < need) { uint32_t m = (uint32_t)need; kroundup32(m); b->data = (uint8_t *)realloc(b->data, m); b->m_data = m; }
}
}  // namespace

extern "C" {

htsFile *hts_open(const char *fn, const char *mode) {
  htsFile *fp = (htsFile *)calloc(1, sizeof(htsFile));
  Stream *s = new Stream();
  fp->priv = s;
  if (mode[0] == 'w') {
    s->write = true; s->bgzf = strchr(mode, 'b') != NULL;
    s->f = fopen(fn, "wb");
    const char *lv = getenv("BRB_SHIM_LEVEL"); if (lv) s->level = atoi(lv);
    if (strchr(mode, 'u')) s->level = 0;
    const char *nw = getenv("BRB_SHIM_NULL_WRITE"); s->null_write = nw && nw[0] == '1';
    fp->is_write = 1; fp->is_bin = fp->is_bgzf = s->bgzf;
  } else {
    s->f = fopen(fn, "rb");
    if (s->f) {
      int c0 = fgetc(s->f), c1 = fgetc(s->f);
      s->bgzf = (c0 == 0x1f && c1 == 0x8b);
      fseek(s->f, 0, SEEK_SET);
      fp->is_bin = fp->is_bgzf = s->bgzf;
    }
  }
  if (!s->f) { delete s; free(fp); return NULL; }
  return fp;
}

int hts_close(htsFile *fp) {
  if (!fp) return 0;
  Stream *s = (Stream *)fp->priv;
  if (s->write && s->bgzf) {
    while (!s->out.empty()) s->flush_block();
    static const uint8_t eof[28] = {0x1f,0x8b,8,4,0,0,0,0,0,0xff,6,0,'B','C',2,0,0x1b,0,3,0,0,0,0,0,0,0,0,0};
    fwrite(eof, 1, 28, s->f);
  } else if (s->write) {
    fwrite(s->out.data(), 1, s->out.size(), s->f);
  }
  fclose(s->f); delete s; free(fp);
  return 0;
}
int hts_set_threads(htsFile *, int) { return 0; }
int hts_set_opt(htsFile *, enum hts_fmt_option, ...) { return 0; }

sam_hdr_t *sam_hdr_read(htsFile *fp) {
  Stream *s = (Stream *)fp->priv;
  sam_hdr_t *h = hdr_new();
  if (s->bgzf) {
    char magic[4]; int32_t l_text, n_ref;
    if (s->read(magic, 4) != 4 || memcmp(magic, "BAM\1", 4)) { free(h); return NULL; }
    s->read(&l_text, 4);
    std::string text(l_text, 0);
    if (l_text) s->read(&text[0], l_text);
    while (!text.empty() && text.back() == 0) text.pop_back();
    hdr_set_text(h, text);
    s->read(&n_ref, 4);
    h->n_targets = n_ref;
    h->target_name = (char **)calloc(n_ref ? n_ref : 1, sizeof(char *));
    h->target_len = (uint32_t *)calloc(n_ref ? n_ref : 1, 4);
    for (int i = 0; i < n_ref; i++) {
      int32_t l_name; s->read(&l_name, 4);
      h->target_name[i] = (char *)calloc(l_name + 1, 1);
      s->read(h->target_name[i], l_name);
      s->read(&h->target_len[i], 4);
    }
  } else {  // SAM text: header lines only
    std::string text, line; int c;
    std::vector<std::pair<std::string, uint32_t>> sq;
    while ((c = fgetc(s->f)) != EOF) {
      if (line.empty() && c != '@') { ungetc(c, s->f); break; }
      line.push_back((char)c);
      if (c == '\n') {
        if (line.compare(0, 3, "@SQ") == 0) {
          std::string l2 = line.substr(0, line.size() - 1);
          sq.emplace_back(field_of(l2, "SN:"), (uint32_t)strtoul(field_of(l2, "LN:").c_str(), NULL, 10));
        }
        text += line; line.clear();
      }
    }
    hdr_set_text(h, text);
    h->n_targets = (int32_t)sq.size();
    h->target_name = (char **)calloc(sq.size() ? sq.size() : 1, sizeof(char *));
    h->target_len = (uint32_t *)calloc(sq.size() ? sq.size() : 1, 4);
    for (size_t i = 0; i < sq.size(); i++) { h->target_name[i] = strdup(sq[i].first.c_str()); h->target_len[i] = sq[i].second; }
  }
  hdr_build_map(h);
  return h;
}

int sam_hdr_write(htsFile *fp, const sam_hdr_t *h) {
  Stream *s = (Stream *)fp->priv;
  if (!s->bgzf) { s->put(h->text, h->l_text); return 0; }
  s->put("BAM\1", 4);
  int32_t l_text = (int32_t)h->l_text; s->put(&l_text, 4); s->put(h->text, h->l_text);
  int32_t n = h->n_targets; s->put(&n, 4);
  for (int i = 0; i < n; i++) {
    int32_t l = (int32_t)strlen(h->target_name[i]) + 1;
    s->put(&l, 4); s->put(h->target_name[i], l); s->put(&h->target_len[i], 4);
  }
  while (!s->out.empty()) s->flush_block();  // htslib flushes after the header
  return 0;
}

sam_hdr_t *sam_hdr_dup(const sam_hdr_t *h0) {
  sam_hdr_t *h = hdr_new();
  hdr_set_text(h, std::string(h0->text ? h0->text : "", h0->l_text));
  h->n_targets = h0->n_targets;
  h->target_name = (char **)calloc(h->n_targets ? h->n_targets : 1, sizeof(char *));
  h->target_len = (uint32_t *)calloc(h->n_targets ? h->n_targets : 1, 4);
  for (int i = 0; i < h->n_targets; i++) { h->target_name[i] = strdup(h0->target_name[i]); h->target_len[i] = h0->target_len[i]; }
  hdr_build_map(h);
  return h;
}

void sam_hdr_destroy(sam_hdr_t *h) {
  if (!h) return;
  for (int i = 0; i < h->n_targets; i++) free(h->target_name[i]);
  free(h->target_name); free(h->target_len); free(h->text);
  delete (NameMap *)h->priv;
  free(h);
}

const char *sam_hdr_str(sam_hdr_t *h) { return h->text ? h->text : ""; }

int sam_hdr_name2tid(sam_hdr_t *h, const char *ref) {
  NameMap *m = (NameMap *)h->priv;
  auto it = m->find(ref);
  return it == m->end() ? -1 : it->second;
}

// htslib sam_hdr_add_pg: unique ID (name, name.1, ...), PP chained to the end of every
// existing @PG chain (one new line per chain end), new line(s) appended after all others.
// Field order: ID, PN, PP, then the caller's key/value pairs.
int sam_hdr_add_pg(sam_hdr_t *h, const char *name, ...) {
  std::string text(h->text ? h->text : "", h->l_text);
  std::vector<std::string> ids, pps;
  for (size_t p = 0; p < text.size();) {
    size_t e = text.find('\n', p); if (e == std::string::npos) e = text.size();
    std::string line = text.substr(p, e - p);
    if (line.compare(0, 3, "@PG") == 0) { ids.push_back(field_of(line, "ID:")); pps.push_back(field_of(line, "PP:")); }
    p = e + 1;
  }
  std::vector<std::string> ends;
  for (size_t i = 0; i < ids.size(); i++) {
    bool referenced = false;
    for (size_t j = 0; j < pps.size(); j++) if (j != i && pps[j] == ids[i]) referenced = true;
    if (!referenced) ends.push_back(ids[i]);
  }
  std::string extra;
  va_list ap; va_start(ap, name);
  for (;;) {
    const char *k = va_arg(ap, const char *); if (!k) break;
    const char *v = va_arg(ap, const char *); if (!v) break;
    extra += "\t"; extra += k; extra += ":"; extra += v;
  }
  va_end(ap);
  auto unique_id = [&](void) {
    std::string id = name; int n = 0;
    for (;;) {
      bool taken = false;
      for (auto &x : ids) if (x == id) taken = true;
      if (!taken) break;
      id = std::string(name) + "." + std::to_string(++n);
    }
    ids.push_back(id);
    return id;
  };
  if (!text.empty() && text.back() != '\n') text += "\n";
  if (ends.empty()) {
    std::string id = unique_id();
    text += "@PG\tID:" + id + "\tPN:" + name + extra + "\n";
  } else {
    for (auto &pp : ends) {
      std::string id = unique_id();
      text += "@PG\tID:" + id + "\tPN:" + name + "\tPP:" + pp + extra + "\n";
    }
  }
  hdr_set_text(h, text);
  return 0;
}

bam1_t *bam_init1(void) { return (bam1_t *)calloc(1, sizeof(bam1_t)); }
void bam_destroy1(bam1_t *b) { if (b) { free(b->data); free(b); } }
bam1_t *bam_dup1(const bam1_t *b0) {
  bam1_t *b = bam_init1();
  b->core = b0->core; b->id = b0->id; b->l_data = b0->l_data;
  ensure(b, b0->l_data > 0 ? b0->l_data : 1);
  memcpy(b->data, b0->data, b0->l_data);
  return b;
}

int bam_set1(bam1_t *b, size_t l_qname, const char *qname, uint16_t flag, int32_t tid,
             hts_pos_t pos, uint8_t mapq, size_t n_cigar, const uint32_t *cigar,
             int32_t mtid, hts_pos_t mpos, hts_pos_t isize, size_t l_seq, const char *seq,
             const char *qual, size_t l_aux) {
  size_t need = l_qname + 1 + n_cigar * 4 + (l_seq + 1) / 2 + l_seq + l_aux;
  ensure(b, need);
  b->core.pos = pos; b->core.tid = tid; b->core.bin = 0; b->core.qual = mapq; b->core.l_extranul = 0;
  b->core.flag = flag; b->core.l_qname = (uint16_t)(l_qname + 1); b->core.n_cigar = (uint32_t)n_cigar;
  b->core.l_qseq = (int32_t)l_seq; b->core.mtid = mtid; b->core.mpos = mpos; b->core.isize = isize;
  uint8_t *p = b->data;
  memcpy(p, qname, l_qname); p[l_qname] = 0; p += l_qname + 1;
  if (n_cigar) memcpy(p, cigar, n_cigar * 4);
  p += n_cigar * 4;
  memset(p, 0, (l_seq + 1) / 2);
  for (size_t i = 0; i < l_seq; i++) p[i >> 1] |= seq_nt16_table[(uint8_t)seq[i]] << ((~i & 1) << 2);
  p += (l_seq + 1) / 2;
  if (qual) for (size_t i = 0; i < l_seq; i++) p[i] = qual[i] - 33; else memset(p, 0xff, l_seq);
  b->l_data = (int)(need - l_aux);
  return b->l_data;
}

int sam_read1(htsFile *fp, sam_hdr_t *, bam1_t *b) {
  Stream *s = (Stream *)fp->priv;
  if (!s->bgzf) return -1;
  int32_t block_len; uint32_t x[8];
  if (s->read(&block_len, 4) != 4) return -1;
  if (s->read(x, 32) != 32) return -2;
  bam1_core_t *c = &b->core;
  c->tid = x[0]; c->pos = (int32_t)x[1];
  c->bin = x[2] >> 16; c->qual = x[2] >> 8 & 0xff; c->l_qname = x[2] & 0xff; c->l_extranul = 0;
  c->flag = x[3] >> 16; c->n_cigar = x[3] & 0xffff;
  c->l_qseq = x[4]; c->mtid = x[5]; c->mpos = (int32_t)x[6]; c->isize = (int32_t)x[7];
  b->l_data = block_len - 32;
  ensure(b, b->l_data > 0 ? b->l_data : 1);
  if (s->read(b->data, b->l_data) != (size_t)b->l_data) return -3;
  return 4 + block_len;
}

int sam_write1(htsFile *fp, const sam_hdr_t *, const bam1_t *b) {
  Stream *s = (Stream *)fp->priv;
  if (s->null_write) return 0;
  const bam1_core_t *c = &b->core;
  uint32_t x[8], block_len = b->l_data - c->l_extranul + 32;
  x[0] = c->tid; x[1] = (uint32_t)c->pos;
  x[2] = (uint32_t)c->bin << 16 | c->qual << 8 | (c->l_qname - c->l_extranul);
  x[3] = (uint32_t)c->flag << 16 | (c->n_cigar & 0xffff);
  x[4] = c->l_qseq; x[5] = c->mtid; x[6] = (uint32_t)c->mpos; x[7] = (uint32_t)c->isize;
  s->flush_try(4 + block_len);
  s->put(&block_len, 4); s->put(x, 32);
  s->put(b->data, c->l_qname - c->l_extranul);
  s->put(b->data + c->l_qname, b->l_data - c->l_qname);
  return 4 + block_len;
}

uint8_t *bam_aux_get(const bam1_t *b, const char tag[2]) {
  const uint8_t *s = bam_get_aux(b), *end = b->data + b->l_data;
  while (s && end - s >= 3) {
    if (s[0] == (uint8_t)tag[0] && s[1] == (uint8_t)tag[1]) return (uint8_t *)s + 2;
    s = aux_skip(s + 2, end);
  }
  return NULL;
}
int bam_aux_del(bam1_t *b, uint8_t *s) {
  uint8_t *end = b->data + b->l_data;
  const uint8_t *next = aux_skip(s, end);
  if (!next) return -1;
  uint8_t *start = s - 2;
  memmove(start, next, end - next);
  b->l_data -= (int)(next - start);
  return 0;
}
int bam_aux_append(bam1_t *b, const char tag[2], char type, int len, const uint8_t *data) {
  ensure(b, (size_t)b->l_data + 3 + len);
  uint8_t *p = b->data + b->l_data;
  p[0] = tag[0]; p[1] = tag[1]; p[2] = type; memcpy(p + 3, data, len);
  b->l_data += 3 + len;
  return 0;
}
int bam_aux_update_int(bam1_t *b, const char tag[2], int64_t val) {
  uint8_t *s = bam_aux_get(b, tag);
  if (s) bam_aux_del(b, s);
  int32_t v = (int32_t)val;
  return bam_aux_append(b, tag, 'i', 4, (const uint8_t *)&v);
}
int64_t bam_aux2i(const uint8_t *s) {
  int t = *s++;
  switch (t) {
    case 'c': return (int8_t)*s; case 'C': return *s;
    case 's': { int16_t v; memcpy(&v, s, 2); return v; } case 'S': { uint16_t v; memcpy(&v, s, 2); return v; }
    case 'i': { int32_t v; memcpy(&v, s, 4); return v; } case 'I': { uint32_t v; memcpy(&v, s, 4); return v; }
  }
  return 0;
}
double bam_aux2f(const uint8_t *s) {
  int t = *s++;
  if (t == 'f') { float v; memcpy(&v, s, 4); return v; }
  if (t == 'd') { double v; memcpy(&v, s, 8); return v; }
  return (double)bam_aux2i(s - 1);
}
char bam_aux2A(const uint8_t *s) { int t = *s++; return t == 'A' ? (char)*s : 0; }
char *bam_aux2Z(const uint8_t *s) { int t = *s++; return (t == 'Z' || t == 'H') ? (char *)s : NULL; }

}  // extern "C"
