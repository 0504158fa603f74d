// Raw BAM record helpers for the drop-in CLI: field access on the on-disk record body, aux-tag editing with
// htslib's semantics (bam_aux_get = first occurrence, bam_aux_del, bam_aux_append), ingest arithmetic
// (GSamRecord::setupCoordinates gclib/GSam.cpp:197-291, spliceStrand :338-349) and the per-record output
// assembly of write_to_bam (reference src/core.cpp:96-212, src/bam.cpp:502-702).
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

namespace brbhost {

inline uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
inline int32_t rdi32(const uint8_t* p) { int32_t v; memcpy(&v, p, 4); return v; }
inline uint16_t rd16(const uint8_t* p) { uint16_t v; memcpy(&v, p, 2); return v; }
inline void wr32(uint8_t* p, uint32_t v) { memcpy(p, &v, 4); }
inline void wr16(uint8_t* p, uint16_t v) { memcpy(p, &v, 2); }

struct RecView {                         // body of one BAM record (after block_size)
  const uint8_t* b; uint32_t len;
  int32_t tid() const { return rdi32(b); }
  int32_t pos() const { return rdi32(b + 4); }
  uint32_t l_name() const { return b[8]; }
  uint32_t n_cigar() const { return rd16(b + 12); }
  uint32_t flag() const { return rd16(b + 14); }
  int32_t l_seq() const { return rdi32(b + 16); }
  int32_t mtid() const { return rdi32(b + 20); }
  int32_t mpos() const { return rdi32(b + 24); }
  const char* name() const { return (const char*)b + 32; }
  const uint8_t* cigar() const { return b + 32 + l_name(); }
  const uint8_t* seq() const { return cigar() + 4 * n_cigar(); }
  const uint8_t* qual() const { return seq() + (l_seq() + 1) / 2; }
  const uint8_t* aux() const { return qual() + l_seq(); }
  uint32_t l_aux() const { return len - (uint32_t)(aux() - b); }
};

inline int aux_type_size(int t) {
  switch (t) { case 'A': case 'c': case 'C': return 1; case 's': case 'S': return 2; case 'i': case 'I': case 'f': return 4; case 'd': return 8; default: return 0; }
}
// end of the value that starts at the type byte `s`; nullptr if malformed
inline const uint8_t* aux_skip(const uint8_t* s, const uint8_t* end) {
  if (s >= end) return nullptr;
  const int t = *s++;
  if (t == 'Z' || t == 'H') { while (s < end && *s) s++; return s < end ? s + 1 : nullptr; }
  if (t == 'B') { if (end - s < 5) return nullptr; const int sz = aux_type_size(*s); uint32_t n; memcpy(&n, s + 1, 4); s += 5; if (!sz || (size_t)(end - s) < (size_t)sz * n) return nullptr; return s + (size_t)sz * n; }
  const int sz = aux_type_size(t);
  if (!sz || end - s < sz) return nullptr;
  return s + sz;
}
// offset (from aux start) of the TYPE byte of the first `tag`, or -1  (bam_aux_get)
inline long aux_find(const uint8_t* aux, size_t l_aux, const char tag[2]) {
  const uint8_t* s = aux; const uint8_t* end = aux + l_aux;
  while (s && end - s >= 3) {
    if (s[0] == (uint8_t)tag[0] && s[1] == (uint8_t)tag[1]) return (long)(s + 2 - aux);
    s = aux_skip(s + 2, end);
  }
  return -1;
}
inline void aux_del(std::vector<uint8_t>& aux, const char tag[2]) {          // bam_aux_get + bam_aux_del
  const long o = aux_find(aux.data(), aux.size(), tag);
  if (o < 0) return;
  const uint8_t* e = aux_skip(aux.data() + o, aux.data() + aux.size());
  if (!e) return;
  aux.erase(aux.begin() + (o - 2), aux.begin() + (e - aux.data()));
}
inline void aux_append_i32(std::vector<uint8_t>& aux, const char tag[2], int32_t v) {
  const size_t o = aux.size(); aux.resize(o + 7);
  aux[o] = (uint8_t)tag[0]; aux[o + 1] = (uint8_t)tag[1]; aux[o + 2] = 'i'; memcpy(&aux[o + 3], &v, 4);
}
inline int64_t aux_to_int(const uint8_t* s) {                                 // bam_aux2i
  const int t = *s++;
  switch (t) {
    case 'c': return (int8_t)*s; case 'C': return *s;
    case 's': { int16_t v; memcpy(&v, s, 2); return v; } case 'S': { uint16_t v; memcpy(&v, s, 2); return v; }
    case 'i': { int32_t v; memcpy(&v, s, 4); return v; } case 'I': { uint32_t v; memcpy(&v, s, 4); return v; }
    default: return 0;
  }
}
// tag_char1 (gclib/GSam.cpp:316-325): first char of an A or Z tag, else 0
inline char aux_char1(const RecView& r, const char tag[2]) {
  const long o = aux_find(r.aux(), r.l_aux(), tag);
  if (o < 0) return 0;
  const uint8_t* s = r.aux() + o;
  return (s[0] == 'A' || s[0] == 'Z') ? (char)s[1] : 0;
}

inline char splice_strand(const RecView& r) {                                 // GSamRecord::spliceStrand
  char c = aux_char1(r, "XS");
  if (c == 0) {
    const char m = aux_char1(r, "ts");
    if (m == '+' || m == '-') c = (r.flag() & 16) ? (m == '+' ? '-' : '+') : m;
  }
  return (c == '+' || c == '-') ? c : '.';
}

inline char get_strand(const RecView& r, bool fr, bool rf) {                  // src/bramble.cpp:213-244
  char strand = splice_strand(r);
  if (strand == '.' && (fr || rf)) {
    const bool rev = r.flag() & 16, hit = (rf && rev) || (fr && !rev);
    if (r.flag() & 1) {
      const int order = (r.flag() & 0x40) ? 1 : ((r.flag() & 0x80) ? 2 : 0);
      if (order == 1) strand = hit ? '-' : '+'; else strand = hit ? '+' : '-';
    } else strand = hit ? '-' : '+';
  }
  return strand;
}

// setupCoordinates + process_exons' end++ : appends [start,end) read exons; returns false for "invalid CIGAR"
inline bool read_exons(const RecView& r, std::vector<uint32_t>& es, std::vector<uint32_t>& ee) {
  const int32_t pos0 = r.pos(); uint32_t l = 0; int32_t exstart = pos0; uint32_t ex_s = 0, ex_e = 0;
  bool started = false, intron = false, ins = false;
  const uint8_t* cg = r.cigar();
  for (uint32_t i = 0; i < r.n_cigar(); i++) {
    const uint32_t w = rd32(cg + 4 * i), op = w & 0xf, len = w >> 4;
    switch (op) {
      case 0: case 7: case 8: started = true; l += len; intron = false; ins = false; break;
      case 2: l += len; ins = false; break;
      case 1: ins = true; break;
      case 3:
        if (!started) break;
        if (!ins || !intron) { ex_e = (uint32_t)(pos0 + (int32_t)l); ex_s = (uint32_t)(exstart + 1); es.push_back(ex_s); ee.push_back(ex_e + 1); }
        l += len; exstart = pos0 + (int32_t)l; intron = true; break;
      case 4: case 5: ins = false; break;
      default: break;
    }
  }
  if (!intron) { ex_s = (uint32_t)(exstart + 1); ex_e = (uint32_t)(pos0 + (int32_t)l); es.push_back(ex_s); ee.push_back(ex_e + 1); }
  return ex_e != 0;
}

struct OutInfo {                         // everything write_to_bam knows about one output record
  uint32_t tid, pos; char strand; bool paired, same_tx, primary, long_reads; uint32_t mapq; int32_t nh, hi;
  double sim; int32_t clip; uint32_t mate_tid, mate_pos; const uint32_t* cigar; uint32_t n_cigar;
};

// One output record (block_size + body) appended to `out`.
inline void assemble_record(const RecView& r, const OutInfo& o, std::vector<uint8_t>& aux, std::vector<uint8_t>& out) {
  // source-record tag edits (once per read in the reference, idempotent here): NH re-set, XS/ts dropped (src/core.cpp:115-124)
  aux.assign(r.aux(), r.aux() + r.l_aux());
  aux_del(aux, "NH"); aux_append_i32(aux, "NH", o.nh);
  aux_del(aux, o.long_reads ? "ts" : "XS");
  if (o.long_reads) {                                                          // set_as_tag, src/bam.cpp:622-634
    const long ao = aux_find(aux.data(), aux.size(), "AS");
    int32_t gn_as = 0;
    if (ao >= 0) { gn_as = (int32_t)aux_to_int(aux.data() + ao); aux_del(aux, "AS"); }
    const double score = (static_cast<double>(gn_as) + static_cast<double>(o.clip)) * o.sim;
    aux_append_i32(aux, "AS", static_cast<int32_t>(score));
  }
  aux_del(aux, "HI"); aux_append_i32(aux, "HI", o.hi);                         // set_hi_tag
  const uint32_t lname = r.l_name(); const int32_t lq = r.l_seq(); const uint32_t seqb = (uint32_t)((lq + 1) / 2);
  const uint32_t body = 32 + lname + 4 * o.n_cigar + seqb + (uint32_t)lq + (uint32_t)aux.size();
  const size_t base = out.size(); out.resize(base + 4 + body);
  uint8_t* b = out.data() + base + 4;
  wr32(out.data() + base, body);
  memcpy(b, r.b, 32 + lname);
  uint32_t flag = r.flag();
  flag = o.primary ? (flag & ~0x100u) : (flag | 0x100u);
  memcpy(b + 32 + lname, o.cigar, 4 * (size_t)o.n_cigar);                      // already merged (and reversed for '-') on the GPU
  uint8_t* seq = b + 32 + lname + 4 * o.n_cigar; uint8_t* qual = seq + seqb;
  if (o.strand == '-') {                                                       // reverse_complement_bam, src/bam.cpp:636-702
    if (lq > 0) {
      static const uint8_t comp[16] = {15, 8, 4, 15, 2, 15, 15, 15, 1, 15, 15, 15, 15, 15, 15, 15};
      const uint8_t* s = r.seq();
      memset(seq, 0, seqb);
      for (int32_t i = 0; i < lq; i++) {
        const int32_t j = lq - 1 - i; const uint8_t nt = (s[j >> 1] >> ((~j & 1) << 2)) & 0xf;
        seq[i >> 1] |= comp[nt] << ((~i & 1) << 2);
      }
      const uint8_t* q = r.qual();
      if (q[0] != 0xff) for (int32_t i = 0; i < lq; i++) qual[i] = q[lq - 1 - i]; else memcpy(qual, q, lq);
    }
    flag ^= 16;
  } else { memcpy(seq, r.seq(), seqb + (size_t)lq); }
  memcpy(qual + lq, aux.data(), aux.size());
  wr32(b, o.tid); wr32(b + 4, o.pos); b[9] = (uint8_t)o.mapq; wr16(b + 12, (uint16_t)o.n_cigar);
  int32_t mtid, mpos, isize;                                                   // set_mate_info, src/bam.cpp:531-588
  if (!o.paired) { flag &= ~(1u | 2u | 32u); mtid = -1; mpos = -1; isize = 0; }
  else {
    flag |= 1u;
    if (o.strand == '-') flag |= 32u;                                          // MREVERSE from the record's OWN transcript strand
    if (o.same_tx) {
      mtid = (int32_t)o.tid; mpos = (int32_t)o.mate_pos; flag |= 2u;
      const int32_t my = (int32_t)o.pos, mate = (int32_t)o.mate_pos;
      isize = my <= mate ? (mate + lq) - my : -((my + lq) - mate);
    } else { mtid = (int32_t)o.mate_tid; mpos = (int32_t)o.mate_pos; isize = 0; flag &= ~2u; }
  }
  wr16(b + 14, (uint16_t)flag); wr32(b + 20, (uint32_t)mtid); wr32(b + 24, (uint32_t)mpos); wr32(b + 28, (uint32_t)isize);
}

}  // namespace brbhost
