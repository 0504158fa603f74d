// Annotation loading for the drop-in CLI: GTF/GFF3 transcripts -> the reference's transcript order, output
// header @SQ lines and the flattened g2t table (brb_index_desc).  Restates what the reference obtains from
// gclib: GffReader::readAll(keepAttr, mergeCloseExons=false) with transcripts only + setRefAlphaSorted()
// (reference src/bramble.cpp:528-535): exons of a transcript sorted and overlapping/adjacent ones merged
// (gclib/gff.cpp:988-1079), transcripts ordered by gfo_cmpByLoc (gclib/gff.cpp:75-91: chromosome NAME strcmp,
// start, level, end, ID), refid = order of first appearance of the chromosome in the file; then
// build_g2t_tree (src/bramble.cpp:132-211).  Supported input: GTF (GENCODE / Ensembl / StringTie style: gene, transcript,
// exon, CDS, UTR, codon rows; transcript rows optional) and GFF3 with ID=/Parent= on gene, mRNA|transcript|*RNA and exon
// rows.  Two gclib details that decide the transcript ORDER (hence tids and the @SQ lines) are kept:
//   * gff_level (gclib/gff.cpp:1436-1450): a transcript whose ROW names a parent gene that was seen earlier in the file is
//     level 1, every other transcript (no gene rows in the file, or built from exon rows only) level 0 -- and the level
//     sorts before the end coordinate and the ID (gfo_cmpByLoc);
//   * an exon row with several parents (GFF3) is added to the parents already seen; if none is, only the first one is
//     created from it (gclib/gff.cpp:1743-1760).
// Not handled: CDS-only models (no exon rows), trans-spliced / discontinuous features with a repeated ID.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/bramble_b200.h"

namespace brbhost {

struct Transcript {
  std::string id, chrom; int refid = -1; char strand = '.';
  uint32_t start = 0, end = 0; int level = 0; bool has_line = false;
  std::vector<std::pair<uint32_t, uint32_t>> exons;     // 1-based inclusive, ascending after finalize()
  uint32_t tlen = 0;
};

struct Annotation {
  std::vector<std::string> chrom_names;                 // refid order (first appearance in the GTF)
  std::unordered_map<std::string, int> refid_of;
  std::vector<Transcript> tx;                           // tid order after load()
};

inline std::string attr_value(const std::string& attrs, const char* key, bool gff3) {
  const size_t kl = strlen(key);
  size_t p = 0;
  while (p < attrs.size()) {
    while (p < attrs.size() && (attrs[p] == ' ' || attrs[p] == ';' || attrs[p] == '\t')) p++;
    size_t e = p; bool inq = false;
    while (e < attrs.size() && (inq || attrs[e] != ';')) { if (attrs[e] == '"') inq = !inq; e++; }
    if (attrs.compare(p, kl, key) == 0 && p + kl < e && (attrs[p + kl] == ' ' || attrs[p + kl] == '=')) {
      size_t v = p + kl + 1;
      while (v < e && attrs[v] == ' ') v++;
      size_t ve = e;
      while (ve > v && attrs[ve - 1] == ' ') ve--;
      if (v < ve && attrs[v] == '"') { v++; if (ve > v && attrs[ve - 1] == '"') ve--; }
      (void)gff3;
      return attrs.substr(v, ve - v);
    }
    p = e + 1;
  }
  return std::string();
}

inline void load_annotation(const std::string& path, Annotation& A) {
  FILE* f = fopen(path.c_str(), "r");
  if (!f) throw std::runtime_error("could not open reference annotation file " + path);
  std::unordered_map<std::string, size_t> by_id;
  std::unordered_map<std::string, bool> gene_seen;      // IDs of the gene rows read so far (GTF: gene_id; GFF3: ID)
  std::vector<Transcript> tx;
  char* line = nullptr; size_t cap = 0; ssize_t n;
  auto get_tx = [&](const std::string& id, const std::string& chrom, char strand) -> Transcript& {
    auto it = by_id.find(id);
    if (it == by_id.end()) {
      by_id.emplace(id, tx.size()); tx.emplace_back(); Transcript& t = tx.back(); t.id = id; t.chrom = chrom; t.strand = strand;
      auto r = A.refid_of.find(chrom);
      if (r == A.refid_of.end()) { t.refid = (int)A.chrom_names.size(); A.refid_of.emplace(chrom, t.refid); A.chrom_names.push_back(chrom); } else t.refid = r->second;
      return t;
    }
    return tx[it->second];
  };
  while ((n = getline(&line, &cap, f)) > 0) {
    if (line[0] == '#') continue;
    while (n > 0 && (line[n - 1] == '\n' || line[n - 1] == '\r')) line[--n] = 0;
    char* col[9]; int nc = 0; char* p = line;
    while (nc < 9) { col[nc++] = p; char* t = strchr(p, '\t'); if (!t) break; *t = 0; p = t + 1; }
    if (nc < 9) continue;
    const std::string feat = col[2], attrs = col[8];
    const bool gff3 = attrs.find('=') != std::string::npos && attrs.find(" \"") == std::string::npos;
    const uint32_t s = (uint32_t)strtoul(col[3], nullptr, 10), e = (uint32_t)strtoul(col[4], nullptr, 10);
    const char strand = col[6][0];
    if (feat == "exon") {
      std::string id = gff3 ? attr_value(attrs, "Parent", true) : attr_value(attrs, "transcript_id", false);
      if (id.empty()) continue;
      std::vector<std::string> parents;              // GFF3: an exon may list several parents
      for (size_t start = 0; start <= id.size();) {
        size_t c = gff3 ? id.find(',', start) : std::string::npos;
        parents.push_back(id.substr(start, c == std::string::npos ? std::string::npos : c - start));
        if (c == std::string::npos) break;
        start = c + 1;
      }
      bool any_known = false;
      for (auto& one : parents) if (by_id.count(one)) any_known = true;
      for (size_t k = 0; k < parents.size(); k++) {
        if (any_known ? !by_id.count(parents[k]) : k > 0) continue;      // parents not seen yet do not get the exon; with none seen the first is created from it
        Transcript& t = get_tx(parents[k], col[0], strand);
        t.exons.emplace_back(std::min(s, e), std::max(s, e));
      }
    } else if (feat == "gene") {
      const std::string id = gff3 ? attr_value(attrs, "ID", true) : attr_value(attrs, "gene_id", false);
      if (!id.empty()) gene_seen[id] = true;
    } else if (feat == "transcript" || feat == "mRNA" || feat.find("RNA") != std::string::npos || feat == "primary_transcript") {
      std::string id = gff3 ? attr_value(attrs, "ID", true) : attr_value(attrs, "transcript_id", false);
      if (id.empty()) continue;
      Transcript& t = get_tx(id, col[0], strand);
      t.start = std::min(s, e); t.end = std::max(s, e); t.has_line = true;
      const std::string gene = gff3 ? attr_value(attrs, "Parent", true) : attr_value(attrs, "gene_id", false);
      t.level = (!gene.empty() && gene_seen.count(gene)) ? 1 : 0;        // updateParent: level of the parent gene + 1
    }
  }
  free(line); fclose(f);
  for (auto& t : tx) {                                 // exon list: sorted, overlapping or adjacent merged (gff.cpp:988-1079)
    if (t.exons.empty()) { if (!t.has_line) continue; t.exons.emplace_back(t.start, t.end); }   // src/bramble.cpp:576-584 implicit exon
    std::sort(t.exons.begin(), t.exons.end());
    std::vector<std::pair<uint32_t, uint32_t>> m;
    for (auto& x : t.exons) { if (!m.empty() && x.first <= m.back().second + 1) m.back().second = std::max(m.back().second, x.second); else m.push_back(x); }
    t.exons.swap(m);
    t.start = t.exons.front().first; t.end = t.exons.back().second;
    t.tlen = 0; for (auto& x : t.exons) t.tlen += x.second - x.first + 1;
  }
  tx.erase(std::remove_if(tx.begin(), tx.end(), [](const Transcript& t) { return t.exons.empty(); }), tx.end());
  std::sort(tx.begin(), tx.end(), [](const Transcript& a, const Transcript& b) {       // gfo_cmpByLoc
    if (a.refid != b.refid) return strcmp(a.chrom.c_str(), b.chrom.c_str()) < 0;
    if (a.start != b.start) return a.start < b.start;
    if (a.level != b.level) return a.level < b.level;
    if (a.end != b.end) return a.end < b.end;
    return strcmp(a.id.c_str(), b.id.c_str()) < 0;
  });
  A.tx.swap(tx);
}

// FASTA -> base codes per refid (A,C,G,T -> 0..3, everything else 4; GFaSeqGet::copyRange upper-cases, src/g2t.cpp:50-55), packed two per
// byte for the upload (brb_index_desc::genome_format 1: base o in byte o / 2, low nibble = even o; chrom_off counts bases)
inline void load_genome(const std::string& path, const Annotation& A, std::vector<uint64_t>& chrom_off, std::vector<uint8_t>& codes) {
  FILE* f = fopen(path.c_str(), "r");
  if (!f) throw std::runtime_error("could not open genome FASTA " + path);
  std::vector<std::vector<uint8_t>> seqs(A.chrom_names.size());
  uint8_t map[256]; memset(map, 4, sizeof map);
  map['A'] = map['a'] = 0; map['C'] = map['c'] = 1; map['G'] = map['g'] = 2; map['T'] = map['t'] = 3;
  char* line = nullptr; size_t cap = 0; ssize_t n; int cur = -1;
  while ((n = getline(&line, &cap, f)) > 0) {
    while (n > 0 && (line[n - 1] == '\n' || line[n - 1] == '\r')) line[--n] = 0;
    if (line[0] == '>') {
      std::string name(line + 1); size_t sp = name.find_first_of(" \t"); if (sp != std::string::npos) name.resize(sp);
      auto it = A.refid_of.find(name); cur = it == A.refid_of.end() ? -1 : it->second;
    } else if (cur >= 0) {
      auto& v = seqs[cur]; size_t o = v.size(); v.resize(o + n);
      for (ssize_t i = 0; i < n; i++) v[o + i] = map[(uint8_t)line[i]];
    }
  }
  free(line); fclose(f);
  chrom_off.assign(1, 0); codes.clear();
  uint64_t nb = 0;
  for (auto& v : seqs) {
    codes.resize((nb + v.size() + 1) / 2, 0);
    for (size_t i = 0; i < v.size(); i++, nb++) { if (nb & 1) codes[nb >> 1] |= (uint8_t)(v[i] << 4); else codes[nb >> 1] = v[i]; }
    std::vector<uint8_t>().swap(v);
    chrom_off.push_back(nb);
  }
  if (nb & 1) codes[nb >> 1] |= 4 << 4;
}

// build_g2t_tree (src/bramble.cpp:132-211) into flat columns; tid = position in A.tx (== output header order).
struct IndexColumns {
  std::vector<uint64_t> table_off; std::vector<int32_t> st, en; std::vector<uint32_t> tid, pos_start, prev_st, prev_en, next_st, next_en, tlen;
  std::vector<uint8_t> exon_id, hpn; std::vector<uint64_t> chrom_off; std::vector<uint8_t> genome;
  brb_index_desc desc{};
};

inline void build_index_columns(const Annotation& A, IndexColumns& C) {
  struct Row { int32_t st, en; uint32_t tid, pos_start, prev_st, prev_en, next_st, next_en, tlen; uint8_t exon_id, hpn; };
  const size_t n_ref = A.chrom_names.size();
  std::vector<std::vector<Row>> tables(2 * n_ref);
  for (size_t tid = 0; tid < A.tx.size(); tid++) {
    const Transcript& t = A.tx[tid];
    if (t.strand != '+' && t.strand != '-') continue;
    const int nx = (int)t.exons.size();
    std::vector<Row> iv(nx);
    uint32_t pos = 0;
    for (int k = 0; k < nx; k++) {
      const int idx = t.strand == '-' ? nx - k - 1 : k;
      Row& r = iv[k]; r.st = (int32_t)t.exons[idx].first; r.en = (int32_t)t.exons[idx].second + 1; r.exon_id = (uint8_t)idx; r.pos_start = pos; r.tid = (uint32_t)tid;
      pos += (uint32_t)(r.en - r.st);
    }
    for (int k = 0; k < nx; k++) {
      Row r = iv[k];
      r.prev_st = k > 0 ? iv[k - 1].st : 0; r.prev_en = k > 0 ? iv[k - 1].en : 0;
      r.next_st = k < nx - 1 ? iv[k + 1].st : 0; r.next_en = k < nx - 1 ? iv[k + 1].en : 0;
      r.hpn = (k > 0 ? 1 : 0) | (k < nx - 1 ? 2 : 0); r.tlen = pos;
      tables[2 * t.refid + (t.strand == '-')].push_back(r);
    }
  }
  C.table_off.assign(1, 0);
  for (auto& tb : tables) {
    std::vector<int32_t> st(tb.size()); std::vector<uint32_t> perm(tb.size());
    for (size_t i = 0; i < tb.size(); i++) st[i] = tb[i].st;
    brb_sort_by_start(tb.size(), st.data(), perm.data());
    for (size_t i = 0; i < tb.size(); i++) {
      const Row& r = tb[perm[i]];
      C.st.push_back(r.st); C.en.push_back(r.en); C.tid.push_back(r.tid); C.pos_start.push_back(r.pos_start);
      C.prev_st.push_back(r.prev_st); C.prev_en.push_back(r.prev_en); C.next_st.push_back(r.next_st); C.next_en.push_back(r.next_en);
      C.tlen.push_back(r.tlen); C.exon_id.push_back(r.exon_id); C.hpn.push_back(r.hpn);
    }
    C.table_off.push_back(C.st.size());
  }
  brb_index_desc& d = C.desc;
  d.n_tables = (uint32_t)(2 * n_ref); d.table_off = C.table_off.data(); d.st = C.st.data(); d.en = C.en.data(); d.tid = C.tid.data();
  d.exon_id = C.exon_id.data(); d.pos_start = C.pos_start.data(); d.prev_st = C.prev_st.data(); d.prev_en = C.prev_en.data();
  d.next_st = C.next_st.data(); d.next_en = C.next_en.data(); d.has_prev_next = C.hpn.data(); d.transcript_len = C.tlen.data();
  d.n_chrom = C.genome.empty() ? 0 : (uint32_t)n_ref; d.chrom_off = C.genome.empty() ? nullptr : C.chrom_off.data();
  d.genome_codes = C.genome.empty() ? nullptr : C.genome.data(); d.genome_format = 1;
}

}  // namespace brbhost
