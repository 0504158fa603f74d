// bramble_b200 -- drop-in CLI for the reference's projection path (same surface as `bramble`,
// reference src/bramble.cpp:449-485): positional BAM, -G/-S/-o/-p, --fr/--rf/--lr/--lr-hq/--strict,
// the --max-* / --similarity-threshold overrides, --quiet, -V; output = transcriptome-coordinate BAM.
// Host keeps BAM/BGZF decode+encode, GTF/FASTA loading, ingest (process_reads src/bramble.cpp:329-435) and
// the flush()/write_to_bam bookkeeping (src/core.cpp:96-332, incl. the libstdc++ unordered_map order and
// mt19937_64 tie-break); everything between (convert_reads' evaluate/pair/merge) runs behind the C ABI.
// Extra flags: --gpus N (bundles dealt round-robin to N devices, ordered merge), --level L (BGZF level).
#include <chrono>
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <limits>
#include <memory>
#include <mutex>
#include <optional>
#include <random>
#include <string>
#include <string_view>
#include <thread>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "annot.hpp"
#include "bamrec.hpp"
#include "bgzf.hpp"
#include "engine_if.hpp"

using namespace brbhost;

namespace {

const char* VERSION = "0.1.6";     // reference include/bramble.h VERSION: appears in the @PG line

struct Options {
  std::string in_bam, gtf, out_bam, fasta, cmdline;
  bool quiet = false, fr = false, rf = false, lr = false, lr_hq = false, strict = false;
  std::optional<uint32_t> max_clip, max_ins, max_gap, max_err; std::optional<float> sim;
  int threads = 1, gpus = 1, level = 6; uint32_t chunk = 100000;
};

[[noreturn]] void usage(const char* msg) {
  fprintf(stderr, "%s\nusage: bramble_b200 <in.bam> -G <gtf> -o <out.bam> [-S genome.fa] [-p N] [--fr|--rf] [--lr|--lr-hq] [--strict]\n"
                  "       [--max-soft-clip N] [--max-junction-insertion N] [--max-junction-deletion N] [--max-error-exon N]\n"
                  "       [--similarity-threshold F] [--quiet] [--gpus N] [--level L] [-V]\n", msg);
  exit(msg[0] ? 106 : 0);          // CLI11 exits with its own error codes on a parse error
}

Options parse(int argc, char** argv) {
  Options o;
  for (int i = 0; i < argc; i++) { if (i) o.cmdline += " "; o.cmdline += argv[i]; }
  auto need = [&](int& i) -> const char* { if (i + 1 >= argc) usage("missing value"); return argv[++i]; };
  for (int i = 1; i < argc; i++) {
    std::string a = argv[i], v; size_t eq = a.find('=');
    bool has_v = false;
    if (a.rfind("--", 0) == 0 && eq != std::string::npos) { v = a.substr(eq + 1); a = a.substr(0, eq); has_v = true; }
    auto val = [&]() -> std::string { return has_v ? v : std::string(need(i)); };
    if (a == "--quiet") o.quiet = true; else if (a == "--fr") o.fr = true; else if (a == "--rf") o.rf = true;
    else if (a == "--lr") o.lr = true; else if (a == "--lr-hq") o.lr_hq = true; else if (a == "--strict") o.strict = true;
    else if (a == "--max-soft-clip") o.max_clip = (uint32_t)std::stoul(val());
    else if (a == "--max-junction-insertion") o.max_ins = (uint32_t)std::stoul(val());
    else if (a == "--max-junction-deletion") o.max_gap = (uint32_t)std::stoul(val());
    else if (a == "--max-error-exon") o.max_err = (uint32_t)std::stoul(val());
    else if (a == "--similarity-threshold") o.sim = std::stof(val());
    else if (a == "-G" || a == "--guide") o.gtf = val(); else if (a == "-S" || a == "--genome") o.fasta = val();
    else if (a == "-o" || a == "--out") o.out_bam = val(); else if (a == "-p") o.threads = std::stoi(val());
    else if (a == "--gpus") o.gpus = std::stoi(val()); else if (a == "--level") o.level = std::stoi(val());
    else if (a == "--chunk") o.chunk = (uint32_t)std::stoul(val());
    else if (a == "-V" || a == "--version") { printf("version: %s\n", VERSION); exit(0); }
    else if (a == "-h" || a == "--help") usage("");
    else if (a[0] == '-' && a.size() > 1) usage(("unknown option " + a).c_str());
    else if (o.in_bam.empty()) o.in_bam = a; else usage("more than one positional argument");
  }
  if (const char* g = getenv("BRB_GPUS")) o.gpus = atoi(g);      // same as --gpus, without changing the @PG CL: text
  if (const char* t = getenv("BRB_THREADS")) o.threads = atoi(t);   // same as -p, without changing the @PG CL: text
  if (const char* l = getenv("BRB_LEVEL")) o.level = atoi(l);
  if (const char* c = getenv("BRB_CHUNK")) o.chunk = (uint32_t)strtoul(c, nullptr, 10);   // same as --chunk, without changing the @PG CL: text
  if (o.in_bam.empty()) usage("in.bam is required");
  if (o.out_bam.empty()) usage("-o is required");         // required in code, src/bramble.cpp:477-478
  return o;
}

brb_params resolve(const Options& o) {                     // src/evaluate.cpp:1136-1221
  brb_params p{}; const bool lng = o.lr || o.lr_hq;
  uint32_t clip, ins, gap, err; float sim;
  if (!lng) { clip = o.strict ? 0 : 5; ins = 0; gap = 0; sim = 1.0f; err = 0; }
  else if (o.lr) { clip = 40; ins = 40; gap = 40; sim = 0.60f; err = 35; }
  else { clip = 5; ins = 10; gap = 10; sim = 0.90f; err = 35; }
  p.max_clip = o.max_clip.value_or(clip); p.max_junc_ins = o.max_ins.value_or(ins); p.max_junc_gap = o.max_gap.value_or(gap);
  p.max_error_exon = o.max_err.value_or(err); p.similarity_threshold = o.sim.value_or(sim);
  p.long_reads = lng; p.use_fasta = !o.fasta.empty(); p.ignore_small_exons = p.max_error_exon > 0;
  p.filter_by_similarity = p.similarity_threshold < 1.0f;
  return p;
}

std::string field_of(const std::string& line, const char* key) {
  size_t p = 0; const size_t kl = strlen(key);
  while ((p = line.find('\t', p)) != std::string::npos) {
    p++;
    if (line.compare(p, kl, key) == 0) { size_t e = line.find('\t', p); return line.substr(p + kl, e == std::string::npos ? std::string::npos : e - p - kl); }
  }
  return std::string();
}

// htslib sam_hdr_add_pg(hdr, "bramble", "VN", v, "CL", cl, NULL): unique ID, PP chained to every chain end.
std::string add_pg(std::string text, const std::string& name, const std::string& extra) {
  std::vector<std::string> ids, pps;
  for (size_t p = 0; p < text.size();) {
    size_t e = text.find('\n', p); if (e == std::string::npos) e = text.size();
    const std::string line = text.substr(p, e - p);
    if (line.compare(0, 3, "@PG") == 0) { ids.push_back(field_of(line, "ID:")); pps.push_back(field_of(line, "PP:")); }
    p = e + 1;
  }
  std::vector<std::string> ends;
  for (size_t i = 0; i < ids.size(); i++) {
    bool referenced = false;
    for (size_t j = 0; j < pps.size(); j++) if (j != i && pps[j] == ids[i]) referenced = true;
    if (!referenced) ends.push_back(ids[i]);
  }
  auto unique_id = [&]() { std::string id = name; int n = 0; for (;;) { bool taken = false; for (auto& x : ids) if (x == id) taken = true; if (!taken) break; id = name + "." + std::to_string(++n); } ids.push_back(id); return id; };
  if (!text.empty() && text.back() != '\n') text += "\n";
  if (ends.empty()) text += "@PG\tID:" + unique_id() + "\tPN:" + name + extra + "\n";
  else for (auto& pp : ends) text += "@PG\tID:" + unique_id() + "\tPN:" + name + "\tPP:" + pp + extra + "\n";
  return text;
}

struct InHeader { std::string text; std::vector<std::string> ref_names; };

InHeader read_bam_header(BgzfReader& in) {
  InHeader h; char magic[4]; int32_t l_text = 0, n_ref = 0;
  if (in.read(magic, 4) != 4 || memcmp(magic, "BAM\1", 4)) throw std::runtime_error("input is not a BAM file");
  in.read(&l_text, 4); h.text.resize(l_text); if (l_text) in.read(&h.text[0], l_text);
  while (!h.text.empty() && h.text.back() == 0) h.text.pop_back();
  in.read(&n_ref, 4);
  for (int i = 0; i < n_ref; i++) { int32_t ln = 0, len = 0; in.read(&ln, 4); std::string nm(ln, 0); in.read(&nm[0], ln); nm.resize(strlen(nm.c_str())); in.read(&len, 4); h.ref_names.push_back(nm); }
  return h;
}

// output header text, reference src/bramble.cpp:552-623
std::string make_header(const InHeader& ih, const Annotation& A, const Options& o) {
  const std::string with_pg = add_pg(ih.text, "bramble", std::string("\tVN:") + VERSION + "\tCL:" + o.cmdline);
  std::string out;
  auto each = [&](auto fn) { for (size_t p = 0; p < with_pg.size();) { size_t e = with_pg.find('\n', p); if (e == std::string::npos) e = with_pg.size(); fn(with_pg.substr(p, e - p)); p = e + 1; } };
  each([&](const std::string& l) { if (l.compare(0, 3, "@HD") == 0) out += l + "\n"; });
  for (auto& t : A.tx) if (t.tlen > 0) out += "@SQ\tSN:" + t.id + "\tLN:" + std::to_string((int)t.tlen) + "\n";
  each([&](const std::string& l) { if (l.compare(0, 3, "@SQ") != 0 && l.compare(0, 3, "@HD") != 0) out += l + "\n"; });
  out += "@CO\tGenerated from GTF: " + o.gtf + "\n";
  return out;
}

// ------------------------------------------------------------------------------------------------ bundles
struct Bundle {
  uint64_t index = 0;
  std::vector<uint8_t> raw; std::vector<uint64_t> rec_off; std::vector<uint32_t> rec_len;       // record bodies
  std::vector<uint32_t> group_off, exon_off, exon_st, exon_en, cigar_off, cigar; std::vector<int32_t> refid, mate; std::vector<int8_t> strand;
  std::vector<uint64_t> seq_off; std::vector<uint8_t> seq;
  std::vector<uint64_t> name_off; std::vector<uint8_t> names;                // read name per group (the engine's on-device flush bookkeeping hashes them)
  std::vector<int32_t> hi, nh; std::vector<uint8_t> mapq, prim; bool have_tags = false;   // BRB_OPT_FLUSH_TAGS results, when the engine provides them
  // result (copied out of the engine's staging)
  std::vector<uint32_t> a_roff, a_rcnt, src, tid, pos, coff, cig; std::vector<int8_t> tstrand; std::vector<uint8_t> flags; std::vector<double> sim; std::vector<int32_t> clip;
  // assembled output records of the bundle (uncompressed BAM bytes) and its share of the final report
  // flush chunks of the bundle (reference CHUNK_SIZE = 5000 pairs, src/core.cpp:27,421-423): group ranges, and the output records
  // of each chunk (uncompressed BAM bytes), assembled independently of one another
  std::vector<std::pair<uint32_t, uint32_t>> chunks; std::vector<size_t> chunk_buckets; std::vector<std::vector<uint8_t>> parts; uint32_t parts_left = 0;
  // device record assembly (brb_emit_bgzf): the engine slot the rows are resident in, and per chunk the emitted units in output order
  // (4 words each: first row, HI, NH, MAPQ | primary << 8) instead of the record bytes
  int slot = -1; std::vector<std::vector<uint32_t>> unit_parts; bool emit_claimed = false, emitted = false; std::vector<uint8_t> zout;
  uint64_t dropped = 0, total_complete = 0, total_unique = 0; bool assembled = false;
  RecView view(uint32_t a) const { return RecView{raw.data() + rec_off[a], rec_len[a]}; }
  uint32_t n_aln() const { return (uint32_t)refid.size(); }
};

struct Stats { uint64_t total_reads = 0, unmapped = 0, dropped = 0, total_complete = 0, total_unique = 0; };

void fill_batch(const Bundle& b, bool use_fasta, brb_batch_in& in) {
  in.n_aln = b.n_aln(); in.n_groups = (uint32_t)b.group_off.size() - 1; in.group_off = b.group_off.data(); in.refid = b.refid.data();
  in.strand = b.strand.data(); in.exon_off = b.exon_off.data(); in.exon_st = b.exon_st.data(); in.exon_en = b.exon_en.data();
  in.cigar_off = b.cigar_off.data(); in.cigar = b.cigar.data(); in.mate_idx = b.mate.data();
  in.group_seq_off = use_fasta ? b.seq_off.data() : nullptr; in.seq_codes = use_fasta ? (b.seq.empty() ? (const uint8_t*)"" : b.seq.data()) : nullptr;
  in.group_name_off = b.name_off.data(); in.name_bytes = b.names.empty() ? (const uint8_t*)"" : b.names.data();
}

void store_result(Bundle& b, const brb_batch_out& o) {
  const uint32_t R = o.n_rec, G = (uint32_t)b.group_off.size() - 1;
  (void)G; b.a_roff.assign(o.aln_rec_off, o.aln_rec_off + b.n_aln()); b.a_rcnt.assign(o.aln_rec_cnt, o.aln_rec_cnt + b.n_aln()); b.src.assign(o.src_aln, o.src_aln + R); b.tid.assign(o.tid, o.tid + R);
  b.pos.assign(o.pos, o.pos + R); b.tstrand.assign(o.tx_strand, o.tx_strand + R); b.flags.assign(o.flags, o.flags + R);
  if (o.sim_score) b.sim.assign(o.sim_score, o.sim_score + R); else b.sim.assign(R, 0.0);      // NULL: every score is 0 (header)
  if (o.clip_score) b.clip.assign(o.clip_score, o.clip_score + R); else b.clip.assign(R, 0);
  b.coff.assign(o.cigar_off, o.cigar_off + R + 1); b.cig.assign(o.cigar, o.cigar + o.n_cigar_words);
  b.have_tags = o.hi && o.nh && o.mapq && o.primary;
  if (b.have_tags) { b.hi.assign(o.hi, o.hi + R); b.nh.assign(o.nh, o.nh + R); b.mapq.assign(o.mapq, o.mapq + R); b.prim.assign(o.primary, o.primary + R); }
}

uint32_t get_mapq(uint32_t nh, bool lng) {                 // src/core.cpp:46-58
  if (!lng) return nh == 1 ? 255 : nh == 2 ? 3 : (nh == 3 || nh == 4) ? 1 : 0;
  return nh > 1 ? 0 : 3;
}
int32_t get_rand(uint32_t x, uint64_t seed_key) {          // src/core.cpp:214-218 (libstdc++ mt19937_64 + uniform_int_distribution)
  std::mt19937_64 gen(seed_key); std::uniform_int_distribution<uint32_t> dis(0, x - 1); return dis(gen);
}

// The flush chunks of a bundle: the reference appends a group's pairs and flushes once 5000 or more have accumulated, checking only at
// the end of a group (src/core.cpp:384-423); the last chunk is whatever is left.
// The records of a chunk leave in the iteration order of the reference's name -> pairs map, ONE map per bundle that is clear()ed
// after each flush (src/core.cpp:234,330): clear() keeps the bucket array, so the order inside chunk k depends on how far the
// earlier chunks grew it.  A stand-in map with the same key sequence is run here to record the bucket count each chunk starts
// from; the chunks are then assembled independently, each in a map rehash()ed to that count.
void plan_chunks(Bundle& b) {
  const uint32_t G = (uint32_t)b.group_off.size() - 1;
  uint32_t n_pairs = 0, g0 = 0;
  b.chunks.clear(); b.chunk_buckets.clear();
  std::unordered_map<std::string_view, char> names;
  size_t buckets = names.bucket_count();
  for (uint32_t g = 0; g < G; g++) {
    bool any = false;
    for (uint32_t a = b.group_off[g]; a < b.group_off[g + 1]; a++)
      for (uint32_t r = b.a_roff[a], hi = r + b.a_rcnt[a]; r < hi;) {
        if (!any) names.try_emplace(std::string_view(b.view(b.src[r]).name()), 0);       // one name per group
        any = true; n_pairs++; r += (b.flags[r] & BRB_REC_PAIRED) ? 2 : 1;
      }
    if (n_pairs >= 5000) { b.chunks.emplace_back(g0, g + 1); b.chunk_buckets.push_back(buckets); g0 = g + 1; n_pairs = 0; names.clear(); buckets = names.bucket_count(); }
  }
  if (g0 < G) { b.chunks.emplace_back(g0, G); b.chunk_buckets.push_back(buckets); }
}

// flush() of the reference for the groups [g0, g1) of a bundle (one flush chunk): emit(unit) is called once per BamInfo, in the
// order write_to_bam sees them.
struct Unit { uint32_t row; bool paired; uint32_t group; int32_t hi_r = 0, hi_m = 0; bool primary = false; uint32_t nh = 0, mapq = 0; };
template <typename Emit>
void replay_chunk(const Bundle& b, uint32_t g0, uint32_t g1, size_t start_buckets, bool lng, Stats& st, Emit&& emit) {
  std::unordered_map<std::string, std::vector<Unit>> pairs_by_name;
  if (start_buckets != pairs_by_name.bucket_count()) pairs_by_name.rehash(start_buckets);    // the bucket array the earlier chunks left behind
  for (uint32_t g = g0; g < g1; g++) {
    bool any = false;
    for (uint32_t a = b.group_off[g]; a < b.group_off[g + 1]; a++) {       // unit blocks of the group, in alignment order = emit order
      const uint32_t lo = b.a_roff[a], hi = lo + b.a_rcnt[a];
      for (uint32_t r = lo; r < hi;) {
        const bool paired = b.flags[r] & BRB_REC_PAIRED;
        const RecView v = b.view(b.src[r]);
        pairs_by_name[std::string(v.name())].push_back(Unit{r, paired, g});
        r += paired ? 2 : 1; any = true;
      }
    }
    if (!any) st.dropped++;
  }
  std::vector<Unit*> filtered;
  for (auto& [read_name, pairs] : pairs_by_name) {
    // HI / NH / MAPQ / primary: from the device when the engine computed them (brb_batch_out::hi ...) -- unless the name came in as
    // more than one group inside this chunk (input not name-collated) or the device left the tie to us (primary == 2)
    bool from_device = b.have_tags;
    for (size_t i = 0; from_device && i < pairs.size(); i++) if (pairs[i].group != pairs[0].group || b.prim[pairs[i].row] == 2) from_device = false;
    if (from_device) {
      const int total = b.nh[pairs[0].row];
      for (auto& u : pairs) { u.hi_r = b.hi[u.row]; if (u.paired) u.hi_m = b.hi[u.row + 1]; u.primary = b.prim[u.row] == 1; u.nh = (uint32_t)total; u.mapq = b.mapq[u.row]; filtered.push_back(&u); }
      st.total_complete += total; if (total == 1) st.total_unique++;
      continue;
    }
    double best_score = -std::numeric_limits<double>::infinity(); int count_at_best = 0; size_t best = pairs.size(); int32_t hit_index = 1; int total = 0;
    auto score_of = [&](const Unit& u) { double s = b.sim[u.row]; if (u.paired) s = std::max(s, b.sim[u.row + 1]); return s; };
    for (size_t i = 0; i < pairs.size(); i++) {
      Unit& u = pairs[i]; u.hi_r = hit_index++; total++;
      if (u.paired) { u.hi_m = hit_index++; total++; }
      const double ps = score_of(u);
      if (ps > best_score) { best_score = ps; best = i; count_at_best = 1; } else if (ps == best_score) count_at_best++;
    }
    if (best != pairs.size()) {
      if (count_at_best == 1) pairs[best].primary = true;
      else {
        std::vector<size_t> tied;
        for (size_t i = 0; i < pairs.size(); i++) if (score_of(pairs[i]) == best_score) tied.push_back(i);
        const uint64_t seed_key = std::hash<std::string>{}(read_name);
        pairs[tied[get_rand((uint32_t)tied.size(), seed_key)]].primary = true;
      }
    }
    st.total_complete += total; if (total == 1) st.total_unique++;
    for (auto& u : pairs) { u.nh = (uint32_t)total; u.mapq = get_mapq((uint32_t)total, lng); filtered.push_back(&u); }
  }
  for (Unit* u : filtered) emit(*u);
}

// ... + write_to_bam on the host: the record bytes of the chunk
void assemble_chunk(const Bundle& b, uint32_t g0, uint32_t g1, size_t start_buckets, bool lng, Stats& st, std::vector<uint8_t>& out) {
  std::vector<uint8_t> aux;
  replay_chunk(b, g0, g1, start_buckets, lng, st, [&](const Unit& un) {
    const Unit* u = &un;
    for (int k = 0; k < (u->paired ? 2 : 1); k++) {
      const uint32_t r = u->row + k, other = u->row + (1 - k);
      OutInfo o{}; o.tid = b.tid[r]; o.pos = b.pos[r]; o.strand = (char)b.tstrand[r]; o.paired = u->paired; o.same_tx = b.flags[r] & BRB_REC_SAMETX;
      o.primary = u->primary; o.long_reads = lng; o.mapq = u->mapq; o.nh = (int32_t)u->nh; o.hi = k == 0 ? u->hi_r : u->hi_m;
      o.sim = b.sim[r]; o.clip = b.clip[r]; o.cigar = b.cig.data() + b.coff[r]; o.n_cigar = b.coff[r + 1] - b.coff[r];
      if (u->paired) { o.mate_tid = b.tid[other]; o.mate_pos = b.pos[other]; }
      assemble_record(b.view(b.src[r]), o, aux, out);
    }
  });
}

// ... or only the ORDER and the tag values, for the engine to build the bytes from its resident rows (brb_emit_in::units)
void units_of_chunk(const Bundle& b, uint32_t g0, uint32_t g1, size_t start_buckets, bool lng, Stats& st, std::vector<uint32_t>& units) {
  replay_chunk(b, g0, g1, start_buckets, lng, st, [&](const Unit& u) {
    if (u.paired && u.hi_m != u.hi_r + 1) throw std::runtime_error("internal: HI of a mate record is not HI + 1");
    units.push_back(u.row); units.push_back((uint32_t)u.hi_r); units.push_back(u.nh); units.push_back((u.mapq & 0xffu) | (u.primary ? 0x100u : 0u));
  });
}

}  // namespace

int main(int argc, char** argv) {
  Options o = parse(argc, argv);
  try {
    const bool lng = o.lr || o.lr_hq; const brb_params params = resolve(o);
    if (!o.quiet) printf("\n[bramble_b200] starting (reference surface version %s)\n", VERSION);
    Annotation A;
    try { load_annotation(o.gtf, A); } catch (const std::exception& e) { fprintf(stderr, "error: %s\n", e.what()); return 0; }   // reference returns 0, src/bramble.cpp:520-525
    if (A.tx.empty()) { fprintf(stderr, "error: could not find valid reference transcripts in %s\n", o.gtf.c_str()); return 0; }
    IndexColumns cols;
    if (!o.fasta.empty()) load_genome(o.fasta, A, cols.chrom_off, cols.genome);
    build_index_columns(A, cols);
    const int n_gpu = std::max(1, o.gpus);
    // BRB_GPU_MAP=d0,d1,...: CUDA device of dealer slot k (default: slot k -> device k).  "0,0" runs the whole --gpus 2 path
    // (dealer, one worker and one index copy per slot, ordered merge) on a single device.
    std::vector<int> gpu_map(n_gpu); for (int g = 0; g < n_gpu; g++) gpu_map[g] = g;
    if (const char* gm = getenv("BRB_GPU_MAP")) { int k = 0; for (const char* q = gm; *q && k < n_gpu; k++) { gpu_map[k] = atoi(q); q = strchr(q, ','); if (!q) break; q++; } }
    std::vector<std::unique_ptr<EngineIf>> engines;
    for (int g = 0; g < n_gpu; g++) { engines.emplace_back(make_engine()); engines.back()->open(gpu_map[g], cols.desc); }

    // the input: inflated by a host pool, or -- BRB_DEVICE_INFLATE=1 -- in pieces by engine 0 (brb_bgzf_inflate).  The device decoder does a
    // whole file at ~27 GB/s but a call takes 5 ms however small (one warp per 64 KiB block), so for inputs of a few bundles the host pool
    // is as fast (measured: producer 0.13 s vs 0.09 s on a 55 MB BAM); it pays on long inputs and on hosts with few cores.
    // (the reader is declared after the engines: its threads are gone before the engines are)
    const int n_inflate = getenv("BRB_INFLATE_THREADS") ? std::max(1, atoi(getenv("BRB_INFLATE_THREADS"))) : std::min(o.threads, 8);
    BgzfReader::PieceInflater dev_inflate = nullptr;
    if (getenv("BRB_DEVICE_INFLATE")) {
      std::vector<uint8_t> probe;
      if (engines[0]->bgzf_inflate(0, nullptr, 0, probe)) {
        EngineIf* e0 = engines[0].get();
        dev_inflate = [e0](int w, const uint8_t* z, size_t n, std::vector<uint8_t>& out) { return e0->bgzf_inflate(w, z, n, out); };
      }
    }
    if (!dev_inflate && getenv("BRB_TEST_PIECES")) dev_inflate = [](int, const uint8_t*, size_t, std::vector<uint8_t>&) { return false; };   // tests: the piece reader with every piece declined (inflated by the host, block by block)
    BgzfReader in(o.in_bam, n_inflate, dev_inflate);
    const InHeader ih = read_bam_header(in);
    const std::string header = make_header(ih, A, o);
    std::vector<int> ref2refid(ih.ref_names.size(), -1);
    for (size_t i = 0; i < ih.ref_names.size(); i++) { auto it = A.refid_of.find(ih.ref_names[i]); if (it != A.refid_of.end()) ref2refid[i] = it->second; }

    BgzfWriter out(o.out_bam, o.level, std::max(1, o.threads));
    { std::vector<uint8_t> hb; auto put = [&](const void* p, size_t n) { hb.insert(hb.end(), (const uint8_t*)p, (const uint8_t*)p + n); };
      put("BAM\1", 4); int32_t lt = (int32_t)header.size(); put(&lt, 4); put(header.data(), header.size());
      int32_t nt = 0; for (auto& t : A.tx) if (t.tlen > 0) nt++; put(&nt, 4);
      for (auto& t : A.tx) if (t.tlen > 0) { int32_t l = (int32_t)t.id.size() + 1; put(&l, 4); put(t.id.c_str(), l); uint32_t ln = t.tlen; put(&ln, 4); }
      out.write(hb.data(), hb.size()); out.flush_block(); }

    // pipeline: producer (this thread: record walk + bundle cut) -> column extraction (pool) -> per-GPU worker -> flush replay +
    // record assembly (pool) -> ordered writer (BGZF blocks compressed on the device, or by the host pool)
    const bool timing = getenv("BRB_CLI_TIMING") != nullptr;      // stage seconds on stderr at the end
    using clk = std::chrono::steady_clock;
    auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
    double t_read = 0, t_ingest = 0, t_wait_push = 0, t_project = 0, t_store = 0, t_assemble = 0, t_write = 0, t_extract = 0, t_emit = 0;
    const auto t_begin = clk::now();
    Stats st; std::mutex m; std::condition_variable cv;
    std::deque<std::shared_ptr<Bundle>> to_ingest;            // cut by the producer, waiting for their columns
    std::vector<std::deque<std::shared_ptr<Bundle>>> todo(n_gpu); std::deque<std::shared_ptr<Bundle>> inflight; bool eof = false; std::string err;
    std::deque<std::shared_ptr<Bundle>> to_assemble;          // projected, waiting for flush replay + record assembly
    uint64_t n_cut = 0, n_ingested = 0;
    int workers_left = n_gpu;
    const int n_asm = std::max(1, o.threads);
    const bool dev_deflate = getenv("BRB_HOST_DEFLATE") == nullptr;      // BGZF blocks of the output are compressed by the engine when it can
    // ... and the records themselves are assembled there too (brb_emit_bgzf): the host only replays flush() for the order and the tags
    const bool dev_emit = dev_deflate && engines[0]->can_emit() && getenv("BRB_HOST_ASSEMBLY") == nullptr;
    std::vector<std::vector<int>> free_slots(n_gpu);
    for (int g = 0; g < n_gpu; g++) for (int s2 = engines[g]->slots(); s2-- > 0;) free_slots[g].push_back(s2);
    int planning = 0;                                         // bundles whose flush chunks are being cut (their chunk tasks are about to appear)
    // ingest of one bundle (reference process_reads, src/bramble.cpp:329-435, per record): read exons, CIGAR words, strand,
    // annotated chromosome, mate links (process_pairs :272-311), group names and -- with -S -- the group-shared sequence
    auto ingest_bundle = [&](Bundle& b) -> std::string {
      const uint32_t n = (uint32_t)b.rec_off.size();
      std::unordered_map<std::string, uint32_t> hashread;
      b.exon_off.assign(1, 0); b.cigar_off.assign(1, 0); b.seq_off.assign(1, 0); b.name_off.assign(1, 0);
      b.refid.reserve(n); b.mate.assign(n, -1); b.strand.reserve(n);
      for (uint32_t a = 0; a < n; a++) {
        const RecView w = b.view(a);
        if (!read_exons(w, b.exon_st, b.exon_en)) return std::string("Error: invalid CIGAR record for ") + w.name() + " !";
        b.exon_off.push_back((uint32_t)b.exon_st.size());
        for (uint32_t i = 0; i < w.n_cigar(); i++) b.cigar.push_back(rd32(w.cigar() + 4 * i));
        b.cigar_off.push_back((uint32_t)b.cigar.size());
        b.strand.push_back(lng ? '.' : get_strand(w, o.fr, o.rf));
        const int32_t t = w.tid(); b.refid.push_back(t >= 0 && (size_t)t < ref2refid.size() ? ref2refid[t] : -1);
        if ((w.flag() & 1) && w.tid() == w.mtid()) {          // process_pairs, src/bramble.cpp:272-311
          const int32_t start = w.pos() + 1, mate_start = w.mpos() < 0 ? 0 : w.mpos() + 1;
          const std::string name = w.name();
          auto it = hashread.find(name + '-' + std::to_string(mate_start));
          if (it != hashread.end()) { const uint32_t mt = it->second; if (b.mate[a] < 0) b.mate[a] = (int32_t)mt; if (b.mate[mt] < 0) b.mate[mt] = (int32_t)a; hashread.erase(it); }
          else hashread[name + '-' + std::to_string(start)] = a;
        }
      }
      static const char nt16[] = "=ACMGRSVTWYHKDBN";
      for (size_t g = 0; g + 1 < b.group_off.size(); g++) {
        if (params.use_fasta) {                              // group-shared sequence = first record of the group with l_qseq > 0 (src/core.cpp:353-378)
          for (uint32_t a = b.group_off[g]; a < b.group_off[g + 1]; a++) {
            const RecView v = b.view(a); const int32_t lq = v.l_seq();
            if (lq > 0) { const uint8_t* sq = v.seq(); for (int32_t i = 0; i < lq; i++) { const char c = nt16[(sq[i >> 1] >> ((~i & 1) << 2)) & 0xf]; b.seq.push_back(c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : 4); } break; }
          }
          b.seq_off.push_back(b.seq.size());
        }
        const RecView v = b.view(b.group_off[g]); const char* nm = v.name(); b.names.insert(b.names.end(), nm, nm + strlen(nm)); b.name_off.push_back(b.names.size());
      }
      return std::string();
    };
    auto worker = [&](int g) {
      for (;;) {
        std::shared_ptr<Bundle> b;
        { std::unique_lock<std::mutex> lk(m); cv.wait(lk, [&] { return (eof && n_ingested == n_cut) || !todo[g].empty() || !err.empty(); });
          if (!err.empty() || (todo[g].empty() && eof && n_ingested == n_cut)) { workers_left--; cv.notify_all(); return; }
          if (todo[g].empty()) continue;
          b = todo[g].front(); todo[g].pop_front(); }
        // the rows stay resident in one of the engine's slots until the writer has had the records built from them
        { std::unique_lock<std::mutex> lk(m); cv.wait(lk, [&] { return !free_slots[g].empty() || !err.empty(); });
          if (!err.empty()) { workers_left--; cv.notify_all(); return; }
          b->slot = free_slots[g].back(); free_slots[g].pop_back(); }
        brb_batch_in bi{}; fill_batch(*b, params.use_fasta, bi); brb_batch_out bo{};
        const auto t0 = clk::now();
        const int rc = engines[g]->project_slot(b->slot, params, bi, bo);
        const auto t1 = clk::now();
        if (rc == 0) store_result(*b, bo);
        const auto t2 = clk::now();
        std::unique_lock<std::mutex> lk(m);
        t_project += secs(t0, t1); t_store += secs(t1, t2);
        if (rc != 0) { err = engines[g]->last_error(); workers_left--; cv.notify_all(); return; }
        if (!dev_emit) { free_slots[g].push_back(b->slot); b->slot = -1; }
        to_assemble.push_back(b); cv.notify_all();
      }
    };
    // pool threads: column extraction of freshly cut bundles, and flush replay + record assembly of projected ones -- one task per
    // flush chunk (a read name never spans two chunks, the chunks of a bundle are independent); the writer below keeps the order
    std::deque<std::pair<std::shared_ptr<Bundle>, uint32_t>> chunk_tasks;
    auto assembler = [&]() {
      for (;;) {
        std::shared_ptr<Bundle> b; int what = 0; uint32_t k = 0;       // 1 ingest, 2 plan the chunks of a projected bundle, 3 one chunk
        { std::unique_lock<std::mutex> lk(m); cv.wait(lk, [&] { return !err.empty() || !chunk_tasks.empty() || !to_assemble.empty() || !to_ingest.empty() || (workers_left == 0 && planning == 0); });
          if (!err.empty()) return;
          if (!chunk_tasks.empty()) { b = chunk_tasks.front().first; k = chunk_tasks.front().second; chunk_tasks.pop_front(); what = 3; }
          else if (!to_assemble.empty()) { b = to_assemble.front(); to_assemble.pop_front(); what = 2; planning++; }
          else if (!to_ingest.empty()) { b = to_ingest.front(); to_ingest.pop_front(); what = 1; }
          else { if (workers_left == 0 && planning == 0) return; continue; } }
        const auto t0 = clk::now();
        if (what == 1) {
          const std::string e = ingest_bundle(*b);
          std::unique_lock<std::mutex> lk(m);
          t_extract += secs(t0, clk::now());
          if (!e.empty()) { if (err.empty()) err = e; cv.notify_all(); return; }
          todo[b->index % n_gpu].push_back(b); n_ingested++; cv.notify_all();
        } else if (what == 2) {
          plan_chunks(*b);
          if (dev_emit) b->unit_parts.assign(b->chunks.size(), std::vector<uint32_t>()); else b->parts.assign(b->chunks.size(), std::vector<uint8_t>());
          std::unique_lock<std::mutex> lk(m);
          t_assemble += secs(t0, clk::now());
          planning--;
          b->parts_left = (uint32_t)b->chunks.size();
          if (b->chunks.empty()) b->assembled = true;
          for (uint32_t i = 0; i < b->chunks.size(); i++) chunk_tasks.emplace_back(b, i);
          cv.notify_all();
        } else {
          Stats bs; std::string e;
          try {
            if (dev_emit) units_of_chunk(*b, b->chunks[k].first, b->chunks[k].second, b->chunk_buckets[k], lng, bs, b->unit_parts[k]);
            else assemble_chunk(*b, b->chunks[k].first, b->chunks[k].second, b->chunk_buckets[k], lng, bs, b->parts[k]);
          } catch (const std::exception& ex) { e = ex.what(); }
          std::unique_lock<std::mutex> lk(m);
          t_assemble += secs(t0, clk::now());
          if (!e.empty()) { if (err.empty()) err = e; cv.notify_all(); return; }
          b->dropped += bs.dropped; b->total_complete += bs.total_complete; b->total_unique += bs.total_unique;
          if (--b->parts_left == 0) { b->assembled = true; if (!dev_emit) { b->raw.clear(); b->raw.shrink_to_fit(); } }
          cv.notify_all();
        }
      }
    };
    // device record path: one emitter thread per engine turns its oldest assembled bundle into BGZF blocks (brb_emit_bgzf on the rows the
    // bundle left in its slot); the writer only appends the blocks in bundle order
    auto emitter = [&](int g) {                                   // the bundles of engine g, oldest first
      for (;;) {
        std::shared_ptr<Bundle> b;
        auto next = [&]() -> std::shared_ptr<Bundle> {
          for (auto& x : inflight) if ((int)(x->index % n_gpu) == g && !x->emit_claimed) return x->assembled ? x : nullptr;
          return nullptr; };
        { std::unique_lock<std::mutex> lk(m);
          cv.wait(lk, [&] { return !err.empty() || (eof && inflight.empty()) || next() != nullptr; });
          if (!err.empty() || (eof && inflight.empty())) return;
          b = next();
          if (!b) continue;
          b->emit_claimed = true; }
        const auto t1 = clk::now();
        try {
          std::vector<uint32_t> units;
          for (auto& part : b->unit_parts) { units.insert(units.end(), part.begin(), part.end()); std::vector<uint32_t>().swap(part); }
          if (!units.empty()) {
            brb_emit_in ei{}; ei.raw = b->raw.data(); ei.raw_bytes = b->raw.size(); ei.rec_off = b->rec_off.data(); ei.rec_len = b->rec_len.data();
            ei.n_units = (uint32_t)(units.size() / 4); ei.units = units.data();
            const uint8_t* z = nullptr; size_t zn = 0;
            engines[g]->emit(b->slot, params, ei, z, zn);
            b->zout.assign(z, z + zn);
          }
        } catch (const std::exception& e) { std::unique_lock<std::mutex> lk(m); if (err.empty()) err = e.what(); cv.notify_all(); return; }
        std::unique_lock<std::mutex> lk(m);
        t_emit += secs(t1, clk::now());
        free_slots[g].push_back(b->slot); b->slot = -1; b->emitted = true; b->raw.clear(); b->raw.shrink_to_fit();
        cv.notify_all();
      }
    };
    auto writer = [&]() {
      std::vector<uint8_t> joined;
      for (;;) {
        std::shared_ptr<Bundle> b;
        { std::unique_lock<std::mutex> lk(m); cv.wait(lk, [&] { return !err.empty() || (!inflight.empty() && (dev_emit ? inflight.front()->emitted : inflight.front()->assembled)) || (eof && inflight.empty()); });
          if (!err.empty()) return; if (inflight.empty()) return; b = inflight.front(); inflight.pop_front(); cv.notify_all(); }
        const auto t1 = clk::now();
        try {
          if (dev_emit) { if (!b->zout.empty()) out.write_raw(b->zout.data(), b->zout.size()); }
          else if (dev_deflate) {                              // a few chunks per call: one upload, one kernel sequence, one download
            size_t k = 0;
            while (k < b->parts.size()) {
              joined.clear();
              while (k < b->parts.size() && (joined.empty() || joined.size() + b->parts[k].size() <= ((size_t)48 << 20))) { joined.insert(joined.end(), b->parts[k].begin(), b->parts[k].end()); std::vector<uint8_t>().swap(b->parts[k]); k++; }
              const uint8_t* z = nullptr; size_t zn = 0;
              if (!joined.empty() && engines[b->index % n_gpu]->bgzf_compress(joined.data(), joined.size(), z, zn)) out.write_raw(z, zn);
              else out.write(joined.data(), joined.size());
            }
          } else for (auto& part : b->parts) { out.write(part.data(), part.size()); std::vector<uint8_t>().swap(part); }
        } catch (const std::exception& e) { std::unique_lock<std::mutex> lk(m); if (err.empty()) err = e.what(); cv.notify_all(); return; }
        st.dropped += b->dropped; st.total_complete += b->total_complete; st.total_unique += b->total_unique;
        t_write += secs(t1, clk::now());
      }
    };
    std::vector<std::thread> threads; for (int g = 0; g < n_gpu; g++) threads.emplace_back(worker, g);
    for (int i = 0; i < n_asm; i++) threads.emplace_back(assembler);
    if (dev_emit) for (int g = 0; g < n_gpu; g++) threads.emplace_back(emitter, g);
    std::thread wthread(writer);

    auto bundle = std::make_shared<Bundle>(); uint64_t n_bundles = 0;
    std::string prev_name;
    auto new_bundle = [&]() { auto b = std::make_shared<Bundle>(); b->group_off.push_back(0); return b; };
    bundle = new_bundle();
    auto push_bundle = [&]() {
      if (bundle->rec_off.empty()) return;
      Bundle& b = *bundle; b.group_off.push_back((uint32_t)b.rec_off.size());
      b.index = n_bundles++;
      const auto tw = clk::now();
      { std::unique_lock<std::mutex> lk(m); cv.wait(lk, [&] { return inflight.size() < (size_t)(2 * n_gpu + 2 + n_asm) || !err.empty(); }); to_ingest.push_back(bundle); inflight.push_back(bundle); n_cut++; }
      cv.notify_all();
      t_wait_push += secs(tw, clk::now());
      bundle = new_bundle();
    };
    std::string producer_err;
    try {
    // The inflated stream is appended to the current bundle's buffer block by block and the records are walked in place
    // (rec_off points at the record body, behind its block_size word); when a record opens the next bundle, the unparsed tail
    // of the buffer -- less than a BGZF block and a record -- moves over.
    size_t scan = 0, last_raw = (size_t)o.chunk * 400;
    bundle->raw.reserve(last_raw); bundle->rec_off.reserve(o.chunk + 64); bundle->rec_len.reserve(o.chunk + 64);
    for (;;) {
      { std::unique_lock<std::mutex> lk(m); if (!err.empty()) break; }     // a worker failed: stop reading, nothing would consume the bundles
      const uint8_t* blk = nullptr; size_t blk_n = 0;
      const auto tr0 = timing ? clk::now() : clk::time_point();
      if (!in.next_chunk(blk, blk_n)) break;
      bundle->raw.insert(bundle->raw.end(), blk, blk + blk_n);
      if (timing) t_read += secs(tr0, clk::now());
      for (;;) {
        std::vector<uint8_t>& raw = bundle->raw;
        if (scan + 4 > raw.size()) break;
        const int32_t bs = rdi32(raw.data() + scan);
        if (bs < 32) throw std::runtime_error("malformed BAM record (block_size < 32)");
        if (scan + 4 + (size_t)bs > raw.size()) break;
        const RecView v{raw.data() + scan + 4, (uint32_t)bs};
        if (v.flag() & 4) { st.total_reads++; st.unmapped++; scan += 4 + (size_t)bs; continue; }
        const uint32_t id = (uint32_t)bundle->rec_off.size();
        const bool new_name = (id == 0) || prev_name != v.name();
        if (id >= o.chunk && new_name) {                       // cut here: this record is parsed again as the first one of the next bundle
          std::vector<uint8_t> tail(raw.begin() + scan, raw.end());
          raw.resize(scan); last_raw = raw.size();
          push_bundle();
          bundle->raw.reserve(last_raw + last_raw / 8 + (1 << 20)); bundle->rec_off.reserve(o.chunk + 64); bundle->rec_len.reserve(o.chunk + 64);
          bundle->raw.insert(bundle->raw.end(), tail.begin(), tail.end()); scan = 0;
          continue;
        }
        Bundle& b = *bundle;
        if (id > 0 && new_name) b.group_off.push_back(id);
        b.rec_off.push_back(scan + 4); b.rec_len.push_back((uint32_t)bs);
        if (new_name) prev_name = v.name();
        st.total_reads++;
        scan += 4 + (size_t)bs;
      }
    }
    if (producer_err.empty() && scan != bundle->raw.size()) { std::unique_lock<std::mutex> lk(m); if (err.empty()) throw std::runtime_error("truncated BAM record"); }
    bundle->raw.resize(scan);
    push_bundle();
    } catch (const std::exception& e) { producer_err = e.what(); }        // (truncated record, inflate / CRC error): report it after the threads are joined
    const auto t_loop_end = clk::now();
    { std::unique_lock<std::mutex> lk(m); eof = true; if (!producer_err.empty() && err.empty()) err = producer_err; } cv.notify_all();
    for (auto& t : threads) t.join();
    wthread.join();
    if (!err.empty()) { fprintf(stderr, err.compare(0, 6, "Error:") == 0 ? "%s\n" : "error: %s\n", err.c_str()); return 1; }
    const auto t_joined = clk::now();
    out.close();
    if (timing) {
      const double total = secs(t_begin, clk::now());
      t_ingest = secs(t_begin, t_loop_end) - t_read - t_wait_push;
      fprintf(stderr, "[timing] pipeline %.3fs: producer read+inflate %.3f, record walk %.3f, blocked on queue %.3f | pool: column extraction %.3f, "
                      "flush replay%s %.3f | engine project %.3f, copy-out %.3f | writer (%s) %.3f%s | final drain %.3f\n",
              total, t_read, t_ingest, t_wait_push, t_extract, dev_emit ? "" : " + assembly", t_assemble, t_project, t_store,
              dev_emit ? "records + BGZF on the device" : dev_deflate ? "BGZF on the device" : "BGZF host pool", t_write,
              dev_emit ? (" + brb_emit_bgzf (emitter threads) " + std::to_string(t_emit)).c_str() : "", secs(t_joined, clk::now()));
    }
    if (!o.quiet) {
      printf("\n[bramble_b200] final report:\n# input alignments:   %llu\n# unmapped reads:     %llu\n# dropped alignments: %llu\n# total alignments:   %llu\n# unique alignments:  %llu\n\n",
             (unsigned long long)st.total_reads, (unsigned long long)st.unmapped, (unsigned long long)st.dropped, (unsigned long long)st.total_complete, (unsigned long long)st.total_unique);
    }
  } catch (const std::exception& e) { fprintf(stderr, "error: %s\n", e.what()); return 1; }
  return 0;
}
