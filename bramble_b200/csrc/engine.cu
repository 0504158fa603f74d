// bramble_b200 engine: C ABI (include/bramble_b200.h) over the sm_100a kernels in this directory.
// Host side = memory plumbing, launch sequencing and the four count->scan->fill hand-offs; all
// projection arithmetic runs on the device.  There is no CPU path: without a device every compute
// entry point returns BRB_E_NO_DEVICE.
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "dev_common.cuh"
#include "scan.cuh"
#include "k_lookup.cuh"
#include "k_chain.cuh"
#include "k_ksw.cuh"
#include "k_pair.cuh"
#include "k_merge.cuh"

using namespace brb;

namespace {

thread_local std::string g_err;
int fail(int code, const char* fmt, ...) {
  char buf[512]; va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
  g_err = buf; return code;
}
#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return fail(BRB_E_CUDA, "%s: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

struct Buf {                       // growable device allocation
  void* p = nullptr; size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) { cudaFree(p); p = nullptr; cap = 0; }
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { cudaGetLastError(); e = cudaMalloc(&p, bytes); want = bytes; }
    if (e != cudaSuccess) { p = nullptr; return fail(BRB_E_NOMEM, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e)); }
    cap = want; return 0;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};
struct HBuf {                      // growable pinned host allocation
  void* p = nullptr; size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t e = cudaMallocHost(&p, want);
    if (e != cudaSuccess) { p = nullptr; return fail(BRB_E_NOMEM, "cudaMallocHost(%zu): %s", want, cudaGetErrorString(e)); }
    cap = want; return 0;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};
#define ENS(buf, bytes) do { int r_ = (buf).ensure(bytes); if (r_) return r_; } while (0)

inline unsigned grid_for(size_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }

}  // namespace

struct brb_index {
  int device = 0;
  DevIndex d{};
  Buf table_off, st, pmax, rec, chrom_off, genome;
  uint64_t bytes = 0;
};

struct brb_ctx {
  brb_index* idx = nullptr;
  cudaStream_t st = nullptr;
  cudaEvent_t ev[8]{};
  cudaEvent_t tev[2]{};
  // resident batch
  DevBatch b{};
  Buf in_group_off, in_refid, in_strand, in_exon_off, in_exon_st, in_exon_en, in_cigar_off, in_cigar, in_mate, in_seq_off, in_seq;
  uint64_t in_bytes = 0, n_mates = 0, seq_bytes = 0, rescue_bytes = 0;
  bool have_batch = false, have_result = false;
  // intermediates
  Buf q_lo, q_hi, t_ncand, t_coff, scan_tmp, counters;
  Buf c_rec, c_task, c_first, c_last, c_nseg, c_nguide, c_cap, c_ioff, ipool;
  Buf r_left, r_right, rpool, r_prob;
  Buf m_valid, m_tid, m_pos, m_sim, m_clip, m_ilen, c_partner;
  Buf u_nrows, u_roff, u_case, g_roff;
  Buf o_src, o_tid, o_pos, o_strand, o_flags, o_sim, o_clip, o_cand, o_ccap, o_coff, o_clen, opool;
  HBuf h_small, h_out;
  uint32_t n_cand = 0, n_rec = 0; uint64_t n_ideal_words = 0, n_out_words = 0;
  brb_stats stats{};
};

extern "C" {

int brb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
const char* brb_last_error(void) { return g_err.c_str(); }
const char* brb_version(void) { return "bramble_b200 0.1 (sm_100a)"; }

int brb_host_alloc(size_t bytes, void** out) {
  if (!out) return fail(BRB_E_ARG, "brb_host_alloc: null out");
  if (brb_device_count() <= 0) return fail(BRB_E_NO_DEVICE, "no CUDA device");
  CU(cudaMallocHost(out, bytes ? bytes : 1));
  return BRB_OK;
}
int brb_host_free(void* p) { if (p) cudaFreeHost(p); return BRB_OK; }

// ---- index: the g2t tree of the reference (src/bramble.cpp:132-211, src/g2t.cpp:34-62), flattened.
int brb_index_create(int device, const brb_index_desc* d, brb_index** out) {
  if (!d || !out) return fail(BRB_E_ARG, "brb_index_create: null argument");
  if (brb_device_count() <= device || device < 0) return fail(BRB_E_NO_DEVICE, "no CUDA device %d", device);
  if (d->n_tables & 1) return fail(BRB_E_ARG, "n_tables must be 2 * n_refids");
  if (d->n_tables && !d->table_off) return fail(BRB_E_ARG, "table_off is null");
  const uint64_t n = d->n_tables ? d->table_off[d->n_tables] : 0;
  if (n >= 0xfffffff0ull) return fail(BRB_E_ARG, "too many index records");
  if (n && (!d->st || !d->en || !d->tid || !d->exon_id || !d->pos_start || !d->prev_st || !d->prev_en || !d->next_st ||
            !d->next_en || !d->has_prev_next || !d->transcript_len)) return fail(BRB_E_ARG, "null index column");
  CU(cudaSetDevice(device));
  brb_index* ix = new brb_index(); ix->device = device;
  std::vector<uint32_t> toff(d->n_tables + 1, 0);
  std::vector<int32_t> pmax(n);
  std::vector<Rec> rec(n);
  for (uint32_t t = 0; t < d->n_tables; t++) {
    toff[t] = (uint32_t)d->table_off[t];
    if (d->table_off[t + 1] < d->table_off[t]) { delete ix; return fail(BRB_E_ARG, "table_off not monotone"); }
    int32_t m = INT32_MIN, prev = INT32_MIN;
    for (uint64_t i = d->table_off[t]; i < d->table_off[t + 1]; i++) {
      if (d->st[i] < prev) { delete ix; return fail(BRB_E_ARG, "table %u is not sorted by start", t); }
      prev = d->st[i];
      m = std::max(m, d->en[i]); pmax[i] = m;
      Rec& r = rec[i];
      r.st = d->st[i]; r.en = d->en[i]; r.tid = d->tid[i]; r.pos_start = d->pos_start[i];
      r.prev_st = d->prev_st[i]; r.prev_en = d->prev_en[i]; r.next_st = d->next_st[i]; r.next_en = d->next_en[i];
      r.tlen = d->transcript_len[i];
      r.meta = (uint32_t)d->exon_id[i] | ((uint32_t)(d->has_prev_next[i] & 1) << 8) | ((uint32_t)((d->has_prev_next[i] >> 1) & 1) << 9);
      r.pad0 = r.pad1 = 0;
    }
  }
  toff[d->n_tables] = (uint32_t)n;
  int rc = 0;
  auto up = [&](Buf& b, const void* src, size_t bytes) -> int {
    int r = b.ensure(bytes ? bytes : 16); if (r) return r;
    if (bytes && cudaMemcpy(b.p, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) return fail(BRB_E_CUDA, "index upload failed");
    ix->bytes += bytes; return 0;
  };
  if (!rc) rc = up(ix->table_off, toff.data(), toff.size() * 4);
  if (!rc) rc = up(ix->st, d->st, n * 4);
  if (!rc) rc = up(ix->pmax, pmax.data(), n * 4);
  if (!rc) rc = up(ix->rec, rec.data(), n * sizeof(Rec));
  ix->d.n_tables = d->n_tables; ix->d.n_rec = n; ix->d.n_chrom = 0;
  if (!rc && d->genome_codes && d->chrom_off && d->n_chrom) {
    rc = up(ix->chrom_off, d->chrom_off, (size_t)(d->n_chrom + 1) * 8);
    if (!rc) rc = up(ix->genome, d->genome_codes, d->chrom_off[d->n_chrom]);
    ix->d.n_chrom = d->n_chrom;
  }
  if (rc) { ix->table_off.release(); ix->st.release(); ix->pmax.release(); ix->rec.release(); ix->chrom_off.release(); ix->genome.release(); delete ix; return rc; }
  ix->d.table_off = ix->table_off.as<uint32_t>(); ix->d.st = ix->st.as<int32_t>(); ix->d.pmax = ix->pmax.as<int32_t>();
  ix->d.rec = ix->rec.as<Rec>();
  ix->d.chrom_off = ix->d.n_chrom ? ix->chrom_off.as<uint64_t>() : nullptr;
  ix->d.genome = ix->d.n_chrom ? ix->genome.as<uint8_t>() : nullptr;
  *out = ix;
  return BRB_OK;
}

int brb_index_destroy(brb_index* ix) {
  if (!ix) return BRB_OK;
  cudaSetDevice(ix->device);
  ix->table_off.release(); ix->st.release(); ix->pmax.release(); ix->rec.release(); ix->chrom_off.release(); ix->genome.release();
  delete ix;
  return BRB_OK;
}

int brb_ctx_create(brb_index* ix, brb_ctx** out) {
  if (!ix || !out) return fail(BRB_E_ARG, "brb_ctx_create: null argument");
  CU(cudaSetDevice(ix->device));
  brb_ctx* c = new brb_ctx(); c->idx = ix;
  CU(cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking));
  for (auto& e : c->ev) CU(cudaEventCreate(&e));
  for (auto& e : c->tev) CU(cudaEventCreate(&e));
  ENS(c->counters, 64); ENS(c->h_small, 256);
  *out = c;
  return BRB_OK;
}

int brb_ctx_destroy(brb_ctx* c) {
  if (!c) return BRB_OK;
  cudaSetDevice(c->idx->device);
  cudaStreamSynchronize(c->st);
  Buf* all[] = {&c->in_group_off, &c->in_refid, &c->in_strand, &c->in_exon_off, &c->in_exon_st, &c->in_exon_en, &c->in_cigar_off,
                &c->in_cigar, &c->in_mate, &c->in_seq_off, &c->in_seq, &c->q_lo, &c->q_hi, &c->t_ncand, &c->t_coff, &c->scan_tmp,
                &c->counters, &c->c_rec, &c->c_task, &c->c_first, &c->c_last, &c->c_nseg, &c->c_nguide, &c->c_cap, &c->c_ioff,
                &c->ipool, &c->r_left, &c->r_right, &c->rpool, &c->r_prob, &c->m_valid, &c->m_tid, &c->m_pos, &c->m_sim, &c->m_clip,
                &c->m_ilen, &c->c_partner, &c->u_nrows, &c->u_roff, &c->u_case, &c->g_roff, &c->o_src, &c->o_tid, &c->o_pos,
                &c->o_strand, &c->o_flags, &c->o_sim, &c->o_clip, &c->o_cand, &c->o_ccap, &c->o_coff, &c->o_clen, &c->opool};
  for (Buf* b : all) b->release();
  c->h_small.release(); c->h_out.release();
  for (auto& e : c->ev) if (e) cudaEventDestroy(e);
  for (auto& e : c->tev) if (e) cudaEventDestroy(e);
  if (c->st) cudaStreamDestroy(c->st);
  delete c;
  return BRB_OK;
}

// ---- batch upload: one bundle (reference BundleData::reads) as structure-of-arrays.
int brb_batch_upload(brb_ctx* c, const brb_batch_in* in) {
  if (!c || !in) return fail(BRB_E_ARG, "brb_batch_upload: null argument");
  CU(cudaSetDevice(c->idx->device));
  c->have_batch = c->have_result = false;
  const uint32_t n = in->n_aln, ng = in->n_groups;
  if (n && (!in->group_off || !in->refid || !in->strand || !in->exon_off || !in->exon_st || !in->exon_en || !in->cigar_off ||
            !in->cigar || !in->mate_idx)) return fail(BRB_E_ARG, "null batch column");
  if (n >= 0x7ffffff0u) return fail(BRB_E_ARG, "batch too large");
  const uint32_t n_exon = n ? in->exon_off[n] : 0, n_cigar = n ? in->cigar_off[n] : 0;
  if (n && (in->group_off[0] != 0 || in->group_off[ng] != n || in->exon_off[0] != 0 || in->cigar_off[0] != 0))
    return fail(BRB_E_ARG, "inconsistent CSR offsets");
  uint64_t n_mates = 0;
  for (uint32_t a = 0; a < n; a++) {        // process_pairs links are single and symmetric (src/bramble.cpp:272-311)
    const int32_t m = in->mate_idx[a];
    if (m < 0) continue;
    if ((uint32_t)m >= n || (uint32_t)m == a || in->mate_idx[m] != (int32_t)a) return fail(BRB_E_ARG, "mate_idx[%u]=%d is not a symmetric link", a, m);
    n_mates++;
  }
  c->in_bytes = 0;
  auto up = [&](Buf& b, const void* src, size_t bytes) -> int {
    int r = b.ensure(bytes ? bytes : 16); if (r) return r;
    if (bytes && cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->st) != cudaSuccess) return fail(BRB_E_CUDA, "batch upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    c->in_bytes += bytes; return 0;
  };
  int rc = 0;
  static const uint32_t zero_off[1] = {0};
  if (!rc) rc = up(c->in_group_off, n ? in->group_off : zero_off, n ? (size_t)(ng + 1) * 4 : 4);
  if (!rc) rc = up(c->in_refid, in->refid, (size_t)n * 4);
  if (!rc) rc = up(c->in_strand, in->strand, (size_t)n);
  if (!rc) rc = up(c->in_exon_off, n ? in->exon_off : zero_off, n ? (size_t)(n + 1) * 4 : 4);
  if (!rc) rc = up(c->in_exon_st, in->exon_st, (size_t)n_exon * 4);
  if (!rc) rc = up(c->in_exon_en, in->exon_en, (size_t)n_exon * 4);
  if (!rc) rc = up(c->in_cigar_off, n ? in->cigar_off : zero_off, n ? (size_t)(n + 1) * 4 : 4);
  if (!rc) rc = up(c->in_cigar, in->cigar, (size_t)n_cigar * 4);
  if (!rc) rc = up(c->in_mate, in->mate_idx, (size_t)n * 4);
  c->seq_bytes = 0;
  const bool have_seq = n && in->group_seq_off && in->seq_codes;
  if (!rc && have_seq) {
    rc = up(c->in_seq_off, in->group_seq_off, (size_t)(ng + 1) * 8);
    if (!rc) rc = up(c->in_seq, in->seq_codes, (size_t)in->group_seq_off[ng]);
    c->seq_bytes = in->group_seq_off[ng];
  }
  if (rc) return rc;
  DevBatch& b = c->b;
  b.n_aln = n; b.n_groups = n ? ng : 0; b.n_exon = n_exon; b.n_cigar = n_cigar;
  b.group_off = c->in_group_off.as<uint32_t>(); b.refid = c->in_refid.as<int32_t>(); b.strand = c->in_strand.as<int8_t>();
  b.exon_off = c->in_exon_off.as<uint32_t>(); b.exon_st = c->in_exon_st.as<uint32_t>(); b.exon_en = c->in_exon_en.as<uint32_t>();
  b.cigar_off = c->in_cigar_off.as<uint32_t>(); b.cigar = c->in_cigar.as<uint32_t>(); b.mate_idx = c->in_mate.as<int32_t>();
  b.group_seq_off = have_seq ? c->in_seq_off.as<uint64_t>() : nullptr; b.seq = have_seq ? c->in_seq.as<uint8_t>() : nullptr;
  b.aln_group = nullptr;
  c->n_mates = n_mates;
  c->stats.h2d_bytes = c->in_bytes;
  c->have_batch = true;
  return BRB_OK;
}

}  // extern "C"

namespace {

int run_rescue(brb_ctx* c, const brb_params& p);   // k_ksw host driver, below

// read back `n` uint32 values at device pointer `src` (synchronises the stream)
int read_u32(brb_ctx* c, const uint32_t* src, uint32_t* dst, size_t n) {
  CU(cudaMemcpyAsync(c->h_small.p, src, n * 4, cudaMemcpyDeviceToHost, c->st));
  CU(cudaStreamSynchronize(c->st));
  memcpy(dst, c->h_small.p, n * 4);
  return 0;
}

float ms_between(cudaEvent_t a, cudaEvent_t b) { float ms = 0; cudaEventElapsedTime(&ms, a, b); return ms; }

int project_resident(brb_ctx* c, const brb_params& p) {
  if (!c->have_batch) return fail(BRB_E_ARG, "no batch resident: call brb_batch_upload first");
  const bool rescue = p.long_reads && p.use_fasta;
  if (rescue && !c->idx->d.genome) return fail(BRB_E_UNSUPPORTED, "use_fasta requested but the index holds no genome");
  if (rescue && c->b.n_aln && !c->b.seq) return fail(BRB_E_UNSUPPORTED, "use_fasta requested but the batch carries no read sequences");
  const DevBatch& b = c->b; const DevIndex& ix = c->idx->d;
  cudaStream_t st = c->st;
  brb_stats& S = c->stats;
  uint64_t L = 0;
  const uint32_t n = b.n_aln, n_task = 2 * n;
  c->have_result = false; c->rescue_bytes = 0; c->n_cand = 0; c->n_rec = 0; c->n_out_words = 0; c->n_ideal_words = 0;
  S.n_candidates = S.n_matches = S.n_rescue = S.n_rec = 0;
  S.ms_lookup = S.ms_chain = S.ms_rescue = S.ms_finalize = S.ms_pair = S.ms_merge = 0; S.kernel_ms = 0;
  ENS(c->g_roff, (size_t)(b.n_groups + 1) * 4);
  if (n == 0) {
    CU(cudaMemsetAsync(c->g_roff.p, 0, 4, st)); CU(cudaStreamSynchronize(st));
    S.n_launches = 0; S.alg_bytes = 0; c->have_result = true; return BRB_OK;
  }
  unsigned long long* counters = c->counters.as<unsigned long long>();   // [0] n_hits [1] out cigar words [2] rescue bytes
  CU(cudaMemsetAsync(counters, 0, 64, st));
  ENS(c->q_lo, (size_t)2 * b.n_exon * 4 + 16); ENS(c->q_hi, (size_t)2 * b.n_exon * 4 + 16);
  ENS(c->t_ncand, (size_t)n_task * 4); ENS(c->t_coff, (size_t)(n_task + 1) * 4);
  ENS(c->u_nrows, (size_t)n * 4); ENS(c->u_roff, (size_t)(n + 1) * 4); ENS(c->u_case, n);
  ENS(c->scan_tmp, scan_scratch_words(std::max<size_t>(n_task, 1)) * 4);

  // ---- phase 1: lookup
  CU(cudaEventRecord(c->ev[0], st));
  k_lookup<<<grid_for(n_task, 256), 256, 0, st>>>(ix, b, p, c->q_lo.as<uint32_t>(), c->q_hi.as<uint32_t>(), c->t_ncand.as<uint32_t>(), counters); L++;
  exclusive_scan(c->t_ncand.as<uint32_t>(), c->t_coff.as<uint32_t>(), n_task, c->scan_tmp.as<uint32_t>(), st, &L);
  uint32_t C = 0;
  { int r = read_u32(c, c->t_coff.as<uint32_t>() + n_task, &C, 1); if (r) return r; }
  c->n_cand = C; S.n_candidates = C;
  const size_t Cz = std::max<uint32_t>(C, 1);
  ENS(c->c_rec, Cz * 4); ENS(c->c_task, Cz * 4); ENS(c->c_first, Cz * 4); ENS(c->c_last, Cz * 4); ENS(c->c_nseg, Cz * 4);
  ENS(c->c_nguide, Cz * 4); ENS(c->c_cap, Cz * 4); ENS(c->c_ioff, (Cz + 1) * 4);
  ENS(c->m_valid, Cz); ENS(c->m_tid, Cz * 4); ENS(c->m_pos, Cz * 4); ENS(c->m_sim, Cz * 8); ENS(c->m_clip, Cz * 4); ENS(c->m_ilen, Cz * 4);
  ENS(c->c_partner, Cz * 4);
  ENS(c->scan_tmp, scan_scratch_words(std::max<size_t>(std::max<size_t>(n_task, Cz), 1)) * 4);
  if (C) { k_fill_cand<<<grid_for(n_task, 256), 256, 0, st>>>(ix, b, p, c->q_lo.as<uint32_t>(), c->q_hi.as<uint32_t>(), c->t_coff.as<uint32_t>(),
                                                              c->c_rec.as<uint32_t>(), c->c_task.as<uint32_t>()); L++; }
  CU(cudaEventRecord(c->ev[1], st));

  // ---- phase 2: chain walk, pass 1
  uint32_t I = 0;
  if (C) {
    k_chain<<<grid_for(C, 256), 256, 0, st>>>(ix, b, p, C, c->q_lo.as<uint32_t>(), c->q_hi.as<uint32_t>(), c->c_rec.as<uint32_t>(),
                                              c->c_task.as<uint32_t>(), c->c_first.as<uint32_t>(), c->c_last.as<uint32_t>(),
                                              c->c_nseg.as<uint32_t>(), c->c_nguide.as<uint32_t>(), c->c_cap.as<uint32_t>()); L++;
  }
  CU(cudaEventRecord(c->ev[2], st));

  // ---- phase 3: long-read clip rescue (ksw2), adds its CIGAR lengths to c_cap
  if (C && rescue) { S.n_launches = L; int r = run_rescue(c, p); if (r) return r; L = S.n_launches; }
  CU(cudaEventRecord(c->ev[3], st));

  // ---- phase 4: pass 2 (ideal CIGAR, scores, similarity filter)
  if (C) {
    exclusive_scan(c->c_cap.as<uint32_t>(), c->c_ioff.as<uint32_t>(), C, c->scan_tmp.as<uint32_t>(), st, &L);
    { int r = read_u32(c, c->c_ioff.as<uint32_t>() + C, &I, 1); if (r) return r; }
    c->n_ideal_words = I;
    ENS(c->ipool, (size_t)std::max<uint32_t>(I, 1) * 4);
    k_build<<<grid_for(C, 256), 256, 0, st>>>(ix, b, p, C, c->q_lo.as<uint32_t>(), c->q_hi.as<uint32_t>(), c->c_task.as<uint32_t>(),
                                              c->c_first.as<uint32_t>(), c->c_last.as<uint32_t>(), c->c_nseg.as<uint32_t>(),
                                              c->c_nguide.as<uint32_t>(), c->c_cap.as<uint32_t>(), c->c_ioff.as<uint32_t>(),
                                              rescue ? c->r_left.as<RescueSeg>() : nullptr, rescue ? c->r_right.as<RescueSeg>() : nullptr,
                                              rescue ? c->rpool.as<uint32_t>() : nullptr, c->ipool.as<uint32_t>(),
                                              c->m_valid.as<uint8_t>(), c->m_tid.as<uint32_t>(), c->m_pos.as<uint32_t>(),
                                              c->m_sim.as<double>(), c->m_clip.as<int32_t>(), c->m_ilen.as<uint32_t>(), counters + 3); L++;
  }
  CU(cudaEventRecord(c->ev[4], st));

  // ---- phase 5: pairing + emission
  PairView pv; pv.t_coff = c->t_coff.as<uint32_t>(); pv.m_valid = c->m_valid.as<uint8_t>(); pv.m_tid = c->m_tid.as<uint32_t>();
  k_pair_count<<<grid_for(n, 256), 256, 0, st>>>(b, pv, c->c_partner.as<uint32_t>(), c->u_nrows.as<uint32_t>(), c->u_case.as<uint8_t>()); L++;
  exclusive_scan(c->u_nrows.as<uint32_t>(), c->u_roff.as<uint32_t>(), n, c->scan_tmp.as<uint32_t>(), st, &L);
  uint32_t R = 0;
  { int r = read_u32(c, c->u_roff.as<uint32_t>() + n, &R, 1); if (r) return r; }
  c->n_rec = R; S.n_rec = R;
  const size_t Rz = std::max<uint32_t>(R, 1);
  ENS(c->o_src, Rz * 4); ENS(c->o_tid, Rz * 4); ENS(c->o_pos, Rz * 4); ENS(c->o_strand, Rz); ENS(c->o_flags, Rz); ENS(c->o_sim, Rz * 8);
  ENS(c->o_clip, Rz * 4); ENS(c->o_cand, Rz * 4); ENS(c->o_ccap, Rz * 4); ENS(c->o_coff, (Rz + 1) * 4); ENS(c->o_clen, Rz * 4);
  ENS(c->scan_tmp, scan_scratch_words(std::max<size_t>(std::max<size_t>(n_task, Cz), Rz)) * 4);
  k_group_rows<<<grid_for(b.n_groups + 1, 256), 256, 0, st>>>(b, c->u_roff.as<uint32_t>(), c->g_roff.as<uint32_t>()); L++;
  uint32_t W = 0;
  if (R) {
    Rows o; o.src = c->o_src.as<uint32_t>(); o.tid = c->o_tid.as<uint32_t>(); o.pos = c->o_pos.as<uint32_t>(); o.strand = c->o_strand.as<int8_t>();
    o.flags = c->o_flags.as<uint8_t>(); o.sim = c->o_sim.as<double>(); o.clip = c->o_clip.as<int32_t>(); o.cand = c->o_cand.as<uint32_t>();
    o.ccap = c->o_ccap.as<uint32_t>();
    k_pair_emit<<<grid_for(n, 256), 256, 0, st>>>(b, pv, c->c_partner.as<uint32_t>(), c->u_roff.as<uint32_t>(), c->u_case.as<uint8_t>(),
                                                  c->c_task.as<uint32_t>(), c->m_pos.as<uint32_t>(), c->m_sim.as<double>(),
                                                  c->m_clip.as<int32_t>(), c->m_ilen.as<uint32_t>(), o); L++;
    exclusive_scan(c->o_ccap.as<uint32_t>(), c->o_coff.as<uint32_t>(), R, c->scan_tmp.as<uint32_t>(), st, &L);
    { int r = read_u32(c, c->o_coff.as<uint32_t>() + R, &W, 1); if (r) return r; }
  }
  c->n_out_words = W;
  CU(cudaEventRecord(c->ev[5], st));

  // ---- phase 6: CIGAR merge
  if (R) {
    ENS(c->opool, (size_t)std::max<uint32_t>(W, 1) * 4);
    k_merge<<<grid_for(R, 256), 256, 0, st>>>(b, R, c->o_src.as<uint32_t>(), c->o_cand.as<uint32_t>(), c->o_strand.as<int8_t>(),
                                              c->o_coff.as<uint32_t>(), c->c_ioff.as<uint32_t>(), c->m_ilen.as<uint32_t>(),
                                              c->ipool.as<uint32_t>(), c->o_clen.as<uint32_t>(), c->opool.as<uint32_t>(), counters + 1); L++;
  }
  CU(cudaEventRecord(c->ev[6], st));
  CU(cudaMemcpyAsync(c->h_small.p, counters, 64, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  unsigned long long hc[8]; memcpy(hc, c->h_small.p, 64);
  S.n_launches = L; S.n_matches = hc[3];
  S.ms_lookup = ms_between(c->ev[0], c->ev[1]); S.ms_chain = ms_between(c->ev[1], c->ev[2]); S.ms_rescue = ms_between(c->ev[2], c->ev[3]);
  S.ms_finalize = ms_between(c->ev[3], c->ev[4]); S.ms_pair = ms_between(c->ev[4], c->ev[5]); S.ms_merge = ms_between(c->ev[5], c->ev[6]);
  S.kernel_ms = ms_between(c->ev[0], c->ev[6]);
  // algorithmic bytes, SURVEY.md 8(d) / DESIGN.md: per-alignment SoA + 56 B per index hit + output rows + rescue bases
  S.alg_bytes = (uint64_t)n * 16 + (uint64_t)b.n_exon * 8 + (uint64_t)b.n_cigar * 4 + c->n_mates * 4 + hc[0] * 56 +
                (uint64_t)R * 40 + hc[1] * 4 + c->rescue_bytes;
  S.n_rescue = hc[4];
  c->have_result = true;
  return BRB_OK;
}

}  // namespace

extern "C" {

int brb_project_resident(brb_ctx* c, const brb_params* p) {
  if (!c || !p) return fail(BRB_E_ARG, "brb_project_resident: null argument");
  if (brb_device_count() <= 0) return fail(BRB_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(c->idx->device));
  return project_resident(c, *p);
}

// Rows come back in one pinned staging block owned by the ctx (valid until the next call on it).
int brb_batch_download(brb_ctx* c, brb_batch_out* out) {
  if (!c || !out) return fail(BRB_E_ARG, "brb_batch_download: null argument");
  if (!c->have_result) return fail(BRB_E_ARG, "no result resident: call brb_project_resident first");
  CU(cudaSetDevice(c->idx->device));
  const size_t R = c->n_rec, W = c->n_out_words, G = c->b.n_groups;
  auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };
  size_t off = 0;
  const size_t o_groff = off; off += al((G + 1) * 4);
  const size_t o_src = off; off += al(R * 4);
  const size_t o_tid = off; off += al(R * 4);
  const size_t o_pos = off; off += al(R * 4);
  const size_t o_sim = off; off += al(R * 8);
  const size_t o_clip = off; off += al(R * 4);
  const size_t o_coff = off; off += al(R * 4);
  const size_t o_clen = off; off += al(R * 4);
  const size_t o_strand = off; off += al(R);
  const size_t o_flags = off; off += al(R);
  const size_t o_cig = off; off += al(W * 4);
  ENS(c->h_out, off + 16);
  uint8_t* h = c->h_out.as<uint8_t>();
  uint64_t bytes = 0;
  auto dn = [&](size_t dst, const Buf& b, size_t n) -> int {
    if (!n) return 0;
    if (cudaMemcpyAsync(h + dst, b.p, n, cudaMemcpyDeviceToHost, c->st) != cudaSuccess) return fail(BRB_E_CUDA, "download failed: %s", cudaGetErrorString(cudaGetLastError()));
    bytes += n; return 0;
  };
  int rc = dn(o_groff, c->g_roff, (G + 1) * 4);
  if (!rc) rc = dn(o_src, c->o_src, R * 4);
  if (!rc) rc = dn(o_tid, c->o_tid, R * 4);
  if (!rc) rc = dn(o_pos, c->o_pos, R * 4);
  if (!rc) rc = dn(o_sim, c->o_sim, R * 8);
  if (!rc) rc = dn(o_clip, c->o_clip, R * 4);
  if (!rc) rc = dn(o_coff, c->o_coff, R * 4);
  if (!rc) rc = dn(o_clen, c->o_clen, R * 4);
  if (!rc) rc = dn(o_strand, c->o_strand, R);
  if (!rc) rc = dn(o_flags, c->o_flags, R);
  if (!rc) rc = dn(o_cig, c->opool, W * 4);
  if (rc) return rc;
  CU(cudaStreamSynchronize(c->st));
  out->n_rec = (uint32_t)R; out->n_cigar_words = W;
  out->group_rec_off = (const uint32_t*)(h + o_groff); out->src_aln = (const uint32_t*)(h + o_src); out->tid = (const uint32_t*)(h + o_tid);
  out->pos = (const uint32_t*)(h + o_pos); out->tx_strand = (const int8_t*)(h + o_strand); out->flags = h + o_flags;
  out->sim_score = (const double*)(h + o_sim); out->clip_score = (const int32_t*)(h + o_clip);
  out->cigar_off = (const uint32_t*)(h + o_coff); out->cigar_len = (const uint32_t*)(h + o_clen); out->cigar = (const uint32_t*)(h + o_cig);
  c->stats.d2h_bytes = bytes;
  return BRB_OK;
}

int brb_project(brb_ctx* c, const brb_params* p, const brb_batch_in* in, brb_batch_out* out) {
  int rc = brb_batch_upload(c, in); if (rc) return rc;
  rc = brb_project_resident(c, p); if (rc) return rc;
  return brb_batch_download(c, out);
}

int brb_ctx_timer_start(brb_ctx* c) {
  if (!c) return fail(BRB_E_ARG, "null ctx");
  CU(cudaSetDevice(c->idx->device));
  CU(cudaEventRecord(c->tev[0], c->st));
  return BRB_OK;
}
int brb_ctx_timer_stop(brb_ctx* c, float* ms) {
  if (!c || !ms) return fail(BRB_E_ARG, "null argument");
  CU(cudaSetDevice(c->idx->device));
  CU(cudaEventRecord(c->tev[1], c->st));
  CU(cudaEventSynchronize(c->tev[1]));
  CU(cudaEventElapsedTime(ms, c->tev[0], c->tev[1]));
  return BRB_OK;
}

int brb_ctx_stats(const brb_ctx* c, brb_stats* out) {
  if (!c || !out) return fail(BRB_E_ARG, "brb_ctx_stats: null argument");
  *out = c->stats; return BRB_OK;
}

// ---- standalone kernels for the parity tests (host buffers in, host buffers out)
int brb_lookup(brb_ctx* c, const brb_params* p, uint32_t nq, const int32_t* refid, const int8_t* strand, const uint32_t* qs,
               const uint32_t* qe, const uint8_t* status, uint32_t cap, uint32_t* hits, uint32_t* n_hits) {
  if (!c || !p || (nq && (!refid || !strand || !qs || !qe || !status || !n_hits))) return fail(BRB_E_ARG, "brb_lookup: null argument");
  if (brb_device_count() <= 0) return fail(BRB_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(c->idx->device));
  if (!nq) return BRB_OK;
  Buf d_ref, d_str, d_qs, d_qe, d_st, d_hits, d_n;
  int rc = 0;
  auto up = [&](Buf& b, const void* src, size_t bytes) { if (rc) return; rc = b.ensure(bytes); if (!rc && cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->st) != cudaSuccess) rc = fail(BRB_E_CUDA, "upload failed"); };
  up(d_ref, refid, (size_t)nq * 4); up(d_str, strand, nq); up(d_qs, qs, (size_t)nq * 4); up(d_qe, qe, (size_t)nq * 4); up(d_st, status, nq);
  if (!rc) rc = d_hits.ensure((size_t)nq * std::max<uint32_t>(cap, 1) * 4);
  if (!rc) rc = d_n.ensure((size_t)nq * 4);
  if (!rc) {
    k_lookup_queries<<<grid_for(nq, 128), 128, 0, c->st>>>(c->idx->d, *p, nq, d_ref.as<int32_t>(), d_str.as<int8_t>(), d_qs.as<uint32_t>(),
                                                           d_qe.as<uint32_t>(), d_st.as<uint8_t>(), cap, d_hits.as<uint32_t>(), d_n.as<uint32_t>());
    if (cap && hits) cudaMemcpyAsync(hits, d_hits.p, (size_t)nq * cap * 4, cudaMemcpyDeviceToHost, c->st);
    cudaMemcpyAsync(n_hits, d_n.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, c->st);
    cudaError_t e = cudaStreamSynchronize(c->st);
    if (e != cudaSuccess) rc = fail(BRB_E_CUDA, "brb_lookup: %s", cudaGetErrorString(e));
  }
  d_ref.release(); d_str.release(); d_qs.release(); d_qe.release(); d_st.release(); d_hits.release(); d_n.release();
  return rc;
}

int brb_merge_cigars(brb_ctx* c, uint32_t n, const uint32_t* real_off, const uint32_t* real, const uint32_t* ideal_off,
                     const uint32_t* ideal, uint32_t* out_off, uint32_t* out_len, uint32_t* out) {
  if (!c || (n && (!real_off || !ideal_off || !out_off || !out_len || !out))) return fail(BRB_E_ARG, "brb_merge_cigars: null argument");
  if (brb_device_count() <= 0) return fail(BRB_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(c->idx->device));
  if (!n) return BRB_OK;
  const size_t nr = real_off[n], ni = ideal_off[n], no = out_off[n];
  Buf d_ro, d_r, d_io, d_i, d_oo, d_ol, d_o;
  int rc = 0;
  auto up = [&](Buf& b, const void* src, size_t bytes) { if (rc) return; rc = b.ensure(bytes ? bytes : 4); if (!rc && bytes && cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->st) != cudaSuccess) rc = fail(BRB_E_CUDA, "upload failed"); };
  up(d_ro, real_off, (size_t)(n + 1) * 4); up(d_r, real, nr * 4); up(d_io, ideal_off, (size_t)(n + 1) * 4); up(d_i, ideal, ni * 4);
  up(d_oo, out_off, (size_t)(n + 1) * 4);
  if (!rc) rc = d_ol.ensure((size_t)n * 4);
  if (!rc) rc = d_o.ensure(std::max<size_t>(no, 1) * 4);
  if (!rc) {
    k_merge_pairs<<<grid_for(n, 128), 128, 0, c->st>>>(n, d_ro.as<uint32_t>(), d_r.as<uint32_t>(), d_io.as<uint32_t>(), d_i.as<uint32_t>(),
                                                       d_oo.as<uint32_t>(), d_ol.as<uint32_t>(), d_o.as<uint32_t>());
    cudaMemcpyAsync(out_len, d_ol.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->st);
    if (no) cudaMemcpyAsync(out, d_o.p, no * 4, cudaMemcpyDeviceToHost, c->st);
    cudaError_t e = cudaStreamSynchronize(c->st);
    if (e != cudaSuccess) rc = fail(BRB_E_CUDA, "brb_merge_cigars: %s", cudaGetErrorString(e));
  }
  d_ro.release(); d_r.release(); d_io.release(); d_i.release(); d_oo.release(); d_ol.release(); d_o.release();
  return rc;
}

}  // extern "C"
#include "rescue_host.inl"
