// bramble_b200 engine: C ABI (include/bramble_b200.h) over the sm_100a kernels in this directory.
// Host side = memory plumbing, launch sequencing and the four count->scan->fill hand-offs; all
// projection arithmetic runs on the device.  There is no CPU path: without a device every compute
// entry point returns BRB_E_NO_DEVICE.
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <mutex>
#include <string>
#include <vector>

#include "dev_common.cuh"
#include "scan.cuh"
#include "k_lookup.cuh"
#include "k_chain.cuh"
#include "k_fused.cuh"
#include "k_front_fast.cuh"
#include "k_ksw.cuh"
#include "k_back.cuh"
#include "k_merge.cuh"
#include "k_flush.cuh"
#include "k_index.cuh"
#include "k_deflate.cuh"
#include "k_emit.cuh"
#include "k_inflate.cuh"
#include "k_expand.cuh"

using namespace brb;

// ---- brb_merge_cigars_static: the scratch of the warp-cooperative CIGAR merge in a fixed array instead of a cudaMalloc'ed pool.
// Every GPU that loads the library holds these 4 MiB (+ 256 bytes) until the process exits.  In HOST code the name below is only a
// stand-in for the symbol: the device address comes from cudaGetSymbolAddress (static_scratch(), once per device).
#define BRB_STATIC_SCRATCH_WORDS (1u << 20)
static __device__ __align__(256) uint32_t g_static_scratch[64 + BRB_STATIC_SCRATCH_WORDS];     // [0..63]: status words (word 0 = pairs merged serially)

namespace {
// The pairs of brb_merge_cigars, persistent warps: warp w owns the slice scr + w * slice_words; a pair that needs more than a slice
// (the caller's bounds were too small) is merged by lane 0 with the serial code, which needs no scratch, and counted in status[0].
__global__ void __launch_bounds__(128) k_merge_pairs_static(uint32_t n, const uint32_t* __restrict__ real_off, const uint32_t* __restrict__ real,
                                                            const uint32_t* __restrict__ ideal_off, const uint32_t* __restrict__ ideal,
                                                            const uint32_t* __restrict__ out_off, uint32_t* __restrict__ out_len, uint32_t* __restrict__ out,
                                                            uint32_t* __restrict__ scr, uint32_t slice_words, uint32_t* __restrict__ status) {
  const unsigned lane = threadIdx.x & 31;
  const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
  uint32_t* mine = scr + (size_t)w * slice_words;
  for (uint32_t i = w; i < n; i += n_warps) {
    const uint32_t r0 = real_off[i], nR = real_off[i + 1] - r0, i0 = ideal_off[i], nI = ideal_off[i + 1] - i0;
    const uint32_t iword = nI == 1 ? ideal[i0] : 0;
    uint32_t len = 0;
    if (merge_warp_scratch(nR, nI) <= slice_words) len = merge_cigars_any_warp(real + r0, nR, ideal + i0, nI, iword, out + out_off[i], mine, lane);
    else {
      if (lane == 0) { len = merge_cigars(real + r0, nR, ideal + i0, nI, iword, out + out_off[i]); atomicAdd(&status[0], 1u); }
      len = __shfl_sync(0xffffffffu, len, 0);
    }
    if (lane == 0) out_len[i] = len;
    __syncwarp();                                           // the slice is reused by the warp's next pair
  }
}
__global__ void k_copy_status(const uint32_t* __restrict__ status, uint32_t* __restrict__ dst) { *dst = status[0]; }
}  // namespace

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

struct RescueBufs;                       // rescue_host.inl

}  // namespace

struct brb_ctx;
namespace { void free_rescue_bufs(brb_ctx* c);
struct RowView { const uint32_t* src; const uint32_t* tid; const uint32_t* pos; const int8_t* strand; const uint8_t* flags; const double* sim; const int32_t* clip; const uint32_t* coff; const uint32_t* pool; };
}
struct brb_index {
  int device = 0;
  DevIndex d{};
  Buf table_off, st, pmax, rec, gprev, bin_off, bins, chrom_off, genome;
  uint64_t bytes = 0;
  void release_all() { for (Buf* b : {&table_off, &st, &pmax, &rec, &gprev, &bin_off, &bins, &chrom_off, &genome}) b->release(); }
};

struct brb_ctx {
  brb_index* idx = nullptr;
  cudaStream_t st = nullptr;
  cudaEvent_t ev[8]{};
  bool ctl_clean = false;           // the device Ctl holds zero counters and ctl_caps (left so by the last sequence's publisher)
  uint32_t ctl_caps[N_CAPS]{};
  int n_sm = 0;                     // cached launch geometry: SM count and resident CTAs per kernel entry
  std::vector<std::pair<const void*, int>> resident_cache;
  cudaEvent_t tev[2]{};
  // resident batch
  DevBatch b{};
  Buf in_group_off, in_refid, in_strand, in_exon_off, in_exon_st, in_exon_en, in_cigar_off, in_cigar, in_mate, in_seq_off, in_seq, in_name_off, in_names;
  Buf f_hi, f_nh, f_mapq, f_prim;   // BRB_OPT_FLUSH_TAGS results per row
  Buf z_in, z_slots, z_sizes, z_crcs, z_tot, z_off, z_out, z_ticket; HBuf z_host;   // brb_bgzf_compress
  Buf e_raw, e_off, e_len, e_units, e_plan, e_sizes, e_offs, e_err;                // brb_emit_bgzf
  Buf i_in, i_meta, i_out, i_ctl; HBuf i_host, i_hmeta;                           // brb_bgzf_inflate
  Buf x_cnt16, x_cnt32, x_md16, x_r16, x_pos, x_err;                              // compact input columns (k_expand.cuh)
  bool expand_pending = false, expand_exons = false, expand_cigar = false; uint32_t expand_n_cigar = 0;   // k_expand_all has run for the resident batch, its totals / verdict not read yet
  bool have_names = false, opt_flush_tags = false, result_tags = false;
  uint64_t in_bytes = 0, n_mates = 0, seq_bytes = 0, rescue_bytes = 0;
  bool have_batch = false, have_result = false, have_scores = false, opt_generic_front = false, opt_no_fastpath = false, opt_ordered = false, result_ordered = false, opt_packed = false, result_packed = false, opt_merge_warp = false, opt_compact = false, result_compact = false; uint32_t compact_level = 0, result_compact_level = 0;
  // control block + pools (capacities persist across calls and only grow)
  Buf ctl;
  uint32_t cap_cand = 0, cap_ideal = 0, cap_rows = 0, cap_tmpw = 0, cap_outw = 0, cap_tb = 0;
  bool validated = false;           // the CSR offsets of the resident batch have been checked on the device
  Buf q_any, a_cand, scan_tmp;
  Buf c_rec, c_task, c_first, c_last, c_nseg, c_nguide, c_cap, ipool;
  Buf r_left, r_right, rpool;
  Buf c_mg, m_rec, m_sim, m_clip, c_stage, t_agg, t_claim, t_off, p_row, p_pool, p_sim, p_clip;
  Buf a_roff, a_rcnt, g_roff;
  Buf o_src, o_tid, o_pos, o_strand, o_flags, o_sim, o_clip, o_coff, tpool, opool;
  HBuf h_small, h_out;
  RescueBufs* rb = nullptr;
  uint32_t n_cand = 0, n_rec = 0; uint64_t n_out_words = 0;
  RowView rv{}; Buf res_slab;         // where the rows of the resident result are: the o_* buffers, or res_slab after brb_ctx_move_result
  uint32_t tune_calls = 0; float tune_ms[2] = {0.f, 0.f};      // long reads: front-kernel time per alignment with 8 / 1 alignments per ticket
  uint32_t res_n_aln = 0; bool res_foreign = false;      // alignments of the batch the resident result belongs to; result moved in from another ctx (brb_ctx_move_result)
  brb_stats stats{};
  std::vector<Buf*> all() {
    return {&in_group_off, &in_refid, &in_strand, &in_exon_off, &in_exon_st, &in_exon_en, &in_cigar_off, &in_cigar, &in_mate, &in_seq_off,
            &in_seq, &in_name_off, &in_names, &f_hi, &f_nh, &f_mapq, &f_prim, &z_in, &z_slots, &z_sizes, &z_crcs, &z_tot, &z_off, &z_out, &z_ticket, &e_raw, &e_off, &e_len, &e_units, &e_plan, &e_sizes, &e_offs, &e_err, &i_in, &i_meta, &i_out, &i_ctl, &x_cnt16, &x_cnt32, &x_md16, &x_r16, &x_pos, &x_err, &res_slab, &ctl, &q_any, &a_cand, &scan_tmp, &c_rec, &c_task, &c_first, &c_last, &c_nseg, &c_nguide, &c_cap, &ipool,
            &r_left, &r_right, &rpool, &c_mg, &m_rec, &m_sim, &m_clip, &c_stage, &t_agg, &t_claim, &t_off, &p_row, &p_pool, &p_sim, &p_clip, &a_roff, &a_rcnt, &g_roff,
            &o_src, &o_tid, &o_pos, &o_strand, &o_flags, &o_sim, &o_clip, &o_coff, &tpool, &opool};
  }
};

extern "C" {

int brb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
const char* brb_last_error(void) { return g_err.c_str(); }
const char* brb_version(void) { return "bramble_b200 0.2 (sm_100a)"; }

int brb_host_alloc(size_t bytes, void** out) {
  if (!out) return fail(BRB_E_ARG, "brb_host_alloc: null out");
  if (brb_device_count() <= 0) return fail(BRB_E_NO_DEVICE, "no CUDA device");
  CU(cudaMallocHost(out, bytes ? bytes : 1));
  return BRB_OK;
}
int brb_host_free(void* p) { if (p) cudaFreeHost(p); return BRB_OK; }

// ---- index: the g2t tree of the reference (src/bramble.cpp:132-211, src/g2t.cpp:34-62), flattened.
// Besides the reference's own columns the device records carry the exon chain of every transcript in
// genomic order (gprev/gnext) and each table gets a coarse bin directory.  Both are derived here; the
// derivation doubles as validation of what the reference's GTF loader guarantees (gclib/gff.cpp:988-1075:
// exons of a transcript are merged when they overlap or touch, hence disjoint and ascending) and of the
// neighbour columns (src/bramble.cpp:161-199).
int brb_index_create(int device, const brb_index_desc* d, brb_index** out) {
  if (!d || !out) return fail(BRB_E_ARG, "brb_index_create: null argument");
  if (brb_device_count() <= device || device < 0) return fail(BRB_E_NO_DEVICE, "no CUDA device %d", device);
  if (d->n_tables & 1) return fail(BRB_E_ARG, "n_tables must be 2 * n_refids");
  if (d->n_tables && !d->table_off) return fail(BRB_E_ARG, "table_off is null");
  const uint64_t n = d->n_tables ? d->table_off[d->n_tables] : 0;
  if (n >= 0xfffffff0ull) return fail(BRB_E_ARG, "too many index records");
  if (n && (!d->st || !d->en || !d->tid || !d->exon_id || !d->pos_start || !d->prev_st || !d->prev_en || !d->next_st ||
            !d->next_en || !d->has_prev_next || !d->transcript_len)) return fail(BRB_E_ARG, "null index column");
  if (d->genome_format > 1) return fail(BRB_E_ARG, "unknown genome_format %u", d->genome_format);
  std::vector<uint32_t> toff(d->n_tables + 1, 0);
  for (uint32_t t = 0; t < d->n_tables; t++) {
    toff[t] = (uint32_t)d->table_off[t];
    if (d->table_off[t + 1] < d->table_off[t]) return fail(BRB_E_ARG, "table_off not monotone");
  }
  toff[d->n_tables] = (uint32_t)n;
  CU(cudaSetDevice(device));
  brb_index* ix = new brb_index(); ix->device = device;
  // the host's share of the build ends here (the order of the records is the caller's: brb_sort_by_start); everything derived
  // from the sorted columns -- running max, exon chains, bin directory, records, validation -- is computed on the device (k_index.cuh)
  Buf c_en, c_tid, c_eid, c_pos, c_ps, c_pe, c_ns, c_ne, c_hpn, c_err, c_cnt, c_gnext, g_tmp;
  std::vector<Buf*> tmp = {&c_en, &c_tid, &c_eid, &c_pos, &c_ps, &c_pe, &c_ns, &c_ne, &c_hpn, &c_err, &c_cnt, &c_gnext, &g_tmp};
  int rc = 0;
  auto done = [&](int code) { for (Buf* b : tmp) b->release(); if (code) { ix->release_all(); delete ix; } return code; };
  auto up = [&](Buf& b, const void* src, size_t bytes, bool resident) -> int {
    int r = b.ensure(bytes ? bytes : 16); if (r) return r;
    if (bytes && cudaMemcpy(b.p, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) return fail(BRB_E_CUDA, "index upload failed");
    if (resident) ix->bytes += bytes;
    return 0;
  };
  if (!rc) rc = up(ix->table_off, toff.data(), toff.size() * 4, true);
  if (!rc) rc = up(ix->st, d->st, n * 4, true);
  if (!rc) rc = up(c_en, d->en, n * 4, false);
  if (!rc) rc = up(c_tid, d->tid, n * 4, false);
  if (!rc) rc = up(c_eid, d->exon_id, n, false);
  if (!rc) rc = up(c_pos, d->pos_start, n * 4, false);
  if (!rc) rc = up(c_ps, d->prev_st, n * 4, false);
  if (!rc) rc = up(c_pe, d->prev_en, n * 4, false);
  if (!rc) rc = up(c_ns, d->next_st, n * 4, false);
  if (!rc) rc = up(c_ne, d->next_en, n * 4, false);
  if (!rc) rc = up(c_hpn, d->has_prev_next, n, false);
  if (!rc) rc = ix->pmax.ensure(n * 4 + 16);
  if (!rc) rc = ix->gprev.ensure(n * 4 + 16);
  if (!rc) rc = c_gnext.ensure(n * 4 + 16);
  if (!rc) rc = ix->rec.ensure(n * sizeof(Rec) + 32);
  if (!rc) rc = c_err.ensure(sizeof(IndexErr));
  if (rc) return done(rc);
  ix->bytes += n * 4 * 2 + n * sizeof(Rec);
  IndexCols C; C.n = (uint32_t)n; C.n_tables = d->n_tables; C.table_off = ix->table_off.as<uint32_t>(); C.st = ix->st.as<int32_t>(); C.en = c_en.as<int32_t>();
  C.tid = c_tid.as<uint32_t>(); C.exon_id = c_eid.as<uint8_t>(); C.pos_start = c_pos.as<uint32_t>(); C.prev_st = c_ps.as<uint32_t>(); C.prev_en = c_pe.as<uint32_t>();
  C.next_st = c_ns.as<uint32_t>(); C.next_en = c_ne.as<uint32_t>(); C.hpn = c_hpn.as<uint8_t>();
  IndexErr* derr = c_err.as<IndexErr>(); IndexErr herr{};
  cudaMemset(derr, 0, sizeof(IndexErr));
  if (d->n_tables) k_index_pmax<<<d->n_tables, 32>>>(C, ix->pmax.as<int32_t>(), derr);
  if (n) k_index_links<<<grid_for(n, 256), 256>>>(C, ix->gprev.as<uint32_t>(), c_gnext.as<uint32_t>(), derr);
  if (cudaMemcpy(&herr, derr, sizeof herr, cudaMemcpyDeviceToHost) != cudaSuccess) return done(fail(BRB_E_CUDA, "index build failed: %s", cudaGetErrorString(cudaGetLastError())));
  if (herr.code == 1) return done(fail(BRB_E_ARG, "table %u is not sorted by start", herr.aux));
  if (herr.code == 2) return done(fail(BRB_E_ARG, "record %u: empty or negative interval", herr.rec));
  if (herr.code == 3) return done(fail(BRB_E_UNSUPPORTED, "transcript %u has overlapping exons (records %u, %u); the reference's GTF loader merges those (gclib/gff.cpp:988)", d->tid[herr.rec], herr.aux, herr.rec));
  if (herr.code == 4) return done(fail(BRB_E_ARG, "record %u: prev/next columns do not describe the neighbouring exons of transcript %u", herr.rec, herr.aux));
  uint32_t max_chain = 0;
  if (n) {
    const uint32_t n_tid = herr.max_tid + 1;
    rc = c_cnt.ensure((size_t)n_tid * 4); if (rc) return done(rc);
    cudaMemset(c_cnt.p, 0, (size_t)n_tid * 4);
    k_index_chain_count<<<grid_for(n, 256), 256>>>(C, ix->gprev.as<uint32_t>(), c_cnt.as<uint32_t>());
    k_index_chain_max<<<grid_for(n_tid, 256), 256>>>(n_tid, c_cnt.as<uint32_t>(), derr);
  }
  // bin directory: 2^shift bases per bin, about two bins per record overall
  const uint64_t span = herr.span;
  uint32_t shift = 4;
  { const uint64_t target = std::max<uint64_t>(2 * n, 1024); while (shift < 24 && (span >> shift) > target) shift++; }
  if (const char* e = getenv("BRB_BIN_SHIFT")) { int v = atoi(e); if (v >= 2 && v <= 28) shift = (uint32_t)v; }
  std::vector<int32_t> tmax(d->n_tables, 0);            // the last running maximum of every table fixes its number of bins
  for (uint32_t t = 0; t < d->n_tables; t++)
    if (toff[t + 1] > toff[t] && cudaMemcpy(&tmax[t], ix->pmax.as<int32_t>() + toff[t + 1] - 1, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return done(fail(BRB_E_CUDA, "index build failed"));
  std::vector<uint32_t> boff(d->n_tables + 1, 0);
  for (uint32_t t = 0; t < d->n_tables; t++) boff[t + 1] = boff[t] + (toff[t + 1] > toff[t] ? ((uint32_t)tmax[t] >> shift) + 2 : 0);
  rc = up(ix->bin_off, boff.data(), boff.size() * 4, true); if (rc) return done(rc);
  rc = ix->bins.ensure((size_t)boff[d->n_tables] * sizeof(uint2) + 16); if (rc) return done(rc);
  ix->bytes += (size_t)boff[d->n_tables] * sizeof(uint2);
  if (boff[d->n_tables]) k_index_bins<<<grid_for(boff[d->n_tables], 256), 256>>>(C, ix->pmax.as<int32_t>(), ix->bin_off.as<uint32_t>(), shift, ix->bins.as<uint2>());
  if (n) k_index_recs<<<grid_for(n, 256), 256>>>(C, ix->gprev.as<uint32_t>(), c_gnext.as<uint32_t>(), ix->rec.as<uint4>(), VAR_SPLIT_REC != 0);
  static_assert(sizeof(Rec) == 32, "a record is a 16-byte key half and a 16-byte link half");
  ix->d.n_tables = d->n_tables; ix->d.n_rec = n; ix->d.n_chrom = 0; ix->d.bin_shift = shift;
  // genome: resident as two 4-bit codes per byte, whichever way it comes in
  if (d->genome_codes && d->chrom_off && d->n_chrom) {
    const uint64_t nb = d->chrom_off[d->n_chrom];
    rc = up(ix->chrom_off, d->chrom_off, (size_t)(d->n_chrom + 1) * 8, true); if (rc) return done(rc);
    if (d->genome_format == 1) { rc = up(ix->genome, d->genome_codes, (nb + 1) / 2, true); if (rc) return done(rc); }
    else {
      rc = up(g_tmp, d->genome_codes, nb, false); if (rc) return done(rc);
      rc = ix->genome.ensure((nb + 1) / 2 + 16); if (rc) return done(rc);
      ix->bytes += (nb + 1) / 2;
      if (nb) k_pack_genome<<<grid_for((nb + 1) / 2, 256), 256>>>(g_tmp.as<uint8_t>(), nb, ix->genome.as<uint8_t>());
    }
    ix->d.n_chrom = d->n_chrom;
  }
  if (cudaMemcpy(&herr, derr, sizeof herr, cudaMemcpyDeviceToHost) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess)
    return done(fail(BRB_E_CUDA, "index build failed: %s", cudaGetErrorString(cudaGetLastError())));
  if (herr.code == 4) return done(fail(BRB_E_ARG, "transcript %u: the prev/next columns do not connect its exons into one chain", herr.aux));
  max_chain = herr.max_chain;
  ix->d.max_chain = max_chain;
  ix->d.table_off = ix->table_off.as<uint32_t>(); ix->d.st = ix->st.as<int32_t>(); ix->d.pmax = ix->pmax.as<int32_t>();
  ix->d.rkey = ix->rec.as<uint4>(); ix->d.rlink = ix->rec.as<uint4>() + (VAR_SPLIT_REC ? n : 0); ix->d.gprev = ix->gprev.as<uint32_t>();
  ix->d.bin_off = ix->bin_off.as<uint32_t>(); ix->d.bins = ix->bins.as<uint2>();
  ix->d.chrom_off = ix->d.n_chrom ? ix->chrom_off.as<uint64_t>() : nullptr;
  ix->d.genome = ix->d.n_chrom ? ix->genome.as<uint8_t>() : nullptr;
  done(0);
  *out = ix;
  return BRB_OK;
}

int brb_ctx_destroy(brb_ctx* c);

int brb_index_destroy(brb_index* ix) {
  if (!ix) return BRB_OK;
  cudaSetDevice(ix->device);
  ix->release_all();
  delete ix;
  return BRB_OK;
}

int brb_ctx_create(brb_index* ix, brb_ctx** out) {
  if (!ix || !out) return fail(BRB_E_ARG, "brb_ctx_create: null argument");
  CU(cudaSetDevice(ix->device));
  brb_ctx* c = new brb_ctx(); c->idx = ix;
  auto init = [&]() -> int {
    CU(cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking));
    for (auto& e : c->ev) CU(cudaEventCreate(&e));
    for (auto& e : c->tev) CU(cudaEventCreate(&e));
    ENS(c->ctl, sizeof(Ctl)); ENS(c->h_small, 1024);      // host images of Ctl: [0] initialisation, [512] result
    int sm = 0;
    CU(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, ix->device));
    c->n_sm = sm > 0 ? sm : 1;
    return BRB_OK;
  };
  const int rc = init();
  if (rc != BRB_OK) { brb_ctx_destroy(c); return rc; }     // (keeps the error text of the failing call)
  *out = c;
  return BRB_OK;
}

int brb_ctx_destroy(brb_ctx* c) {
  if (!c) return BRB_OK;
  cudaSetDevice(c->idx->device);
  if (c->st) cudaStreamSynchronize(c->st);
  for (Buf* b : c->all()) b->release();
  free_rescue_bufs(c);
  c->h_small.release(); c->h_out.release(); c->z_host.release();
  for (auto& e : c->ev) if (e) cudaEventDestroy(e);
  for (auto& e : c->tev) if (e) cudaEventDestroy(e);
  if (c->st) cudaStreamDestroy(c->st);
  delete c;
  return BRB_OK;
}

// What k_expand_all found out about a batch that came with compact columns -- read after a stream synchronisation the caller has done
// anyway (the public brb_batch_upload, or the end of a projection): the true exon total, and the two refusals.
static int finish_expand(brb_ctx* c) {
  if (!c->expand_pending) return BRB_OK;
  c->expand_pending = false;
  uint32_t h[4] = {0, 0, 0, 0};
  CU(cudaMemcpyAsync(h, c->x_err.p, 16, cudaMemcpyDeviceToHost, c->st));     // (the ctx stream is idle here: this returns at once)
  CU(cudaStreamSynchronize(c->st));
  if (c->expand_cigar && (h[3] || h[2] != c->expand_n_cigar)) { c->have_batch = false; return fail(BRB_E_ARG, "cigar_cnt16 sums to %u words, n_cigar_total says %u", h[2], c->expand_n_cigar); }
  if (h[0]) { c->have_batch = false; return fail(BRB_E_ARG, "invalid CIGAR: alignment %u yields no read exon", h[0] - 1); }
  if (c->expand_exons) c->b.n_exon = h[1];
  return BRB_OK;
}

// ---- batch upload: one bundle (reference BundleData::reads) as structure-of-arrays.  Content checks that
// need a pass over the data (symmetric mate links, ascending read exons) run on the device inside k_lookup.
static int batch_upload_async(brb_ctx* c, const brb_batch_in* in) {
  if (!c || !in) return fail(BRB_E_ARG, "brb_batch_upload: null argument");
  CU(cudaSetDevice(c->idx->device));
  c->have_batch = c->have_result = false; c->validated = false;
  const uint32_t n = in->n_aln, ng = in->n_groups;
  const bool cx_exon = in->pos0 != nullptr && !in->exon_off, cx_cig = in->cigar_cnt16 != nullptr && !in->cigar_off;
  const bool cx_mate = in->mate_delta16 != nullptr && !in->mate_idx, cx_ref = in->refid16 != nullptr && !in->refid;
  if (n && (!in->group_off || !(in->refid || cx_ref) || !in->strand || !(cx_exon || (in->exon_off && in->exon_st && in->exon_en)) ||
            !(in->cigar_off || cx_cig) || !in->cigar || !(in->mate_idx || cx_mate))) return fail(BRB_E_ARG, "null batch column");
  if (n >= 0x7ffffff0u) return fail(BRB_E_ARG, "batch too large");
  if (cx_cig && in->n_cigar_total >= 0xfffffff0ull) return fail(BRB_E_ARG, "batch too large");
  uint32_t n_exon = (n && !cx_exon) ? in->exon_off[n] : 0;
  const uint32_t n_cigar = !n ? 0 : cx_cig ? (uint32_t)in->n_cigar_total : in->cigar_off[n];
  if (n && (in->group_off[0] != 0 || in->group_off[ng] != n || (!cx_exon && in->exon_off[0] != 0) || (!cx_cig && in->cigar_off[0] != 0)))
    return fail(BRB_E_ARG, "inconsistent CSR offsets");
  c->in_bytes = 0;
  auto up = [&](Buf& b, const void* src, size_t bytes) -> int {
    int r = b.ensure(bytes ? bytes : 16); if (r) return r;
    if (bytes && cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->st) != cudaSuccess) return fail(BRB_E_CUDA, "batch upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    c->in_bytes += bytes; return 0;
  };
  int rc = 0;
  static const uint32_t zero_off[1] = {0};
  cudaStream_t st = c->st;
  if (!rc) rc = up(c->in_group_off, n ? in->group_off : zero_off, n ? (size_t)(ng + 1) * 4 : 4);
  if (!rc) rc = up(c->in_strand, in->strand, (size_t)n);
  if (!rc && !cx_ref) rc = up(c->in_refid, in->refid, (size_t)n * 4);
  if (!rc && !cx_mate) rc = up(c->in_mate, in->mate_idx, (size_t)n * 4);
  if (!rc && !cx_cig) rc = up(c->in_cigar_off, n ? in->cigar_off : zero_off, n ? (size_t)(n + 1) * 4 : 4);
  if (!rc) rc = up(c->in_cigar, in->cigar, (size_t)n_cigar * 4);
  if (!rc && !cx_exon) {
    rc = up(c->in_exon_off, n ? in->exon_off : zero_off, n ? (size_t)(n + 1) * 4 : 4);
    if (!rc) rc = up(c->in_exon_st, in->exon_st, (size_t)n_exon * 4);
    if (!rc) rc = up(c->in_exon_en, in->exon_en, (size_t)n_exon * 4);
  }
  if (!rc && cx_exon && !n) rc = up(c->in_exon_off, zero_off, 4);
  c->expand_pending = false;
  if (!rc && n && (cx_exon || cx_cig || cx_mate || cx_ref)) {  // the compact columns: uploaded as they are, expanded by ONE kernel (k_expand.cuh), no host round trip
    ExpandJob J{}; J.n = n; J.n_cigar = n_cigar;
    if (cx_ref) { rc = up(c->x_r16, in->refid16, (size_t)n * 2); if (!rc) rc = c->in_refid.ensure((size_t)n * 4); J.refid16 = c->x_r16.as<int16_t>(); J.refid = c->in_refid.as<int32_t>(); }
    if (!rc && cx_mate) { rc = up(c->x_md16, in->mate_delta16, (size_t)n * 2); if (!rc) rc = c->in_mate.ensure((size_t)n * 4); J.mate_delta16 = c->x_md16.as<int16_t>(); J.mate_idx = c->in_mate.as<int32_t>(); }
    if (!rc && cx_cig) { rc = up(c->x_cnt16, in->cigar_cnt16, (size_t)n * 2); if (!rc) rc = c->in_cigar_off.ensure((size_t)(n + 1) * 4); J.cnt16 = c->x_cnt16.as<uint16_t>(); }
    J.cigar_off = c->in_cigar_off.as<uint32_t>(); J.cigar = c->in_cigar.as<uint32_t>();
    if (!rc && cx_exon) {
      // room for the exons: an alignment has at most one exon per CIGAR word, plus one (the host learns the true total with the results)
      const uint64_t cap = (uint64_t)n_cigar + n;
      if (cap >= 0xfffffff0ull) return fail(BRB_E_ARG, "batch too large");
      rc = up(c->x_pos, in->pos0, (size_t)n * 4);
      if (!rc) rc = c->in_exon_off.ensure((size_t)(n + 1) * 4);
      if (!rc) rc = c->in_exon_st.ensure((size_t)cap * 4 + 16);
      if (!rc) rc = c->in_exon_en.ensure((size_t)cap * 4 + 16);
      J.pos0 = c->x_pos.as<int32_t>(); J.exon_off = c->in_exon_off.as<uint32_t>(); J.exon_st = c->in_exon_st.as<uint32_t>(); J.exon_en = c->in_exon_en.as<uint32_t>(); J.exon_cap = (uint32_t)cap;
      n_exon = (uint32_t)cap;                                  // (until the true total is known)
    }
    const uint32_t tiles = (n + 255) / 256;
    if (!rc) rc = c->x_cnt32.ensure((size_t)(2 * (size_t)tiles + 2) * 8);
    if (!rc) rc = c->x_err.ensure(16);
    if (!rc) {
      CU(cudaMemsetAsync(c->x_cnt32.p, 0, (size_t)(2 * (size_t)tiles + 2) * 8, st));
      CU(cudaMemsetAsync(c->x_err.p, 0, 16, st));
      J.look = c->x_cnt32.as<unsigned long long>(); J.res = c->x_err.as<uint32_t>();
      k_expand_all<<<std::min<uint32_t>(tiles, (uint32_t)c->n_sm * 8), 256, 0, st>>>(J);
      c->expand_pending = true; c->expand_exons = cx_exon; c->expand_cigar = cx_cig; c->expand_n_cigar = n_cigar;
    }
  }
  c->seq_bytes = 0;
  const bool have_seq = n && in->group_seq_off && in->seq_codes;
  if (!rc && have_seq) {
    rc = up(c->in_seq_off, in->group_seq_off, (size_t)(ng + 1) * 8);
    if (!rc) rc = up(c->in_seq, in->seq_codes, (size_t)in->group_seq_off[ng]);
    c->seq_bytes = in->group_seq_off[ng];
  }
  c->have_names = n && in->group_name_off && in->name_bytes;
  if (!rc && c->have_names) {
    rc = up(c->in_name_off, in->group_name_off, (size_t)(ng + 1) * 8);
    if (!rc) rc = up(c->in_names, in->name_bytes, (size_t)in->group_name_off[ng]);
  }
  if (rc) return rc;
  DevBatch& b = c->b;
  b.n_aln = n; b.n_groups = n ? ng : 0; b.n_exon = n_exon; b.n_cigar = n_cigar;
  b.group_off = c->in_group_off.as<uint32_t>(); b.refid = c->in_refid.as<int32_t>(); b.strand = c->in_strand.as<int8_t>();
  b.exon_off = c->in_exon_off.as<uint32_t>(); b.exon_st = c->in_exon_st.as<uint32_t>(); b.exon_en = c->in_exon_en.as<uint32_t>();
  b.cigar_off = c->in_cigar_off.as<uint32_t>(); b.cigar = c->in_cigar.as<uint32_t>(); b.mate_idx = c->in_mate.as<int32_t>();
  b.group_seq_off = have_seq ? c->in_seq_off.as<uint64_t>() : nullptr; b.seq = have_seq ? c->in_seq.as<uint8_t>() : nullptr;
  b.aln_group = nullptr; b.back_shift = 5; b.front_shift = 5; b.count_hits = 0; b.compact_rows = 0;
  c->stats.h2d_bytes = c->in_bytes;
  c->have_batch = true;
  return BRB_OK;
}

// The read exons of the resident batch as the engine holds them -- uploaded, or derived on the device from pos0 + CIGAR.
int brb_batch_read_exons(brb_ctx* c, uint32_t* exon_off, uint32_t* exon_st, uint32_t* exon_en, uint64_t cap, uint64_t* n_exon) {
  if (!c || !n_exon) return fail(BRB_E_ARG, "brb_batch_read_exons: null argument");
  if (!c->have_batch) return fail(BRB_E_ARG, "no batch resident: call brb_batch_upload first");
  CU(cudaSetDevice(c->idx->device));
  *n_exon = c->b.n_exon;
  if (c->b.n_exon > cap) return fail(BRB_E_ARG, "brb_batch_read_exons: %u exons, room for %llu", c->b.n_exon, (unsigned long long)cap);
  if (exon_off) CU(cudaMemcpyAsync(exon_off, c->in_exon_off.p, (size_t)(c->b.n_aln + 1) * 4, cudaMemcpyDeviceToHost, c->st));
  if (exon_st && c->b.n_exon) CU(cudaMemcpyAsync(exon_st, c->in_exon_st.p, (size_t)c->b.n_exon * 4, cudaMemcpyDeviceToHost, c->st));
  if (exon_en && c->b.n_exon) CU(cudaMemcpyAsync(exon_en, c->in_exon_en.p, (size_t)c->b.n_exon * 4, cudaMemcpyDeviceToHost, c->st));
  CU(cudaStreamSynchronize(c->st));
  return BRB_OK;
}

// Public form: on return the caller's columns have been read (they may be reused or freed).
int brb_batch_upload(brb_ctx* c, const brb_batch_in* in) {
  const int rc = batch_upload_async(c, in);
  if (rc) return rc;
  CU(cudaStreamSynchronize(c->st));
  CU(cudaGetLastError());
  return finish_expand(c);
}

}  // extern "C"

namespace {

int launch_rescue(brb_ctx* c, const brb_params& p, uint64_t* launches);   // clip-rescue phase, rescue_host.inl
struct RescueBufs;
bool grow_ksw_big(RescueBufs& R, const Ctl* h);
RescueBufs& rescue_bufs(brb_ctx* c);

float ms_between(cudaEvent_t a, cudaEvent_t b) { float ms = 0; cudaEventElapsedTime(&ms, a, b); return ms; }

uint32_t grown(uint64_t need) { const uint64_t w = need + need / 4 + 4096; return (uint32_t)std::min<uint64_t>(w, 0xfffffff0ull); }

// Sizes every pool for the current capacities (Buf::ensure keeps what is already large enough).
int ensure_pools(brb_ctx* c, bool rescue) {
  const DevBatch& b = c->b;
  const size_t n = b.n_aln, C = c->cap_cand, R = c->cap_rows;
  ENS(c->q_any, (size_t)2 * b.n_exon + 16); ENS(c->a_cand, n * 16);
  ENS(c->a_roff, (n + 1) * 4); ENS(c->a_rcnt, (n + 1) * 4); ENS(c->g_roff, (size_t)(b.n_groups + 1) * 4);
  const size_t nbat = (n + ((size_t)1 << b.back_shift) - 1) >> b.back_shift;
  ENS(c->c_stage, C * 16); ENS(c->t_agg, nbat * 8 + 8); ENS(c->t_claim, nbat * 8 + 8); ENS(c->t_off, (nbat + 1) * 8);
  ENS(c->scan_tmp, scan_scratch_words(std::max<size_t>(nbat, 1)) * 8);
  if (c->opt_ordered) { ENS(c->p_row, R * 16); ENS(c->p_pool, (size_t)c->cap_outw * 4); }
  if (c->opt_packed) ENS(c->p_row, R * 16);
  ENS(c->m_rec, C * 16);
  if (c->b.back_shift < 5) ENS(c->c_mg, C * 8);              // long reads: where k_merge_long left the merged CIGAR of every match
  if (c->have_scores) { ENS(c->m_sim, C * 8); ENS(c->m_clip, C * 4); ENS(c->o_sim, R * 8); ENS(c->o_clip, R * 4); if (c->opt_ordered) { ENS(c->p_sim, R * 8); ENS(c->p_clip, R * 4); } }
  if (rescue || c->b.back_shift < 5) { ENS(c->c_rec, C * 4); ENS(c->c_task, C * 4); ENS(c->c_first, C * 4); ENS(c->c_last, C * 4); ENS(c->c_nseg, C * 4); ENS(c->c_nguide, C * 4); ENS(c->c_cap, C * 4); }
  ENS(c->ipool, (size_t)c->cap_ideal * 4);
  ENS(c->o_src, R * 4); ENS(c->o_tid, R * 4); ENS(c->o_pos, R * 4); ENS(c->o_strand, R); ENS(c->o_flags, R);
  ENS(c->o_coff, (R + 1) * 4);
  ENS(c->tpool, (size_t)c->cap_tmpw * 4); ENS(c->opool, (size_t)c->cap_outw * 4);
  return 0;
}

int project_resident(brb_ctx* c, const brb_params& p) {
  if (!c->have_batch) return fail(BRB_E_ARG, "no batch resident: call brb_batch_upload first");
  const bool rescue = p.long_reads && p.use_fasta;
  if (rescue && !c->idx->d.genome) return fail(BRB_E_UNSUPPORTED, "use_fasta requested but the index holds no genome");
  if (rescue && c->b.n_aln && !c->b.seq) return fail(BRB_E_UNSUPPORTED, "use_fasta requested but the batch carries no read sequences");
  c->b.count_hits = c->opt_generic_front ? 1u : 0u;
  c->b.compact_rows = (c->opt_compact && c->opt_packed && !c->opt_ordered && !p.long_reads) ? c->compact_level : 0u;
  // Long reads: the front kernel takes 8 alignments per ticket, or 1 -- which is faster depends on the data (ONT-like bundles: 8, 0.09 vs
  // 0.23 ms; bundles whose work sits in a few reads of 200-isoform genes: 1, 0.45 vs 0.73 ms), and nothing in the batch says which.  So
  // each context measures: first call 8, second call 1, then the faster one per alignment, probed again every 64 calls.  The results do
  // not depend on the choice (BRB_FRONT_SHIFT pins it).
  static const int pin_shift = [] { const char* e = getenv("BRB_FRONT_SHIFT"); return e ? std::max(0, std::min(5, atoi(e))) : -1; }();
  uint32_t fshift = 5;
  if (p.long_reads) {
    const uint32_t k = c->tune_calls % 64;
    fshift = pin_shift >= 0 ? (uint32_t)pin_shift : k == 0 ? 3u : k == 1 ? 0u : (c->tune_ms[1] < c->tune_ms[0] ? 0u : 3u);
  }
  c->b.front_shift = fshift;
  c->b.back_shift = p.long_reads ? 3u : 5u;           // long reads: 8 alignments per k_back batch (rows of 10^2..10^4 CIGAR words)
  const DevBatch& b = c->b; const DevIndex& ix = c->idx->d;
  cudaStream_t st = c->st;
  brb_stats& S = c->stats;
  const uint32_t n = b.n_aln;
  c->have_result = false; c->rescue_bytes = 0; c->n_cand = 0; c->n_rec = 0; c->n_out_words = 0;
  c->have_scores = p.filter_by_similarity || rescue;     // otherwise similarity_score and clip_score are 0 for every record
  S.n_candidates = S.n_matches = S.n_rescue = S.n_rec = 0; S.n_launches = 0; S.alg_bytes = 0;
  S.ms_lookup = S.ms_chain = S.ms_rescue = S.ms_finalize = S.ms_pair = S.ms_merge = 0; S.kernel_ms = 0;
  if (n == 0) {
    ENS(c->g_roff, (size_t)(b.n_groups + 1) * 4); ENS(c->o_coff, 4);
    CU(cudaMemsetAsync(c->g_roff.p, 0, 4, st)); CU(cudaMemsetAsync(c->o_coff.p, 0, 4, st)); CU(cudaStreamSynchronize(st));
    c->have_result = true; c->res_n_aln = c->b.n_aln; c->res_foreign = false; c->rv = RowView{c->o_src.as<uint32_t>(), c->o_tid.as<uint32_t>(), c->o_pos.as<uint32_t>(), c->o_strand.as<int8_t>(), c->o_flags.as<uint8_t>(), c->have_scores ? c->o_sim.as<double>() : nullptr, c->have_scores ? c->o_clip.as<int32_t>() : nullptr, c->o_coff.as<uint32_t>(), c->opool.as<uint32_t>()}; c->result_ordered = c->opt_ordered; c->result_packed = c->opt_packed && !c->opt_ordered; c->result_compact = false; c->result_compact_level = 0; return BRB_OK;
  }
  if (!c->cap_cand) {           // first guess from the batch shape; a pool that overflows grows and the sequence replays
    c->cap_cand = grown((uint64_t)12 * n); c->cap_rows = grown((uint64_t)6 * n);
    const uint64_t slack = std::min<uint64_t>((n + 31) / 32, (uint64_t)c->n_sm * 8 * 8) * POOL_CHUNK;   // unused chunk tails of the persistent warps
    c->cap_cand = grown((uint64_t)12 * n + slack);
    c->cap_ideal = grown((uint64_t)4 * c->cap_cand + b.n_exon + slack); c->cap_tmpw = grown((uint64_t)8 * c->cap_rows + (p.long_reads ? 6 : 2) * (uint64_t)b.n_cigar + slack);   // long rows: merged CIGAR + the cooperative merge's scratch
    c->cap_outw = grown((uint64_t)3 * c->cap_rows + b.n_cigar);
    c->cap_tb = grown((uint64_t)64 * n);
    if (getenv("BRB_TEST_SMALL_POOLS")) c->cap_cand = c->cap_rows = c->cap_ideal = c->cap_tmpw = c->cap_outw = c->cap_tb = 64;   // tests: force the grow-and-replay path
  }
  const int n_sm = c->n_sm;
  const int ncta = n_sm * 8;
  auto resident = [&](const void* kern) {
    for (auto& e : c->resident_cache) if (e.first == kern) return e.second;
    int nb = 0; if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, 256, 0) != cudaSuccess || nb < 1) nb = 1;
    c->resident_cache.emplace_back(kern, n_sm * nb); return n_sm * nb;
  };
  // short-read presets: only exon 0 is looked up, candidates verify the later exons themselves (k_front_short)
  // long reads: lookup + candidate list (k_front<false>), then one thread per candidate for the chain walk (k_chain) and the ideal
  // CIGAR (k_build) -- a read of a 200-isoform, 300-exon gene would otherwise keep one warp busy for the whole kernel
  const bool split = rescue || p.long_reads;
  const bool short_front = !p.long_reads && !p.ignore_small_exons && !rescue && !c->opt_generic_front;
  const bool fast_front = short_front && !p.filter_by_similarity && !c->opt_no_fastpath;
  const void* front_fn = fast_front ? (const void*)k_front_fast
                       : short_front ? (p.filter_by_similarity ? (const void*)k_front_short<true> : (const void*)k_front_short<false>)
                                     : (split ? (const void*)k_front<false> : (const void*)k_front<true>);
  const int ncta_front = resident(front_fn);
  const bool ordered = c->opt_ordered, packed = c->opt_packed && !ordered;      // BRB_OPT_ORDERED_ROWS wins when both are set
  const bool lng = p.long_reads;
  const int ncta_back = lng ? resident(ordered ? (const void*)k_back<true, false, true> : packed ? (const void*)k_back<false, true, true> : (const void*)k_back<false, false, true>)
                            : resident(ordered ? (const void*)k_back<true, false, false> : packed ? (const void*)k_back<false, true, false> : (const void*)k_back<false, false, false>);
  Ctl* ctl = c->ctl.as<Ctl>();
  static_assert(sizeof(Ctl) <= 512, "two Ctl images in h_small");
  Ctl* hctl = c->h_small.as<Ctl>();                                               // initialisation image (host -> device)
  Ctl* hres = reinterpret_cast<Ctl*>(c->h_small.as<char>() + 512);                // result image (written by the device)
  uint64_t L = 0;
  uint32_t n_syncs = 0;                               // host round trips of the call (one per sequence; more only when a pool had to grow)
  bool one_front = false, back_finishes = false;      // which phase events this sequence recorded
  for (int attempt = 0;; attempt++) {
    if (attempt > 12) return fail(BRB_E_NOMEM, "pool sizing did not converge");
    { int r = ensure_pools(c, rescue); if (r) return r; }
    const uint32_t caps[N_CAPS] = {c->cap_cand, c->cap_ideal, c->cap_rows, c->cap_tmpw, c->cap_outw, c->cap_tb};
    if (!c->ctl_clean || memcmp(caps, c->ctl_caps, sizeof(caps)) != 0) {          // steady state: the previous sequence left Ctl ready
      memset(hctl, 0, sizeof(Ctl));
      hctl->cap_cand = c->cap_cand; hctl->cap_ideal = c->cap_ideal; hctl->cap_rows = c->cap_rows; hctl->cap_tmpw = c->cap_tmpw; hctl->cap_outw = c->cap_outw;
      hctl->cap_tb = c->cap_tb;
      CU(cudaMemcpyAsync(ctl, hctl, sizeof(Ctl), cudaMemcpyHostToDevice, st));
      memcpy(c->ctl_caps, caps, sizeof(caps));
    }
    c->ctl_clean = false;
    Matches M; M.rec = c->m_rec.as<uint4>();
    M.sim = c->have_scores ? c->m_sim.as<double>() : nullptr; M.clip = c->have_scores ? c->m_clip.as<int32_t>() : nullptr;
    if (!c->validated) {          // CSR offsets of the batch: once per upload, before any kernel trusts them
      k_validate_batch<<<grid_for(std::max<size_t>(std::max<size_t>(n, b.n_groups) + 1, 1), 256), 256, 0, st>>>(b, ctl); L++;
      c->validated = true;
    }
    // ---- phase 1: lookup + candidate lists
    CU(cudaEventRecord(c->ev[0], st));
    one_front = false; back_finishes = false;
    // ---- phases 1..4: lookup, candidates, chain walk, (clip rescue), ideal CIGAR + scores
    if (short_front) {
      if (fast_front) k_front_fast<<<ncta_front, 256, 0, st>>>(ix, b, p, ctl, c->a_cand.as<uint4>(), c->ipool.as<uint32_t>(), M);
      else if (p.filter_by_similarity) k_front_short<true><<<ncta_front, 256, 0, st>>>(ix, b, p, ctl, c->a_cand.as<uint4>(), c->ipool.as<uint32_t>(), M);
      else k_front_short<false><<<ncta_front, 256, 0, st>>>(ix, b, p, ctl, c->a_cand.as<uint4>(), c->ipool.as<uint32_t>(), M);
      L++;
      CU(cudaEventRecord(c->ev[1], st)); one_front = true;
    } else if (!split) {
      k_front<true><<<ncta_front, 256, 0, st>>>(ix, b, p, ctl, c->q_any.as<uint8_t>(), c->a_cand.as<uint4>(), nullptr, nullptr,
                                          c->ipool.as<uint32_t>(), M); L++;
      CU(cudaEventRecord(c->ev[1], st)); one_front = true;
    } else {
      k_front<false><<<ncta_front, 256, 0, st>>>(ix, b, p, ctl, c->q_any.as<uint8_t>(), c->a_cand.as<uint4>(), c->c_rec.as<uint32_t>(), c->c_task.as<uint32_t>(),
                                           nullptr, M); L++;
      CU(cudaEventRecord(c->ev[1], st));
      CandFacts F; F.first = c->c_first.as<uint32_t>(); F.last = c->c_last.as<uint32_t>(); F.nseg = c->c_nseg.as<uint32_t>();
      F.nguide = c->c_nguide.as<uint32_t>(); F.cap = c->c_cap.as<uint32_t>();
      k_chain<<<ncta, 256, 0, st>>>(ix, b, p, ctl, c->q_any.as<uint8_t>(), c->c_rec.as<uint32_t>(), c->c_task.as<uint32_t>(), F); L++;
      CU(cudaEventRecord(c->ev[2], st));
      if (rescue) { int r = launch_rescue(c, p, &L); if (r) return r; }      // plan -> size-class DP kernels -> clip segments: no host round trip
      CU(cudaEventRecord(c->ev[3], st));
      k_build<<<ncta, 256, 0, st>>>(ix, b, p, ctl, c->q_any.as<uint8_t>(), c->c_task.as<uint32_t>(), F, rescue ? c->r_left.as<RescueSeg>() : nullptr,
                                    rescue ? c->r_right.as<RescueSeg>() : nullptr, c->rpool.as<uint32_t>(), c->ipool.as<uint32_t>(), M); L++;
      CU(cudaEventRecord(c->ev[4], st));
    }
    // ---- phase 5: pairing, CIGAR merge, rows (one kernel; row and CIGAR room claimed per batch in completion order)
    BackIn pv; pv.a_cand = c->a_cand.as<uint4>(); pv.m_rec = c->m_rec.as<uint4>(); pv.m_sim = M.sim; pv.m_clip = M.clip; pv.ipool = c->ipool.as<uint32_t>();
    pv.mg = c->c_mg.as<uint2>();
    Rows o; o.src = c->o_src.as<uint32_t>(); o.tid = c->o_tid.as<uint32_t>(); o.pos = c->o_pos.as<uint32_t>(); o.strand = c->o_strand.as<int8_t>();
    o.flags = c->o_flags.as<uint8_t>(); o.sim = c->o_sim.as<double>(); o.clip = c->o_clip.as<int32_t>(); o.coff = c->o_coff.as<uint32_t>();
    o.pool = c->opool.as<uint32_t>(); o.a_roff = c->a_roff.as<uint32_t>(); o.a_rcnt = c->a_rcnt.as<uint32_t>();
    if (split) { k_count_valid<<<grid_for(n, 256), 256, 0, st>>>(b, ctl, c->a_cand.as<uint4>(), c->m_rec.as<uint4>()); L++; }
    if (lng) { k_merge_long<<<resident((const void*)k_merge_long), 256, 0, st>>>(b, ctl, c->c_task.as<uint32_t>(), pv, c->tpool.as<uint32_t>(), c->c_mg.as<uint2>()); L++; }
    const uint32_t nbat = (uint32_t)((n + (1u << b.back_shift) - 1) >> b.back_shift);
    BatchTotals tl; tl.agg = c->t_agg.as<unsigned long long>(); tl.claim = c->t_claim.as<unsigned long long>();
    ProvRows pr; pr.row = c->p_row.as<uint4>(); pr.pool = c->p_pool.as<uint32_t>();
    pr.sim = c->have_scores ? c->p_sim.as<double>() : nullptr; pr.clip = c->have_scores ? c->p_clip.as<int32_t>() : nullptr;
    if (ordered) {
      if (lng) k_back<true, false, true><<<ncta_back, 256, 0, st>>>(b, ctl, pv, c->c_stage.as<uint4>(), c->tpool.as<uint32_t>(), tl, pr, o, hres);
      else k_back<true, false, false><<<ncta_back, 256, 0, st>>>(b, ctl, pv, c->c_stage.as<uint4>(), c->tpool.as<uint32_t>(), tl, pr, o, hres);
      L++;
      CU(cudaEventRecord(c->ev[5], st));
      // ---- phase 6: batch totals -> offsets in bundle order; rows to their final place
      exclusive_scan(tl.agg, c->t_off.as<unsigned long long>(), nbat, c->scan_tmp.as<unsigned long long>(), st, &L);
      k_place<<<grid_for(nbat, 8), 256, 0, st>>>(b, ctl, tl, c->t_off.as<unsigned long long>(), pr, o); L++;   // one warp per batch
      k_group_rows<<<grid_for(b.n_groups + 1, 256), 256, 0, st>>>(b, ctl, c->a_roff.as<uint32_t>(), c->g_roff.as<uint32_t>()); L++;
    } else {
      if (packed && lng) k_back<false, true, true><<<ncta_back, 256, 0, st>>>(b, ctl, pv, c->c_stage.as<uint4>(), c->tpool.as<uint32_t>(), tl, pr, o, hres);
      else if (packed) k_back<false, true, false><<<ncta_back, 256, 0, st>>>(b, ctl, pv, c->c_stage.as<uint4>(), c->tpool.as<uint32_t>(), tl, pr, o, hres);
      else if (lng) k_back<false, false, true><<<ncta_back, 256, 0, st>>>(b, ctl, pv, c->c_stage.as<uint4>(), c->tpool.as<uint32_t>(), tl, pr, o, hres);
      else k_back<false, false, false><<<ncta_back, 256, 0, st>>>(b, ctl, pv, c->c_stage.as<uint4>(), c->tpool.as<uint32_t>(), tl, pr, o, hres);
      L++;
      CU(cudaEventRecord(c->ev[5], st)); back_finishes = true;       // the last CTA to leave publishes the totals
    }
    if (!back_finishes) { k_publish<<<1, 1, 0, st>>>(ctl, hres); L++; CU(cudaEventRecord(c->ev[6], st)); }
    CU(cudaStreamSynchronize(st));                          // the counters are in hres (mapped pinned memory) once the sequence has drained
    n_syncs++;
    CU(cudaGetLastError());
    c->ctl_clean = true;
    { const int r = finish_expand(c); if (r) return r; }    // a batch that came with compact columns: its exon total, or its refusal
    if (hres->bad_input) break;
    bool again = false;
    if (hres->n_cand > c->cap_cand) { c->cap_cand = grown((uint64_t)hres->n_cand + std::min<uint64_t>((n + 31) / 32, (uint64_t)c->n_sm * 8 * 8) * POOL_CHUNK); c->cap_ideal = std::max(c->cap_ideal, grown((uint64_t)4 * c->cap_cand)); again = true; }
    else if (hres->n_tb > c->cap_tb) { c->cap_tb = grown(hres->n_tb); again = true; }
    else if (rescue && grow_ksw_big(rescue_bufs(c), hres)) again = true;
    else if (hres->n_ideal > c->cap_ideal) { c->cap_ideal = grown(hres->n_ideal); again = true; }
    else if (hres->n_rows > c->cap_rows) { c->cap_rows = grown(hres->n_rows); c->cap_tmpw = std::max(c->cap_tmpw, grown((uint64_t)8 * c->cap_rows));
                                           c->cap_outw = std::max(c->cap_outw, grown((uint64_t)3 * c->cap_rows)); again = true; }
    else if (hres->n_tmpw > c->cap_tmpw) { c->cap_tmpw = grown(hres->n_tmpw); again = true; }
    else if (hres->n_outw > c->cap_outw) { c->cap_outw = grown(hres->n_outw); again = true; }
    if (!again) break;
  }
  c->result_tags = false;
  if (c->opt_flush_tags && !hres->bad_input) {            // per-name HI / NH / MAPQ / primary on the finished rows
    if (packed) return fail(BRB_E_UNSUPPORTED, "BRB_OPT_FLUSH_TAGS is not available together with BRB_OPT_PACKED_ROWS");
    if (!c->have_names) return fail(BRB_E_ARG, "BRB_OPT_FLUSH_TAGS needs brb_batch_in::group_name_off / name_bytes");
    const size_t R = std::max<uint32_t>(hres->n_rows, 1);
    ENS(c->f_hi, R * 4); ENS(c->f_nh, R * 4); ENS(c->f_mapq, R); ENS(c->f_prim, R);
    FlushIn fi; fi.a_roff = c->a_roff.as<uint32_t>(); fi.a_rcnt = c->a_rcnt.as<uint32_t>(); fi.flags = c->o_flags.as<uint8_t>();
    fi.sim = c->have_scores ? c->o_sim.as<double>() : nullptr; fi.name_off = c->in_name_off.as<uint64_t>(); fi.names = c->in_names.as<uint8_t>();
    fi.cap_rows = c->cap_rows;
    FlushOut fo; fo.hi = c->f_hi.as<int32_t>(); fo.nh = c->f_nh.as<int32_t>(); fo.mapq = c->f_mapq.as<uint8_t>(); fo.primary = c->f_prim.as<uint8_t>();
    CU(cudaEventRecord(c->ev[7], st));
    k_flush_tags<<<grid_for(b.n_groups, 128), 128, 0, st>>>(b, fi, fo, p.long_reads != 0); L++;
    CU(cudaStreamSynchronize(st)); n_syncs++;
    CU(cudaGetLastError());
    c->result_tags = true;
  }
  if (hres->bad_input == 3) return fail(BRB_E_UNSUPPORTED, "a batch of 32 alignments emits 2^28 or more CIGAR words");
  if (hres->bad_input == 7) return fail(BRB_E_UNSUPPORTED, "BRB_OPT_PACKED_ROWS = 3: a row does not fit 8 bytes (tid >= 2^22, pos >= 2^24, or an M length / CIGAR word count >= 2^13); project the bundle with the option at 2 or less");
  if (hres->bad_input == 6) return fail(BRB_E_UNSUPPORTED, "BRB_OPT_PACKED_ROWS = 2: a unit of 65536 rows or more; project the bundle with the option at 0 or 1");
  if (hres->bad_input == 4) return fail(BRB_E_UNSUPPORTED, "BRB_OPT_PACKED_ROWS: the bundle pools 2^27 or more CIGAR words; project it without the option");
  if (hres->bad_input == 5) return fail(BRB_E_ARG, "CSR offsets (group_off / exon_off / cigar_off) must start at 0, be non-decreasing and end at the column sizes");
  if (hres->bad_input == 1) return fail(BRB_E_ARG, "mate_idx is not a symmetric single link (src/bramble.cpp:272-311)");
  if (hres->bad_input) return fail(BRB_E_ARG, "read exons of an alignment must be ascending and disjoint");
  c->n_cand = hres->n_cand; c->n_rec = hres->n_rows; c->n_out_words = hres->n_outw;
  S.n_candidates = hres->n_cand; S.n_rec = hres->n_rows; S.n_launches = L; S.n_matches = hres->n_matches; S.n_rescue = hres->n_rescue;
  const cudaEvent_t ev_front_end = one_front ? c->ev[1] : c->ev[4], ev_end = back_finishes ? c->ev[5] : c->ev[6];
  S.ms_lookup = ms_between(c->ev[0], c->ev[1]);
  if (p.long_reads && n) { c->tune_ms[c->b.front_shift == 0 ? 1 : 0] = S.ms_lookup / (float)n; c->tune_calls++; }
  if (!one_front) { S.ms_chain = ms_between(c->ev[1], c->ev[2]); S.ms_rescue = ms_between(c->ev[2], c->ev[3]); S.ms_finalize = ms_between(c->ev[3], c->ev[4]); }
  S.ms_pair = ms_between(ev_front_end, c->ev[5]); S.ms_merge = back_finishes ? 0.f : ms_between(c->ev[5], c->ev[6]);
  S.kernel_ms = ms_between(c->ev[0], ev_end);
  // algorithmic bytes, SURVEY.md 8(d) / DESIGN.md: per-alignment SoA + 56 B per index hit + output rows + rescue bases
  S.alg_bytes_front = (uint64_t)n * 16 + (uint64_t)b.n_exon * 8 + hres->n_mates * 4 + hres->n_hits * 56;
  S.alg_bytes = S.alg_bytes_front + (uint64_t)b.n_cigar * 4 + (uint64_t)hres->n_rows * 40 + hres->out_words * 4 + hres->rescue_bytes;
  S.ksw_cells = hres->ksw_cells; S.ksw_cells_run = hres->ksw_cells_run; S.host_syncs = n_syncs;
  c->have_result = true; c->res_n_aln = c->b.n_aln; c->res_foreign = false; c->rv = RowView{c->o_src.as<uint32_t>(), c->o_tid.as<uint32_t>(), c->o_pos.as<uint32_t>(), c->o_strand.as<int8_t>(), c->o_flags.as<uint8_t>(), c->have_scores ? c->o_sim.as<double>() : nullptr, c->have_scores ? c->o_clip.as<int32_t>() : nullptr, c->o_coff.as<uint32_t>(), c->opool.as<uint32_t>()}; c->result_ordered = ordered; c->result_packed = packed; c->result_compact = packed && c->b.compact_rows; c->result_compact_level = c->result_compact ? c->b.compact_rows : 0u;
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
  if (c->res_foreign) return fail(BRB_E_ARG, "the resident result was moved in from another ctx (brb_ctx_move_result): only brb_emit_bgzf reads it");
  CU(cudaSetDevice(c->idx->device));
  const size_t R = c->n_rec, W = c->n_out_words, G = c->b.n_groups;
  const bool sc = c->have_scores && R;
  auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };
  size_t off = 0;
  const bool ord = c->result_ordered, pk = c->result_packed, cm = c->result_compact;
  const size_t N = c->b.n_aln, NB = (N + 31) >> 5;
  const size_t o_groff = off; off += ord ? al((G + 1) * 4) : 0;
  const size_t o_aoff = off; off += al(cm ? NB * 4 : N * 4);              // compact: one row offset per batch of 32 alignments ...
  const size_t o_acnt = off; off += al(cm ? N * 2 : N * 4);               // ... and 16-bit unit sizes
  const size_t row_bytes = !cm ? 16 : c->result_compact_level == 2 ? 8 : 12;
  const size_t o_prow = off; off += pk ? al(R * row_bytes) : 0;
  const size_t o_src = off; off += pk ? 0 : al(R * 4);
  const size_t o_tid = off; off += pk ? 0 : al(R * 4);
  const size_t o_pos = off; off += pk ? 0 : al(R * 4);
  const size_t o_sim = off; off += sc ? al(R * 8) : 0;
  const size_t o_clip = off; off += sc ? al(R * 4) : 0;
  const size_t o_coff = off; off += pk ? 0 : al((R + 1) * 4);
  const size_t o_strand = off; off += pk ? 0 : al(R);
  const size_t o_flags = off; off += pk ? 0 : al(R);
  const size_t o_cig = off; off += al(W * 4);
  const bool tg = c->result_tags && R;
  const size_t o_hi = off; off += tg ? al(R * 4) : 0;
  const size_t o_nh = off; off += tg ? al(R * 4) : 0;
  const size_t o_mq = off; off += tg ? al(R) : 0;
  const size_t o_pr = off; off += tg ? al(R) : 0;
  ENS(c->h_out, off + 16);
  uint8_t* h = c->h_out.as<uint8_t>();
  uint64_t bytes = 0;
  auto dn = [&](size_t dst, const Buf& b, size_t n) -> int {
    if (!n) return 0;
    if (cudaMemcpyAsync(h + dst, b.p, n, cudaMemcpyDeviceToHost, c->st) != cudaSuccess) return fail(BRB_E_CUDA, "download failed: %s", cudaGetErrorString(cudaGetLastError()));
    bytes += n; return 0;
  };
  int rc = ord ? dn(o_groff, c->g_roff, (G + 1) * 4) : 0;
  if (!rc) rc = dn(o_aoff, c->a_roff, cm ? NB * 4 : N * 4);
  if (!rc) rc = dn(o_acnt, c->a_rcnt, cm ? N * 2 : N * 4);
  if (!rc && pk) rc = dn(o_prow, c->p_row, R * row_bytes);
  if (!rc && !pk) rc = dn(o_src, c->o_src, R * 4);
  if (!rc && !pk) rc = dn(o_tid, c->o_tid, R * 4);
  if (!rc && !pk) rc = dn(o_pos, c->o_pos, R * 4);
  if (!rc && sc) rc = dn(o_sim, c->o_sim, R * 8);
  if (!rc && sc) rc = dn(o_clip, c->o_clip, R * 4);
  if (!rc && !pk) rc = dn(o_coff, c->o_coff, (R + 1) * 4);
  if (!rc && !pk) rc = dn(o_strand, c->o_strand, R);
  if (!rc && !pk) rc = dn(o_flags, c->o_flags, R);
  if (!rc) rc = dn(o_cig, c->opool, W * 4);
  if (!rc && tg) rc = dn(o_hi, c->f_hi, R * 4);
  if (!rc && tg) rc = dn(o_nh, c->f_nh, R * 4);
  if (!rc && tg) rc = dn(o_mq, c->f_mapq, R);
  if (!rc && tg) rc = dn(o_pr, c->f_prim, R);
  if (rc) return rc;
  CU(cudaStreamSynchronize(c->st));
  out->n_rec = (uint32_t)R; out->n_cigar_words = W;
  out->group_rec_off = ord ? (const uint32_t*)(h + o_groff) : nullptr;
  out->aln_rec_off = cm ? nullptr : (const uint32_t*)(h + o_aoff); out->aln_rec_cnt = cm ? nullptr : (const uint32_t*)(h + o_acnt);
  out->batch_rec_off = cm ? (const uint32_t*)(h + o_aoff) : nullptr; out->aln_rec_cnt16 = cm ? (const uint16_t*)(h + o_acnt) : nullptr;
  out->packed_rows = pk ? (const uint32_t*)(h + o_prow) : nullptr; out->packed_row_words = pk ? (uint32_t)(row_bytes / 4) : 0u;
  out->src_aln = pk ? nullptr : (const uint32_t*)(h + o_src); out->tid = pk ? nullptr : (const uint32_t*)(h + o_tid);
  out->pos = pk ? nullptr : (const uint32_t*)(h + o_pos); out->tx_strand = pk ? nullptr : (const int8_t*)(h + o_strand); out->flags = pk ? nullptr : h + o_flags;
  out->sim_score = sc ? (const double*)(h + o_sim) : nullptr; out->clip_score = sc ? (const int32_t*)(h + o_clip) : nullptr;
  out->cigar_off = pk ? nullptr : (const uint32_t*)(h + o_coff); out->cigar = (const uint32_t*)(h + o_cig);
  out->hi = tg ? (const int32_t*)(h + o_hi) : nullptr; out->nh = tg ? (const int32_t*)(h + o_nh) : nullptr;
  out->mapq = tg ? h + o_mq : nullptr; out->primary = tg ? h + o_pr : nullptr;
  c->stats.d2h_bytes = bytes;
  return BRB_OK;
}

int brb_project(brb_ctx* c, const brb_params* p, const brb_batch_in* in, brb_batch_out* out) {
  static const bool timing = getenv("BRB_PROJECT_TIMING") != nullptr;       // host milliseconds of the three stages on stderr, per call
  auto now_ms = [] { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec / 1e6; };
  const double t0 = timing ? now_ms() : 0;
  int rc = batch_upload_async(c, in); if (rc) return rc;        // (the copies are ordered before the kernels on the ctx stream; the call returns after the download)
  const double t1 = timing ? now_ms() : 0;
  rc = brb_project_resident(c, p); if (rc) return rc;
  const double t2 = timing ? now_ms() : 0;
  rc = brb_batch_download(c, out);
  if (timing) fprintf(stderr, "[project %p] start %.3f upload %.3f project %.3f (kernels %.3f) download %.3f ms\n", (void*)c, t0, t1 - t0, t2 - t1, c->stats.kernel_ms, now_ms() - t2);
  return rc;
}

int brb_ctx_set_option(brb_ctx* c, int option, int value) {
  if (!c) return fail(BRB_E_ARG, "null ctx");
  if (option == BRB_OPT_GENERIC_FRONT) { c->opt_generic_front = value != 0; return BRB_OK; }
  if (option == BRB_OPT_NO_FASTPATH) { c->opt_no_fastpath = value != 0; return BRB_OK; }
  if (option == BRB_OPT_ORDERED_ROWS) { c->opt_ordered = value != 0; return BRB_OK; }
  if (option == BRB_OPT_PACKED_ROWS) { c->opt_packed = value != 0; c->opt_compact = value == 2 || value == 3; c->compact_level = value == 3 ? 2u : value == 2 ? 1u : 0u; return BRB_OK; }
  if (option == BRB_OPT_MERGE_WARP) { c->opt_merge_warp = value != 0; return BRB_OK; }
  if (option == BRB_OPT_FLUSH_TAGS) { c->opt_flush_tags = value != 0; return BRB_OK; }
  return fail(BRB_E_ARG, "unknown option %d", option);
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

// ---- BGZF encoder (k_deflate.cuh): n bytes -> complete BGZF blocks (0xff00 input bytes each, the last one shorter), on the ctx stream.
}  // extern "C"
namespace {
// BGZF blocks of the n bytes at z_in (device; 16 readable bytes behind them): kernels + the copy of the blocks to pinned host memory
int deflate_resident(brb_ctx* c, uint64_t n, const uint8_t** out, uint64_t* out_bytes) {
  const uint64_t nb64 = (n + BGZF_IN - 1) / BGZF_IN;
  if (nb64 > 0x3fffffffull) return fail(BRB_E_ARG, "BGZF encoder: input too large for one call");
  const uint32_t nb = (uint32_t)nb64;
  ENS(c->z_slots, (size_t)nb * BGZF_SLOT); ENS(c->z_sizes, (size_t)nb * 4); ENS(c->z_crcs, (size_t)nb * 4);
  ENS(c->z_tot, (size_t)nb * 8 + 8); ENS(c->z_off, (size_t)(nb + 1) * 8); ENS(c->z_ticket, 16);
  ENS(c->scan_tmp, scan_scratch_words(nb) * 8);
  ENS(c->z_out, n + (size_t)nb * 64 + 64);                       // (a stored block adds 5 + 26 bytes to its input)
  cudaStream_t st = c->st;
  CU(cudaMemsetAsync(c->z_ticket.p, 0, 16, st));
  DeflateJob J; J.in = c->z_in.as<uint8_t>(); J.n = n; J.slots = c->z_slots.as<uint8_t>(); J.sizes = c->z_sizes.as<uint32_t>(); J.crcs = c->z_crcs.as<uint32_t>();
  J.n_blocks = nb; J.ticket = c->z_ticket.as<uint32_t>();
  const uint32_t ctas = std::min<uint32_t>((nb + DEFLATE_WARPS - 1) / DEFLATE_WARPS, (uint32_t)c->n_sm * 8);
  uint64_t L = 0;
  k_bgzf_deflate<<<ctas, 32 * DEFLATE_WARPS, 0, st>>>(J);
  k_bgzf_sizes<<<grid_for(nb, 256), 256, 0, st>>>(J, c->z_tot.as<unsigned long long>());
  exclusive_scan(c->z_tot.as<unsigned long long>(), c->z_off.as<unsigned long long>(), nb, c->scan_tmp.as<unsigned long long>(), st, &L);
  k_bgzf_pack<<<nb, 256, 0, st>>>(J, c->z_off.as<unsigned long long>(), c->z_out.as<uint8_t>());
  unsigned long long total = 0;
  CU(cudaMemcpyAsync(&total, c->z_off.as<unsigned long long>() + nb, 8, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  ENS(c->z_host, total + 64);                                    // pinned: sized by what the blocks came to, not by the worst case
  CU(cudaMemcpyAsync(c->z_host.p, c->z_out.p, total, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  *out = c->z_host.as<uint8_t>(); *out_bytes = total;
  return BRB_OK;
}
}  // namespace
extern "C" {

int brb_bgzf_compress(brb_ctx* c, const uint8_t* data, uint64_t n, const uint8_t** out, uint64_t* out_bytes) {
  if (!c || !out || !out_bytes || (n && !data)) return fail(BRB_E_ARG, "brb_bgzf_compress: null argument");
  if (brb_device_count() <= 0) return fail(BRB_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(c->idx->device));
  *out = nullptr; *out_bytes = 0;
  if (!n) return BRB_OK;
  ENS(c->z_in, n + 16);
  CU(cudaMemsetAsync(c->z_in.as<uint8_t>() + (n & ~(uint64_t)3), 0, 16 - (n & 3), c->st));     // the word loads may read up to 7 bytes past the end
  CU(cudaMemcpyAsync(c->z_in.p, data, n, cudaMemcpyHostToDevice, c->st));
  return deflate_resident(c, n, out, out_bytes);
}

// BGZF blocks in, their inflated bytes out (k_inflate.cuh).  The host only walks the block headers (BC subfield -> block size, footer ->
// ISIZE) to place every block; Huffman decoding, LZ77 copies and the CRC-32 / ISIZE verification run on the device.
int brb_bgzf_inflate(brb_ctx* c, const uint8_t* data, uint64_t n, const uint8_t** out, uint64_t* out_bytes) {
  if (!c || !out || !out_bytes || (n && !data)) return fail(BRB_E_ARG, "brb_bgzf_inflate: null argument");
  if (brb_device_count() <= 0) return fail(BRB_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(c->idx->device));
  *out = nullptr; *out_bytes = 0;
  if (!n) return BRB_OK;
  // ---- block directory
  uint64_t nb = 0, total = 0;
  for (uint64_t p = 0; p < n; nb++) {
    if (n - p < 28 || data[p] != 0x1f || data[p + 1] != 0x8b || data[p + 2] != 8 || !(data[p + 3] & 4)) return fail(BRB_E_ARG, "brb_bgzf_inflate: not a BGZF block at byte %llu", (unsigned long long)p);
    const uint32_t xlen = data[p + 10] | (data[p + 11] << 8);
    if (n - p < 12ull + xlen + 8) return fail(BRB_E_ARG, "brb_bgzf_inflate: truncated block at byte %llu", (unsigned long long)p);
    uint32_t bsize = 0; bool found = false;
    for (uint32_t q = 0; q + 4 <= xlen;) {
      const uint8_t* x = data + p + 12 + q; const uint32_t slen = x[2] | (x[3] << 8);
      if (x[0] == 'B' && x[1] == 'C' && slen == 2 && q + 6 <= xlen) { bsize = x[4] | (x[5] << 8); found = true; }
      q += 4 + slen;
    }
    if (!found || (uint64_t)bsize + 1 < 12ull + xlen + 8 || n - p < (uint64_t)bsize + 1) return fail(BRB_E_ARG, "brb_bgzf_inflate: bad or truncated block at byte %llu", (unsigned long long)p);
    p += (uint64_t)bsize + 1;
  }
  if (nb > 0x3fffffffull) return fail(BRB_E_ARG, "brb_bgzf_inflate: too many blocks for one call");
  ENS(c->i_hmeta, nb * 20 + 64);
  uint64_t* h_in = c->i_hmeta.as<uint64_t>(); uint64_t* h_out = h_in + nb; uint32_t* h_clen = reinterpret_cast<uint32_t*>(h_out + nb);
  { uint64_t k = 0;
    for (uint64_t p = 0; p < n; k++) {
      const uint32_t xlen = data[p + 10] | (data[p + 11] << 8);
      uint32_t bsize = 0;
      for (uint32_t q = 0; q + 4 <= xlen;) { const uint8_t* x = data + p + 12 + q; const uint32_t slen = x[2] | (x[3] << 8); if (x[0] == 'B' && x[1] == 'C' && slen == 2) bsize = x[4] | (x[5] << 8); q += 4 + slen; }
      const uint64_t end = p + bsize + 1;
      h_in[k] = p + 12 + xlen; h_clen[k] = (uint32_t)(end - 8 - h_in[k]); h_out[k] = total;
      total += (uint32_t)data[end - 4] | ((uint32_t)data[end - 3] << 8) | ((uint32_t)data[end - 2] << 16) | ((uint32_t)data[end - 1] << 24);
      p = end;
    } }
  ENS(c->i_in, n + 16); ENS(c->i_meta, nb * 20 + 64); ENS(c->i_out, total + 16); ENS(c->i_ctl, 16);
  cudaStream_t st = c->st;
  CU(cudaMemcpyAsync(c->i_in.p, data, n, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(c->i_meta.p, c->i_hmeta.p, nb * 20, cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync(c->i_ctl.p, 0, 16, st));
  InflateJob J; J.in = c->i_in.as<uint8_t>(); J.in_off = c->i_meta.as<uint64_t>(); J.out_off = J.in_off + nb; J.clen = reinterpret_cast<const uint32_t*>(J.out_off + nb);
  J.out = c->i_out.as<uint8_t>(); J.n_blocks = (uint32_t)nb; J.ticket = c->i_ctl.as<uint32_t>(); J.err = J.ticket + 2;
  const uint32_t ctas = std::min<uint32_t>((uint32_t)((nb + INFLATE_WARPS - 1) / INFLATE_WARPS), (uint32_t)c->n_sm * 8);
  static const bool timing = getenv("BRB_INFLATE_TIMING") != nullptr;     // kernel milliseconds on stderr
  if (timing) CU(cudaEventRecord(c->tev[0], st));
  k_bgzf_inflate<<<ctas, 32 * INFLATE_WARPS, 0, st>>>(J);
  if (timing) { CU(cudaEventRecord(c->tev[1], st)); CU(cudaEventSynchronize(c->tev[1])); float ms = 0; cudaEventElapsedTime(&ms, c->tev[0], c->tev[1]);
                fprintf(stderr, "[inflate] %llu blocks, %llu -> %llu bytes: kernel %.3f ms\n", (unsigned long long)nb, (unsigned long long)n, (unsigned long long)total, ms); }
  uint32_t h_ctl[4] = {0, 0, 0, 0};
  CU(cudaMemcpyAsync(h_ctl, c->i_ctl.p, 16, cudaMemcpyDeviceToHost, st));
  if (total) { ENS(c->i_host, total + 64); CU(cudaMemcpyAsync(c->i_host.p, c->i_out.p, total, cudaMemcpyDeviceToHost, st)); }
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  if (h_ctl[2]) {
    static const char* what[] = {"", "bad block type", "bad stored-block length", "bad code lengths", "bad symbol or distance", "output does not match ISIZE", "CRC-32 mismatch", "compressed data ran out"};
    return fail(BRB_E_ARG, "brb_bgzf_inflate: block %u: %s", h_ctl[2] - 1, what[h_ctl[3] < 8 ? h_ctl[3] : 0]);
  }
  *out = total ? c->i_host.as<uint8_t>() : (const uint8_t*)""; *out_bytes = total;
  return BRB_OK;
}

// The resident result rows change hands: a device-to-device copy (a few MB per bundle) into ONE allocation of `to`, so that `from`
// keeps its own buffers for the next bundle and nothing is allocated in steady state.
int brb_ctx_move_result(brb_ctx* from, brb_ctx* to) {
  if (!from || !to || from == to) return fail(BRB_E_ARG, "brb_ctx_move_result: two different contexts are needed");
  if (from->idx != to->idx) return fail(BRB_E_ARG, "brb_ctx_move_result: the contexts belong to different indexes");
  if (!from->have_result) return fail(BRB_E_ARG, "brb_ctx_move_result: no result resident in the source ctx");
  if (from->result_packed || from->result_ordered) return fail(BRB_E_UNSUPPORTED, "brb_ctx_move_result moves the plain row columns (no BRB_OPT_PACKED_ROWS / BRB_OPT_ORDERED_ROWS)");
  CU(cudaSetDevice(from->idx->device));
  const size_t R = from->n_rec, W = from->n_out_words; const bool sc = from->have_scores && R;
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  size_t off = 0;
  const size_t p_src = off; off += al(R * 4); const size_t p_tid = off; off += al(R * 4); const size_t p_pos = off; off += al(R * 4);
  const size_t p_str = off; off += al(R); const size_t p_fl = off; off += al(R);
  const size_t p_sim = off; off += sc ? al(R * 8) : 0; const size_t p_clip = off; off += sc ? al(R * 4) : 0;
  const size_t p_coff = off; off += al((R + 1) * 4); const size_t p_pool = off; off += al(W * 4);
  CU(cudaStreamSynchronize(to->st));                            // (a brb_emit_bgzf on the previous rows of `to` has returned, but be sure)
  ENS(to->res_slab, off + 256);
  uint8_t* d = to->res_slab.as<uint8_t>(); const RowView& v = from->rv; cudaStream_t st = from->st;
  auto cp = [&](size_t dst, const void* src, size_t n) { return (n && R) ? cudaMemcpyAsync(d + dst, src, n, cudaMemcpyDeviceToDevice, st) : cudaSuccess; };   // (no rows: nothing to read)
  CU(cp(p_src, v.src, R * 4)); CU(cp(p_tid, v.tid, R * 4)); CU(cp(p_pos, v.pos, R * 4)); CU(cp(p_str, v.strand, R)); CU(cp(p_fl, v.flags, R));
  if (sc) { CU(cp(p_sim, v.sim, R * 8)); CU(cp(p_clip, v.clip, R * 4)); }
  CU(cp(p_coff, v.coff, (R + 1) * 4)); CU(cp(p_pool, v.pool, W * 4));
  CU(cudaStreamSynchronize(st));
  to->rv = RowView{(const uint32_t*)(d + p_src), (const uint32_t*)(d + p_tid), (const uint32_t*)(d + p_pos), (const int8_t*)(d + p_str), d + p_fl,
                   sc ? (const double*)(d + p_sim) : nullptr, sc ? (const int32_t*)(d + p_clip) : nullptr, (const uint32_t*)(d + p_coff), (const uint32_t*)(d + p_pool)};
  to->n_rec = from->n_rec; to->n_out_words = from->n_out_words; to->res_n_aln = from->res_n_aln; to->have_scores = from->have_scores;
  to->result_ordered = false; to->result_packed = false; to->result_tags = false;
  to->have_result = true; to->res_foreign = true;
  from->have_result = false;
  return BRB_OK;
}

// Output records of the resident result, assembled and BGZF-compressed on the device (k_emit.cuh + k_deflate.cuh).
int brb_emit_bgzf(brb_ctx* c, const brb_params* p, const brb_emit_in* in, const uint8_t** out, uint64_t* out_bytes, uint64_t* raw_out_bytes) {
  if (!c || !p || !in || !out || !out_bytes) return fail(BRB_E_ARG, "brb_emit_bgzf: null argument");
  if (brb_device_count() <= 0) return fail(BRB_E_NO_DEVICE, "no CUDA device");
  if (!c->have_result) return fail(BRB_E_ARG, "brb_emit_bgzf: no result resident (call brb_project first)");
  if (c->result_packed || c->result_ordered) return fail(BRB_E_UNSUPPORTED, "brb_emit_bgzf needs the plain row columns (no BRB_OPT_PACKED_ROWS / BRB_OPT_ORDERED_ROWS)");
  const uint32_t N = c->res_n_aln, U = in->n_units;
  if (N && (!in->raw || !in->rec_off || !in->rec_len)) return fail(BRB_E_ARG, "brb_emit_bgzf: null record arrays");
  if (U && !in->units) return fail(BRB_E_ARG, "brb_emit_bgzf: null units");
  CU(cudaSetDevice(c->idx->device));
  *out = nullptr; *out_bytes = 0; if (raw_out_bytes) *raw_out_bytes = 0;
  if (!U) return BRB_OK;
  cudaStream_t st = c->st;
  static const bool timing = getenv("BRB_EMIT_TIMING") != nullptr;      // stage milliseconds on stderr (adds stream synchronisations)
  double t_prev = 0; std::string tline;
  auto now_ms = [] { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec / 1e6; };
  auto lap = [&](const char* what) { if (!timing) return; cudaStreamSynchronize(st); const double t = now_ms(); char b2[64]; snprintf(b2, sizeof b2, " %s %.2f", what, t - t_prev); tline += b2; t_prev = t; };
  if (timing) t_prev = now_ms();
  ENS(c->e_raw, in->raw_bytes + 16); ENS(c->e_off, (size_t)N * 8 + 16); ENS(c->e_len, (size_t)N * 4 + 16); ENS(c->e_units, (size_t)U * 16);
  ENS(c->e_plan, (size_t)N * sizeof(EmitPlan) + 16); ENS(c->e_sizes, (size_t)U * 8 + 16); ENS(c->e_offs, (size_t)(U + 1) * 8); ENS(c->e_err, 16);
  ENS(c->scan_tmp, scan_scratch_words(U) * 8);
  CU(cudaMemcpyAsync(c->e_raw.p, in->raw, in->raw_bytes, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(c->e_off.p, in->rec_off, (size_t)N * 8, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(c->e_len.p, in->rec_len, (size_t)N * 4, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(c->e_units.p, in->units, (size_t)U * 16, cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync(c->e_err.p, 0, 16, st));
  lap("alloc+h2d");
  EmitIn E{};
  E.raw = c->e_raw.as<uint8_t>(); E.raw_bytes = in->raw_bytes; E.rec_off = c->e_off.as<uint64_t>(); E.rec_len = c->e_len.as<uint32_t>(); E.n_aln = N;
  E.n_units = U; E.units = c->e_units.as<uint4>();
  const RowView& rv = c->rv; const bool sc = c->have_scores && c->n_rec;
  E.n_rows = c->n_rec; E.src = rv.src; E.tid = rv.tid; E.pos = rv.pos; E.strand = rv.strand; E.flags = rv.flags;
  E.sim = sc ? rv.sim : nullptr; E.clip = sc ? rv.clip : nullptr; E.coff = rv.coff; E.pool = rv.pool; E.long_reads = p->long_reads ? 1u : 0u;
  EmitPlan* plan = c->e_plan.as<EmitPlan>(); uint32_t* err = c->e_err.as<uint32_t>();
  unsigned long long* sizes = c->e_sizes.as<unsigned long long>(); unsigned long long* offs = c->e_offs.as<unsigned long long>();
  uint64_t L = 0;
  if (N) k_emit_plan<<<grid_for(N, 128), 128, 0, st>>>(E, plan, err);
  k_emit_sizes<<<grid_for(U, 256), 256, 0, st>>>(E, plan, sizes, err);
  exclusive_scan(sizes, offs, U, c->scan_tmp.as<unsigned long long>(), st, &L);
  struct { unsigned long long total; } h_tot{0}; uint32_t h_err[4] = {0, 0, 0, 0};
  CU(cudaMemcpyAsync(&h_tot.total, offs + U, 8, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(h_err, err, 16, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  lap("plan+sizes+scan");
  if (h_err[0] == 1) return fail(BRB_E_ARG, "brb_emit_bgzf: malformed input record (fixed fields run past block_size, or rec_off / rec_len outside raw)");
  if (h_err[0]) return fail(BRB_E_ARG, "brb_emit_bgzf: a unit names a row or an alignment that does not exist");
  const uint64_t total = h_tot.total;
  if (raw_out_bytes) *raw_out_bytes = total;
  if (!total) return BRB_OK;
  ENS(c->z_in, total + 16);
  CU(cudaMemsetAsync(c->z_in.as<uint8_t>() + (total & ~(uint64_t)3), 0, 16 - (total & 3), st));
  k_emit_write<<<std::min<uint32_t>(grid_for(U, 8), (uint32_t)c->n_sm * 16), 256, 0, st>>>(E, plan, offs, c->z_in.as<uint8_t>());
  lap("write");
  const int rc = deflate_resident(c, total, out, out_bytes);
  if (timing) { lap("deflate+d2h"); fprintf(stderr, "[emit] %u units, %llu -> %llu bytes:%s ms\n", U, (unsigned long long)total, (unsigned long long)*out_bytes, tline.c_str()); }
  return rc;
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
  Buf d_ro, d_r, d_io, d_i, d_oo, d_ol, d_o, d_so, d_s;
  int rc = 0;
  auto up = [&](Buf& b, const void* src, size_t bytes) { if (rc) return; rc = b.ensure(bytes ? bytes : 4); if (!rc && bytes && cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->st) != cudaSuccess) rc = fail(BRB_E_CUDA, "upload failed"); };
  up(d_ro, real_off, (size_t)(n + 1) * 4); up(d_r, real, nr * 4); up(d_io, ideal_off, (size_t)(n + 1) * 4); up(d_i, ideal, ni * 4);
  up(d_oo, out_off, (size_t)(n + 1) * 4);
  std::vector<uint32_t> scr_off(n + 1, 0);
  if (c->opt_merge_warp) {                        // BRB_OPT_MERGE_WARP: the warp-cooperative code (one warp per pair) instead of one thread per pair
    for (uint32_t i = 0; i < n; i++) scr_off[i + 1] = scr_off[i] + merge_warp_scratch(real_off[i + 1] - real_off[i], ideal_off[i + 1] - ideal_off[i]);
    up(d_so, scr_off.data(), (size_t)(n + 1) * 4);
    if (!rc) rc = d_s.ensure((size_t)scr_off[n] * 4 + 4);
  }
  if (!rc) rc = d_ol.ensure((size_t)n * 4);
  if (!rc) rc = d_o.ensure(std::max<size_t>(no, 1) * 4);
  if (!rc) {
    if (c->opt_merge_warp)
      k_merge_pairs_warp<<<grid_for((size_t)n * 32, 128), 128, 0, c->st>>>(n, d_ro.as<uint32_t>(), d_r.as<uint32_t>(), d_io.as<uint32_t>(), d_i.as<uint32_t>(),
                                                                          d_oo.as<uint32_t>(), d_so.as<uint32_t>(), d_ol.as<uint32_t>(), d_o.as<uint32_t>(), d_s.as<uint32_t>());
    else
      k_merge_pairs<<<grid_for(n, 128), 128, 0, c->st>>>(n, d_ro.as<uint32_t>(), d_r.as<uint32_t>(), d_io.as<uint32_t>(), d_i.as<uint32_t>(),
                                                         d_oo.as<uint32_t>(), d_ol.as<uint32_t>(), d_o.as<uint32_t>());
    cudaMemcpyAsync(out_len, d_ol.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->st);
    if (no) cudaMemcpyAsync(out, d_o.p, no * 4, cudaMemcpyDeviceToHost, c->st);
    cudaError_t e = cudaStreamSynchronize(c->st);
    if (e != cudaSuccess) rc = fail(BRB_E_CUDA, "brb_merge_cigars: %s", cudaGetErrorString(e));
  }
  d_so.release(); d_s.release();
  d_ro.release(); d_r.release(); d_io.release(); d_i.release(); d_oo.release(); d_ol.release(); d_o.release();
  return rc;
}

// Device address of g_static_scratch on `device` (made current by the caller): asked for once per device, no allocation involved.
static int static_scratch(int device, uint32_t** out) {
  static std::mutex mu; static uint32_t* addr[64] = {};
  if (device < 0 || device >= 64) return fail(BRB_E_ARG, "brb_merge_cigars_static: device %d out of range", device);
  std::lock_guard<std::mutex> lk(mu);
  if (!addr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) { cudaGetLastError(); return fail(BRB_E_CUDA, "cudaGetSymbolAddress(g_static_scratch) on device %d: %s", device, cudaGetErrorString(e)); }
    addr[device] = static_cast<uint32_t*>(p);
  }
  *out = addr[device];
  return BRB_OK;
}

int brb_merge_cigars_static(int device, void* stream, uint32_t n, uint32_t max_real_ops, uint32_t max_ideal_ops,
                            const uint32_t* d_real_off, const uint32_t* d_real, const uint32_t* d_ideal_off, const uint32_t* d_ideal,
                            const uint32_t* d_out_off, uint32_t* d_out_len, uint32_t* d_out, uint32_t* d_n_serial) {
  if (n && (!d_real_off || !d_real || !d_ideal_off || !d_ideal || !d_out_off || !d_out_len || !d_out)) return fail(BRB_E_ARG, "brb_merge_cigars_static: null argument");
  if (brb_device_count() <= device || device < 0) return fail(BRB_E_NO_DEVICE, "no CUDA device %d", device);
  const uint64_t slice = (uint64_t)max_real_ops + 4ull * max_ideal_ops + 16;                   // merge_warp_scratch of the largest pair
  if (slice > BRB_STATIC_SCRATCH_WORDS)
    return fail(BRB_E_ARG, "brb_merge_cigars_static: max_real_ops + 4 * max_ideal_ops + 16 = %llu scratch words per pair exceed the static scratch of %u words; "
                           "use brb_merge_cigars (allocating) for CIGARs of that size", (unsigned long long)slice, BRB_STATIC_SCRATCH_WORDS);
  CU(cudaSetDevice(device));
  uint32_t* base = nullptr;
  { const int rc = static_scratch(device, &base); if (rc) return rc; }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CU(cudaMemsetAsync(base, 0, 256, st));                     // the status words start at zero; the merge scratch is written before it is read
  if (n) {
    // persistent warps, one slice each: as many as the array holds (>= 1), as the pairs need, and as the device keeps resident (64 per SM)
    int n_sm = 0; CU(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device));
    const uint64_t fit = BRB_STATIC_SCRATCH_WORDS / slice, want = std::min<uint64_t>(n, (uint64_t)std::max(1, n_sm) * 64);
    const unsigned block = fit >= 4 ? 128u : 32u, wpb = block / 32;
    const unsigned ctas = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>((want + wpb - 1) / wpb, fit / wpb));
    k_merge_pairs_static<<<ctas, block, 0, st>>>(n, d_real_off, d_real, d_ideal_off, d_ideal, d_out_off, d_out_len, d_out, base + 64, (uint32_t)slice, base);
    CU(cudaGetLastError());
  }
  if (d_n_serial) { k_copy_status<<<1, 1, 0, st>>>(base, d_n_serial); CU(cudaGetLastError()); }
  return BRB_OK;
}

}  // extern "C"
#include "rescue_host.inl"
