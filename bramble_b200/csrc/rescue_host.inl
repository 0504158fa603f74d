// Host driver of the clip-rescue phase + brb_ksw_extend (included at the end of engine.cu).
// No host round trip inside the phase: the plan kernel sorts the problems into size-class queues on the device, every
// class kernel is launched with a fixed persistent grid and takes its problems by ticket, pools have capacities that
// persist across calls (a bundle that needs more is counted, the host grows the pool and replays the whole sequence).
namespace {

struct KswClassCfg { int G, W, ctas_per_sm; };
constexpr int kKswReg = KSW_CLASSES - 1;          // classes served by the register kernel
constexpr KswClassCfg kKswCfg[kKswReg] = {{8, 2, 4}, {16, 2, 4}, {32, 2, 4}, {32, 4, 2}, {32, 8, 1}};
inline uint32_t ksw_rows_cap(int cls) { const int qmax = kKswCfg[cls].G * kKswCfg[cls].W * 2; return (uint32_t)(2 * qmax + KSW_TAIL - 1); }
inline size_t ksw_group_bytes(int cls) { return (size_t)ksw_rows_cap(cls) * kKswCfg[cls].G * kKswCfg[cls].W * 2; }

struct RescueBufs {
  Buf qlen, tlen, tb_off, queue, tb, kout, aln_group, key, rec, rep, pmat[kKswReg], big_pmat, big_seq, big_state;
  uint32_t ctas[kKswReg] = {};
  cudaStream_t side[KSW_CLASSES] = {};            // the class kernels are independent: each runs on its own stream between two events
  cudaEvent_t ev_fork = nullptr, ev_join[KSW_CLASSES] = {};
  std::vector<Buf*> all() { std::vector<Buf*> v = {&qlen, &tlen, &tb_off, &queue, &tb, &kout, &aln_group, &key, &rec, &rep, &big_pmat, &big_seq, &big_state}; for (auto& b : pmat) v.push_back(&b); return v; }
  void release() {
    for (Buf* b : all()) b->release();
    for (auto& s : side) if (s) { cudaStreamDestroy(s); s = nullptr; }
    if (ev_fork) { cudaEventDestroy(ev_fork); ev_fork = nullptr; }
    for (auto& e : ev_join) if (e) { cudaEventDestroy(e); e = nullptr; }
  }
  size_t big_mat = 0; uint32_t big_len = 0, big_state_len = 0, big_ctas = 0;     // per-warp scratch of k_ksw_big (capacities persist)
};
RescueBufs& rescue_bufs(brb_ctx* c) { if (!c->rb) c->rb = new RescueBufs(); return *c->rb; }
void free_rescue_bufs(brb_ctx* c) {
  if (!c->rb) return;
  c->rb->release();
  delete c->rb; c->rb = nullptr;
}

// Scratch of the class kernels (fixed per class and SM count) and of the big-problem kernel (sized by what the last sequence asked for).
int ensure_ksw_scratch(RescueBufs& R, int n_sm) {
  if (!R.ev_fork) {
    if (cudaEventCreateWithFlags(&R.ev_fork, cudaEventDisableTiming) != cudaSuccess) return fail(BRB_E_CUDA, "cudaEventCreate failed");
    for (int k = 0; k < KSW_CLASSES; k++)
      if (cudaStreamCreateWithFlags(&R.side[k], cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&R.ev_join[k], cudaEventDisableTiming) != cudaSuccess)
        return fail(BRB_E_CUDA, "rescue side streams: %s", cudaGetErrorString(cudaGetLastError()));
  }
  for (int k = 0; k < kKswReg; k++) {
    R.ctas[k] = (uint32_t)(n_sm * kKswCfg[k].ctas_per_sm);
    ENS(R.pmat[k], (size_t)R.ctas[k] * KSW_WARPS * (32 / kKswCfg[k].G) * ksw_group_bytes(k));
  }
  R.big_ctas = (uint32_t)n_sm;
  const size_t warps = (size_t)R.big_ctas * KSW_WARPS;
  ENS(R.big_pmat, std::max<size_t>(R.big_mat, 128) * warps); ENS(R.big_seq, (size_t)std::max<uint32_t>(R.big_len, 16) * 2 * warps);
  ENS(R.big_state, (size_t)std::max<uint32_t>(R.big_state_len, 16) * 4 * warps);
  return 0;
}

template <int G, int W>
void launch_ksw_class(cudaStream_t st, uint32_t ctas, const DevIndex& ix, const DevBatch& b, const KswSrc& S, const KswWork& K, int cls, uint32_t n_direct, Ctl* ctl, uint8_t* pmat) {
  k_ksw_group<G, W><<<ctas, 32 * KSW_WARPS, 0, st>>>(ix, b, S, K, cls, n_direct, ctl, pmat, ksw_rows_cap(cls));
}

// The DP kernels of all classes on the queues that `ctl` / K describe, side by side: the longest problems (fewest, slowest) first.
// n_direct != nullptr: explicit problem counts (brb_ksw_extend).
void launch_ksw_all(cudaStream_t st, RescueBufs& R, const DevIndex& ix, const DevBatch& b, const KswSrc& S, const KswWork& K, Ctl* ctl, const uint32_t* n_direct, uint64_t* L) {
  auto nd = [&](int k) { return n_direct ? n_direct[k] : 0u; };
  cudaEventRecord(R.ev_fork, st);
  for (int k = KSW_CLASSES - 1; k >= 0; k--) {
    if (n_direct && !nd(k)) continue;
    cudaStream_t s = R.side[k];
    cudaStreamWaitEvent(s, R.ev_fork, 0);
    if (k == KSW_BIG) {
      KswBig B; B.pmat = R.big_pmat.as<uint8_t>(); B.pmat_stride = R.big_mat; B.seq = R.big_seq.as<uint8_t>(); B.seq_stride = R.big_len;
      B.state = R.big_state.as<int8_t>(); B.state_stride = R.big_state_len;
      k_ksw_big<<<R.big_ctas, 32 * KSW_WARPS, 0, s>>>(ix, b, S, K, nd(k), ctl, B);
    } else {
      uint8_t* pm = R.pmat[k].as<uint8_t>();
      switch (k) {
        case 0: launch_ksw_class<8, 2>(s, R.ctas[0], ix, b, S, K, 0, nd(0), ctl, pm); break;
        case 1: launch_ksw_class<16, 2>(s, R.ctas[1], ix, b, S, K, 1, nd(1), ctl, pm); break;
        case 2: launch_ksw_class<32, 2>(s, R.ctas[2], ix, b, S, K, 2, nd(2), ctl, pm); break;
        case 3: launch_ksw_class<32, 4>(s, R.ctas[3], ix, b, S, K, 3, nd(3), ctl, pm); break;
        default: launch_ksw_class<32, 8>(s, R.ctas[4], ix, b, S, K, 4, nd(4), ctl, pm); break;
      }
    }
    (*L)++;
    cudaEventRecord(R.ev_join[k], s);
    cudaStreamWaitEvent(st, R.ev_join[k], 0);
  }
}

// After a sequence: did the big-problem kernel skip anything?  Then its scratch grows to what the plan kernel saw.
bool grow_ksw_big(RescueBufs& R, const Ctl* h) {
  if (!h->ksw_big_skipped) return false;
  R.big_mat = std::max<size_t>(R.big_mat, (((size_t)h->ksw_max_mat + 127) & ~(size_t)127));
  R.big_len = std::max<uint32_t>(R.big_len, (uint32_t)((std::max(h->ksw_max_qlen, h->ksw_max_tlen) + 16 + 15) & ~15ull));
  if (h->ksw_max_tlen + 1 > (unsigned long long)KSW_SH_T) R.big_state_len = std::max<uint32_t>(R.big_state_len, (uint32_t)((h->ksw_max_tlen + 1 + 15) & ~15ull));
  return true;
}

// plan -> class kernels -> finish, all on the ctx stream; arrays are sized by the pool capacities.
int launch_rescue(brb_ctx* c, const brb_params& p, uint64_t* L) {
  RescueBufs& R = rescue_bufs(c);
  cudaStream_t st = c->st;
  const size_t C = c->cap_cand, slots = 2 * C;
  DevBatch& b = c->b; const DevIndex& ix = c->idx->d;
  Ctl* ctl = c->ctl.as<Ctl>();
  ENS(R.aln_group, (size_t)b.n_aln * 4);
  ENS(R.qlen, slots * 4); ENS(R.tlen, slots * 4); ENS(R.tb_off, slots * 4); ENS(R.queue, (size_t)KSW_CLASSES * slots * 4);
  ENS(R.key, slots * 8); ENS(R.rec, slots * 4); ENS(R.rep, slots * 4);
  ENS(c->r_left, C * sizeof(RescueSeg)); ENS(c->r_right, C * sizeof(RescueSeg));
  ENS(R.kout, slots * sizeof(KswOut));
  ENS(R.tb, (size_t)c->cap_tb * 4 + 16); ENS(c->rpool, (size_t)c->cap_tb * 4 + 16);
  { int r = ensure_ksw_scratch(R, c->n_sm); if (r) return r; }
  k_aln_group<<<grid_for(b.n_groups, 128), 128, 0, st>>>(b, R.aln_group.as<uint32_t>()); (*L)++;
  b.aln_group = R.aln_group.as<uint32_t>();
  KswWork K; K.qlen = R.qlen.as<uint32_t>(); K.tlen = R.tlen.as<uint32_t>(); K.tb_off = R.tb_off.as<uint32_t>(); K.queue = R.queue.as<uint32_t>();
  K.queue_stride = (uint32_t)slots; K.tb = R.tb.as<uint32_t>(); K.out = R.kout.as<KswOut>();
  K.key = R.key.as<unsigned long long>(); K.rec = R.rec.as<uint32_t>(); K.rep = R.rep.as<uint32_t>();
  KswSrc S{}; S.c_task = c->c_task.as<uint32_t>(); S.c_first = c->c_first.as<uint32_t>(); S.c_last = c->c_last.as<uint32_t>();
  k_rescue_sizes<<<c->n_sm * 8, 256, 0, st>>>(ix, b, p, S.c_task, S.c_first, S.c_last, c->c_cap.as<uint32_t>(), K, ctl); (*L)++;
  k_rescue_plan<<<c->n_sm * 8, 256, 0, st>>>(ix, b, S.c_task, c->a_cand.as<uint4>(), K, ctl); (*L)++;
  launch_ksw_all(st, R, ix, b, S, K, ctl, nullptr, L);
  k_rescue_finish<<<c->n_sm * 8, 128, 0, st>>>(K, c->rpool.as<uint32_t>(), c->r_left.as<RescueSeg>(), c->r_right.as<RescueSeg>(), c->c_cap.as<uint32_t>(), ctl); (*L)++;
  return 0;
}
}  // namespace

// K3/K4 standalone: n independent extension problems (codes 0..4), results as ksw_extz_t fields + forward CIGAR.  The problems go
// through the same size-class kernels as the clip rescue of brb_project (a queue per class, built here on the host).
extern "C" int brb_ksw_extend(brb_ctx* c, uint32_t n, const uint64_t* q_off, const uint8_t* q, const uint64_t* t_off, const uint8_t* t,
                              int32_t* max, int32_t* score, int32_t* max_q, int32_t* max_t, uint8_t* zdropped, uint32_t cigar_cap,
                              uint32_t* n_cigar, uint32_t* cigar_out) {
  if (!c || (n && (!q_off || !t_off || !q || !t || !max || !score || !max_q || !max_t || !zdropped || !n_cigar))) return fail(BRB_E_ARG, "brb_ksw_extend: null argument");
  if (brb_device_count() <= 0) return fail(BRB_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(c->idx->device));
  if (!n) return BRB_OK;
  if (!c->n_sm) { int sm = 0; c->n_sm = (cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, c->idx->device) == cudaSuccess && sm > 0) ? sm : 148; }
  std::vector<uint32_t> qo(n + 1), to(n + 1), tbo(n + 1), queue((size_t)KSW_CLASSES * n);
  uint32_t cnt[KSW_CLASSES] = {};
  RescueBufs R;                                   // a private set: the test entry does not touch the context's rescue state
  tbo[0] = 0;
  for (uint32_t i = 0; i <= n; i++) { qo[i] = (uint32_t)(q_off[i] - q_off[0]); to[i] = (uint32_t)(t_off[i] - t_off[0]); }
  for (uint32_t i = 0; i < n; i++) {
    const unsigned long long ql = qo[i + 1] - qo[i], tl = to[i + 1] - to[i];
    tbo[i + 1] = tbo[i] + (uint32_t)(ql + tl + 2);
    const int cls = (tl > ql + KSW_TAIL) ? KSW_BIG : ksw_class_of((uint32_t)ql);      // (a rescue target is never longer than the query + 40; anything else goes to the general kernel)
    queue[(size_t)cls * n + cnt[cls]++] = i;
    if (cls == KSW_BIG && ql && tl) {
      R.big_mat = std::max<size_t>(R.big_mat, (size_t)(((ql + tl - 1) * std::min(ql, tl) + 127) & ~127ull));
      R.big_len = std::max<uint32_t>(R.big_len, (uint32_t)((std::max(ql, tl) + 16 + 15) & ~15ull));
      if (tl + 1 > (unsigned long long)KSW_SH_T) R.big_state_len = std::max<uint32_t>(R.big_state_len, (uint32_t)((tl + 1 + 15) & ~15ull));
    }
  }
  Buf d_qo, d_to, d_q, d_t, d_ctl;
  int rc = 0; uint64_t L = 0;
  auto up = [&](Buf& b, const void* src, size_t bytes) { if (rc) return; rc = b.ensure(bytes ? bytes : 4); if (!rc && bytes && cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->st) != cudaSuccess) rc = fail(BRB_E_CUDA, "upload failed"); };
  up(d_qo, qo.data(), (size_t)(n + 1) * 4); up(d_to, to.data(), (size_t)(n + 1) * 4); up(R.tb_off, tbo.data(), (size_t)(n + 1) * 4);
  up(R.queue, queue.data(), queue.size() * 4);
  up(d_q, q + q_off[0], qo[n]); up(d_t, t + t_off[0], to[n]);
  if (!rc) rc = R.tb.ensure((size_t)(tbo[n] + 4) * 4);
  if (!rc) rc = R.kout.ensure((size_t)n * sizeof(KswOut));
  if (!rc) rc = d_ctl.ensure(sizeof(Ctl));
  if (!rc) rc = ensure_ksw_scratch(R, c->n_sm);
  std::vector<KswOut> ho(n); std::vector<uint32_t> htb(tbo[n] + 4);
  if (!rc) {
    cudaMemsetAsync(d_ctl.p, 0, sizeof(Ctl), c->st);
    cudaMemsetAsync(R.kout.p, 0, (size_t)n * sizeof(KswOut), c->st);
    KswWork K{}; K.tb_off = R.tb_off.as<uint32_t>(); K.queue = R.queue.as<uint32_t>(); K.queue_stride = n; K.tb = R.tb.as<uint32_t>(); K.out = R.kout.as<KswOut>();
    KswSrc S{}; S.q = d_q.as<uint8_t>(); S.t = d_t.as<uint8_t>(); S.q_off = d_qo.as<uint32_t>(); S.t_off = d_to.as<uint32_t>();
    DevBatch nb{};
    launch_ksw_all(c->st, R, c->idx->d, nb, S, K, d_ctl.as<Ctl>(), cnt, &L);
    cudaMemcpyAsync(ho.data(), R.kout.p, (size_t)n * sizeof(KswOut), cudaMemcpyDeviceToHost, c->st);
    cudaMemcpyAsync(htb.data(), R.tb.p, (size_t)tbo[n] * 4, cudaMemcpyDeviceToHost, c->st);
    cudaError_t e = cudaStreamSynchronize(c->st);
    if (e != cudaSuccess) rc = fail(BRB_E_CUDA, "brb_ksw_extend: %s", cudaGetErrorString(e));
  }
  if (!rc) for (uint32_t i = 0; i < n; i++) {
    const bool empty = qo[i + 1] == qo[i] || to[i + 1] == to[i];          // ksw_reset_extz: nothing aligned
    if (empty) { max[i] = 0; score[i] = BRB_KSW_NEG_INF; max_q[i] = -1; max_t[i] = -1; zdropped[i] = 0; n_cigar[i] = 0; continue; }
    max[i] = ho[i].max; score[i] = ho[i].score; max_q[i] = ho[i].max_q; max_t[i] = ho[i].max_t; zdropped[i] = (uint8_t)ho[i].zdropped;
    n_cigar[i] = ho[i].n_cig;
    for (uint32_t k = 0; k < ho[i].n_cig && k < cigar_cap && cigar_out; k++) cigar_out[(size_t)i * cigar_cap + k] = htb[tbo[i] + ho[i].n_cig - 1 - k];   // ksw_backtrack's final reversal
  }
  d_qo.release(); d_to.release(); d_q.release(); d_t.release(); d_ctl.release();
  R.release();
  return rc;
}
