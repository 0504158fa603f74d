// Host driver of the clip-rescue phase + brb_ksw_extend (included at the end of engine.cu).
namespace {

struct KswScratch { Buf pmat, state; uint32_t warps = 0; size_t pmat_stride = 0; uint32_t state_stride = 0; };

// Sizes the per-warp scratch of k_ksw for the largest problem and launches it.
int launch_ksw(cudaStream_t st, KswProblems P, unsigned long long max_mat, unsigned long long max_tlen, KswScratch& S, uint64_t* launches) {
  uint32_t ctas = 148 * 4;
  if ((uint64_t)ctas * KSW_WARPS > P.n) ctas = (uint32_t)((P.n + KSW_WARPS - 1) / KSW_WARPS);
  if (ctas == 0) ctas = 1;
  size_t stride = ((size_t)max_mat + 127) & ~(size_t)127; if (stride == 0) stride = 128;
  const size_t budget = (size_t)24 << 30;
  while (ctas > 1 && stride * ctas * KSW_WARPS > budget) ctas /= 2;
  S.warps = ctas * KSW_WARPS; S.pmat_stride = stride;
  ENS(S.pmat, stride * S.warps);
  S.state_stride = 0;
  if (max_tlen + 1 > (unsigned long long)KSW_SH_T) {
    S.state_stride = (uint32_t)((max_tlen + 1 + 15) & ~15ull);
    ENS(S.state, (size_t)S.state_stride * 4 * S.warps);
  }
  k_ksw<<<ctas, 32 * KSW_WARPS, 0, st>>>(P, S.pmat.as<uint8_t>(), S.pmat_stride, S.state.as<int8_t>(), S.state_stride);
  (*launches)++;
  return 0;
}

struct RescueBufs { Buf qlen, tlen, tbcap, q_off, t_off, tb_off, qbuf, tbuf, tb, kout, aln_group; KswScratch ks; };
RescueBufs& rescue_bufs(brb_ctx* c) { if (!c->rb) c->rb = new RescueBufs(); return *c->rb; }
void free_rescue_bufs(brb_ctx* c) {
  if (!c->rb) return;
  RescueBufs& R = *c->rb;
  for (Buf* b : {&R.qlen, &R.tlen, &R.tbcap, &R.q_off, &R.t_off, &R.tb_off, &R.qbuf, &R.tbuf, &R.tb, &R.kout, &R.aln_group, &R.ks.pmat, &R.ks.state}) b->release();
  delete c->rb; c->rb = nullptr;
}

int run_rescue(brb_ctx* c, const brb_params& p) {
  RescueBufs& R = rescue_bufs(c);
  cudaStream_t st = c->st;
  uint64_t L = c->stats.n_launches;
  const uint32_t C = c->n_cand, slots = 2 * C;
  DevBatch& b = c->b; const DevIndex& ix = c->idx->d;
  Ctl* ctl = c->ctl.as<Ctl>();
  ENS(R.aln_group, (size_t)b.n_aln * 4);
  ENS(R.qlen, (size_t)slots * 4); ENS(R.tlen, (size_t)slots * 4); ENS(R.tbcap, (size_t)slots * 4);
  ENS(R.q_off, (size_t)(slots + 1) * 4); ENS(R.t_off, (size_t)(slots + 1) * 4); ENS(R.tb_off, (size_t)(slots + 1) * 4);
  ENS(c->r_left, (size_t)C * sizeof(RescueSeg)); ENS(c->r_right, (size_t)C * sizeof(RescueSeg));
  ENS(R.kout, (size_t)slots * sizeof(KswOut));
  ENS(c->scan_tmp, scan_scratch_words(slots) * 4);
  k_aln_group<<<grid_for(b.n_groups, 128), 128, 0, st>>>(b, R.aln_group.as<uint32_t>()); L++;
  b.aln_group = R.aln_group.as<uint32_t>();
  RescuePlan plan; plan.qlen = R.qlen.as<uint32_t>(); plan.tlen = R.tlen.as<uint32_t>(); plan.tbcap = R.tbcap.as<uint32_t>();
  k_rescue_plan<<<grid_for(slots, 128), 128, 0, st>>>(ix, b, p, C, c->c_task.as<uint32_t>(), c->c_first.as<uint32_t>(), c->c_last.as<uint32_t>(),
                                                      c->c_cap.as<uint32_t>(), plan, ctl); L++;
  exclusive_scan(plan.qlen, R.q_off.as<uint32_t>(), slots, c->scan_tmp.as<uint32_t>(), st, &L);
  exclusive_scan(plan.tlen, R.t_off.as<uint32_t>(), slots, c->scan_tmp.as<uint32_t>(), st, &L);
  exclusive_scan(plan.tbcap, R.tb_off.as<uint32_t>(), slots, c->scan_tmp.as<uint32_t>(), st, &L);
  uint32_t* hs = c->h_small.as<uint32_t>();
  CU(cudaMemcpyAsync(hs, R.q_off.as<uint32_t>() + slots, 4, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(hs + 1, R.t_off.as<uint32_t>() + slots, 4, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(hs + 2, R.tb_off.as<uint32_t>() + slots, 4, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(hs + 4, &ctl->ksw_max_mat, 16, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  const uint32_t Q = hs[0], T = hs[1], TB = hs[2];
  unsigned long long mx[2]; memcpy(mx, hs + 4, 16);
  c->rescue_bytes = ((uint64_t)Q + T + 1) / 2;
  ENS(R.qbuf, (size_t)Q + 16); ENS(R.tbuf, (size_t)T + 16); ENS(R.tb, (size_t)(TB + 4) * 4); ENS(c->rpool, (size_t)(TB + 4) * 4);
  if (Q) {
    k_rescue_gather<<<grid_for(slots, 128), 128, 0, st>>>(ix, b, p, C, c->c_task.as<uint32_t>(), c->c_first.as<uint32_t>(), c->c_last.as<uint32_t>(),
                                                          R.q_off.as<uint32_t>(), R.t_off.as<uint32_t>(), R.qbuf.as<uint8_t>(), R.tbuf.as<uint8_t>()); L++;
    KswProblems P; P.n = slots; P.q_off = R.q_off.as<uint32_t>(); P.t_off = R.t_off.as<uint32_t>(); P.tb_off = R.tb_off.as<uint32_t>();
    P.q = R.qbuf.as<uint8_t>(); P.t = R.tbuf.as<uint8_t>(); P.tb = R.tb.as<uint32_t>(); P.out = R.kout.as<KswOut>();
    int r = launch_ksw(st, P, mx[0], mx[1], R.ks, &L); if (r) return r;
  }
  k_rescue_finish<<<grid_for(slots, 128), 128, 0, st>>>(C, R.q_off.as<uint32_t>(), R.tb_off.as<uint32_t>(), R.kout.as<KswOut>(), R.tb.as<uint32_t>(),
                                                        c->rpool.as<uint32_t>(), c->r_left.as<RescueSeg>(), c->r_right.as<RescueSeg>(),
                                                        c->c_cap.as<uint32_t>(), ctl); L++;
  c->stats.n_launches = L;
  return 0;
}
}  // namespace

// K3/K4 standalone: n independent extension problems (codes 0..4), results as ksw_extz_t fields + forward CIGAR.
extern "C" int brb_ksw_extend(brb_ctx* c, uint32_t n, const uint64_t* q_off, const uint8_t* q, const uint64_t* t_off, const uint8_t* t,
                              int32_t* max, int32_t* score, int32_t* max_q, int32_t* max_t, uint8_t* zdropped, uint32_t cigar_cap,
                              uint32_t* n_cigar, uint32_t* cigar_out) {
  if (!c || (n && (!q_off || !t_off || !q || !t || !max || !score || !max_q || !max_t || !zdropped || !n_cigar))) return fail(BRB_E_ARG, "brb_ksw_extend: null argument");
  if (brb_device_count() <= 0) return fail(BRB_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(c->idx->device));
  if (!n) return BRB_OK;
  std::vector<uint32_t> qo(n + 1), to(n + 1), tbo(n + 1);
  unsigned long long max_mat = 0, max_tlen = 0;
  tbo[0] = 0;
  for (uint32_t i = 0; i <= n; i++) { qo[i] = (uint32_t)(q_off[i] - q_off[0]); to[i] = (uint32_t)(t_off[i] - t_off[0]); }
  for (uint32_t i = 0; i < n; i++) {
    const unsigned long long ql = qo[i + 1] - qo[i], tl = to[i + 1] - to[i];
    tbo[i + 1] = tbo[i] + (uint32_t)(ql + tl + 2);
    if (ql && tl) { max_mat = std::max(max_mat, (ql + tl - 1) * std::min(ql, tl)); max_tlen = std::max(max_tlen, tl); }
  }
  Buf d_qo, d_to, d_tbo, d_q, d_t, d_tb, d_out; KswScratch ks;
  int rc = 0; uint64_t L = 0;
  auto up = [&](Buf& b, const void* src, size_t bytes) { if (rc) return; rc = b.ensure(bytes ? bytes : 4); if (!rc && bytes && cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->st) != cudaSuccess) rc = fail(BRB_E_CUDA, "upload failed"); };
  up(d_qo, qo.data(), (size_t)(n + 1) * 4); up(d_to, to.data(), (size_t)(n + 1) * 4); up(d_tbo, tbo.data(), (size_t)(n + 1) * 4);
  up(d_q, q + q_off[0], qo[n]); up(d_t, t + t_off[0], to[n]);
  if (!rc) rc = d_tb.ensure((size_t)(tbo[n] + 4) * 4);
  if (!rc) rc = d_out.ensure((size_t)n * sizeof(KswOut));
  std::vector<KswOut> ho(n); std::vector<uint32_t> htb(tbo[n] + 4);
  if (!rc) {
    KswProblems P; P.n = n; P.q_off = d_qo.as<uint32_t>(); P.t_off = d_to.as<uint32_t>(); P.tb_off = d_tbo.as<uint32_t>();
    P.q = d_q.as<uint8_t>(); P.t = d_t.as<uint8_t>(); P.tb = d_tb.as<uint32_t>(); P.out = d_out.as<KswOut>();
    rc = launch_ksw(c->st, P, max_mat, max_tlen, ks, &L);
    if (!rc) {
      cudaMemcpyAsync(ho.data(), d_out.p, (size_t)n * sizeof(KswOut), cudaMemcpyDeviceToHost, c->st);
      cudaMemcpyAsync(htb.data(), d_tb.p, (size_t)tbo[n] * 4, cudaMemcpyDeviceToHost, c->st);
      cudaError_t e = cudaStreamSynchronize(c->st);
      if (e != cudaSuccess) rc = fail(BRB_E_CUDA, "brb_ksw_extend: %s", cudaGetErrorString(e));
    }
  }
  if (!rc) for (uint32_t i = 0; i < n; i++) {
    max[i] = ho[i].max; score[i] = ho[i].score; max_q[i] = ho[i].max_q; max_t[i] = ho[i].max_t; zdropped[i] = (uint8_t)ho[i].zdropped;
    n_cigar[i] = ho[i].n_cig;
    for (uint32_t k = 0; k < ho[i].n_cig && k < cigar_cap && cigar_out; k++) cigar_out[(size_t)i * cigar_cap + k] = htb[tbo[i] + ho[i].n_cig - 1 - k];   // ksw_backtrack's final reversal
  }
  d_qo.release(); d_to.release(); d_tbo.release(); d_q.release(); d_t.release(); d_tb.release(); d_out.release(); ks.pmat.release(); ks.state.release();
  return rc;
}
