// TEST INFRASTRUCTURE -- binds the host CLI's EngineIf to the CPU oracle (brb_port.cpp) so that tests can
// check the host half of the drop-in (ingest, flush replay, record assembly) against the reference binary
// without a GPU.  The product binary (bramble_b200/bramble_b200_cli) never links this file.
#include <stdexcept>
#include <vector>
#include "../bramble_b200/csrc/host/engine_if.hpp"
#include "brb_port.h"

namespace brbhost {
namespace {
struct PortEngine : EngineIf {
  const brb_index_desc* desc = nullptr; brb_port_out po{}; bool have = false; std::vector<uint32_t> aoff, acnt;
  ~PortEngine() override { if (have) brb_port_free(&po); }
  void open(int, const brb_index_desc& d) override { desc = &d; }
  int project(const brb_params& p, const brb_batch_in& in, brb_batch_out& out) override {
    if (have) { brb_port_free(&po); have = false; }
    int rc = brb_port_project(desc, &p, &in, &po); if (rc) return rc; have = true;
    out.n_rec = po.n_rec; out.n_cigar_words = po.n_cigar_words; out.group_rec_off = po.group_rec_off; out.src_aln = po.src_aln; out.tid = po.tid;
    out.pos = po.pos; out.tx_strand = po.tx_strand; out.flags = po.flags; out.sim_score = po.sim_score; out.clip_score = po.clip_score;
    out.cigar_off = po.cigar_off; out.cigar = po.cigar;   // po.cigar_off has n_rec+1 entries (dense pool)
    // unit blocks: the oracle emits in the reference's order, so the block of owner a starts where its first row is
    aoff.assign(in.n_aln + 1, 0); acnt.assign(in.n_aln + 1, 0);
    for (uint32_t r = 0; r < po.n_rec; r++) {
      if (!(po.flags[r] & BRB_REC_FIRST)) continue;
      const uint32_t a = po.src_aln[r];
      if (acnt[a] == 0) aoff[a] = r;
      acnt[a] += (po.flags[r] & BRB_REC_PAIRED) ? 2 : 1;
    }
    out.aln_rec_off = aoff.data(); out.aln_rec_cnt = acnt.data(); out.packed_rows = nullptr;
    return 0;
  }
  std::string last_error() override { return "oracle port error"; }
};
}  // namespace
EngineIf* make_engine() { return new PortEngine(); }
}  // namespace brbhost
