// Product binding of EngineIf: the CUDA engine behind the C ABI.  No fallback: open() throws without a device.
#include <cstdlib>
#include <stdexcept>
#include <vector>
#include "engine_if.hpp"

namespace brbhost {
namespace {
struct CudaEngine : EngineIf {
  // ctx: the projecting context (the only one whose pools ever grow).  holders: contexts that only KEEP the rows of a projected bundle
  // (brb_ctx_move_result) while the host replays flush() for the order.  ectx: the context of the emitter thread -- record assembly
  // and BGZF scratch, one set per engine.  zctx: the writer thread's context for the plain BGZF encoder (host-assembled records).
  brb_index* ix = nullptr; brb_ctx* ctx = nullptr; brb_ctx* zctx = nullptr; brb_ctx* ectx = nullptr; brb_ctx* ictx[2] = {nullptr, nullptr};   // ictx: the reader's two inflate workers
  std::vector<brb_ctx*> holders;
  ~CudaEngine() override { for (brb_ctx* c : ictx) if (c) brb_ctx_destroy(c); for (brb_ctx* c : holders) brb_ctx_destroy(c); if (ectx) brb_ctx_destroy(ectx); if (zctx) brb_ctx_destroy(zctx); if (ctx) brb_ctx_destroy(ctx); if (ix) brb_index_destroy(ix); }
  static void need(int rc, const char* what) { if (rc != BRB_OK) throw std::runtime_error(std::string(what) + ": " + brb_last_error()); }
  void open(int device, const brb_index_desc& desc) override {
    if (brb_device_count() <= device) throw std::runtime_error("no CUDA device " + std::to_string(device) + " (bramble_b200 has no CPU path)");
    need(brb_index_create(device, &desc, &ix), "brb_index_create");
    need(brb_ctx_create(ix, &ctx), "brb_ctx_create");
    brb_ctx_set_option(ctx, BRB_OPT_FLUSH_TAGS, 1);          // HI / NH / MAPQ / primary tie-break on the device (reference flush(), src/core.cpp:237-332)
    for (int s = 0; s < slots(); s++) { brb_ctx* c = nullptr; need(brb_ctx_create(ix, &c), "brb_ctx_create"); holders.push_back(c); }
    need(brb_ctx_create(ix, &ectx), "brb_ctx_create");
  }
  int project(const brb_params& p, const brb_batch_in& in, brb_batch_out& out) override { return brb_project(ctx, &p, &in, &out); }
  std::string last_error() override { return brb_last_error(); }
  bool bgzf_compress(const uint8_t* data, size_t n, const uint8_t*& out, size_t& out_n) override {
    if (!zctx) need(brb_ctx_create(ix, &zctx), "brb_ctx_create");
    uint64_t zn = 0;
    need(brb_bgzf_compress(zctx, data, n, &out, &zn), "brb_bgzf_compress");
    out_n = (size_t)zn;
    return true;
  }
  bool bgzf_inflate(int w, const uint8_t* z, size_t n, std::vector<uint8_t>& out) override {
    if (w < 0 || w > 1) return false;
    if (!ictx[w]) need(brb_ctx_create(ix, &ictx[w]), "brb_ctx_create");
    const uint8_t* p = nullptr; uint64_t m = 0;
    if (brb_bgzf_inflate(ictx[w], z, n, &p, &m) != BRB_OK) return false;        // (a malformed block: the host decoder names it)
    out.assign(p, p + m);
    return true;
  }
  int slots() override { static const int n = [] { const char* e = getenv("BRB_SLOTS"); const int v = e ? atoi(e) : 4; return v < 1 ? 1 : v > 16 ? 16 : v; }(); return n; }
  // out stays valid (pinned staging of ctx) until the next project; the rows move to the holder
  int project_slot(int s, const brb_params& p, const brb_batch_in& in, brb_batch_out& out) override {
    const int rc = brb_project(ctx, &p, &in, &out);
    return rc ? rc : brb_ctx_move_result(ctx, holders.at((size_t)s));
  }
  bool can_emit() override { return true; }
  void emit(int s, const brb_params& p, const brb_emit_in& in, const uint8_t*& out, size_t& out_n) override {     // one caller at a time per engine
    need(brb_ctx_move_result(holders.at((size_t)s), ectx), "brb_ctx_move_result");
    uint64_t zn = 0;
    need(brb_emit_bgzf(ectx, &p, &in, &out, &zn, nullptr), "brb_emit_bgzf");
    out_n = (size_t)zn;
  }
};
}  // namespace
EngineIf* make_engine() { return new CudaEngine(); }
}  // namespace brbhost
