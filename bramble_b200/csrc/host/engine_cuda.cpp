// Product binding of EngineIf: the CUDA engine behind the C ABI.  No fallback: open() throws without a device.
#include <stdexcept>
#include "engine_if.hpp"

namespace brbhost {
namespace {
struct CudaEngine : EngineIf {
  brb_index* ix = nullptr; brb_ctx* ctx = nullptr; brb_ctx* zctx = nullptr;       // zctx: the writer thread's context (own stream) for the BGZF encoder
  ~CudaEngine() override { if (zctx) brb_ctx_destroy(zctx); if (ctx) brb_ctx_destroy(ctx); if (ix) brb_index_destroy(ix); }
  void open(int device, const brb_index_desc& desc) override {
    if (brb_device_count() <= device) throw std::runtime_error("no CUDA device " + std::to_string(device) + " (bramble_b200 has no CPU path)");
    if (brb_index_create(device, &desc, &ix) != BRB_OK) throw std::runtime_error(std::string("brb_index_create: ") + brb_last_error());
    if (brb_ctx_create(ix, &ctx) != BRB_OK) throw std::runtime_error(std::string("brb_ctx_create: ") + brb_last_error());
    brb_ctx_set_option(ctx, BRB_OPT_FLUSH_TAGS, 1);          // HI / NH / MAPQ / primary tie-break on the device (reference flush(), src/core.cpp:237-332)
  }
  int project(const brb_params& p, const brb_batch_in& in, brb_batch_out& out) override { return brb_project(ctx, &p, &in, &out); }
  std::string last_error() override { return brb_last_error(); }
  bool bgzf_compress(const uint8_t* data, size_t n, const uint8_t*& out, size_t& out_n) override {
    if (!zctx && brb_ctx_create(ix, &zctx) != BRB_OK) throw std::runtime_error(std::string("brb_ctx_create: ") + brb_last_error());
    uint64_t zn = 0;
    if (brb_bgzf_compress(zctx, data, n, &out, &zn) != BRB_OK) throw std::runtime_error(std::string("brb_bgzf_compress: ") + brb_last_error());
    out_n = (size_t)zn;
    return true;
  }
};
}  // namespace
EngineIf* make_engine() { return new CudaEngine(); }
}  // namespace brbhost
