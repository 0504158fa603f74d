// Product binding of EngineIf: the CUDA engine behind the C ABI.  No fallback: open() throws without a device.
#include <stdexcept>
#include "engine_if.hpp"

namespace brbhost {
namespace {
struct CudaEngine : EngineIf {
  brb_index* ix = nullptr; brb_ctx* ctx = nullptr;
  ~CudaEngine() override { if (ctx) brb_ctx_destroy(ctx); if (ix) brb_index_destroy(ix); }
  void open(int device, const brb_index_desc& desc) override {
    if (brb_device_count() <= device) throw std::runtime_error("no CUDA device " + std::to_string(device) + " (bramble_b200 has no CPU path)");
    if (brb_index_create(device, &desc, &ix) != BRB_OK) throw std::runtime_error(std::string("brb_index_create: ") + brb_last_error());
    if (brb_ctx_create(ix, &ctx) != BRB_OK) throw std::runtime_error(std::string("brb_ctx_create: ") + brb_last_error());
    brb_ctx_set_option(ctx, BRB_OPT_FLUSH_TAGS, 1);          // HI / NH / MAPQ / primary tie-break on the device (reference flush(), src/core.cpp:237-332)
  }
  int project(const brb_params& p, const brb_batch_in& in, brb_batch_out& out) override { return brb_project(ctx, &p, &in, &out); }
  std::string last_error() override { return brb_last_error(); }
};
}  // namespace
EngineIf* make_engine() { return new CudaEngine(); }
}  // namespace brbhost
