// The one seam between the host CLI and the projection engine: the C ABI of include/bramble_b200.h.
// The product binary binds it to libbramble_b200.so (engine_cuda.cpp); the CPU oracle build used ONLY by
// tests binds the same interface to oracle/brb_port.cpp (oracle/port_engine.cpp) so that the host half
// (ingest, flush replay, record assembly, BGZF) can be checked byte-for-byte without a GPU.
#pragma once
#include <string>
#include <vector>
#include "../../../include/bramble_b200.h"

namespace brbhost {
struct EngineIf {
  virtual ~EngineIf() {}
  virtual void open(int device, const brb_index_desc& desc) = 0;                 // throws std::runtime_error
  virtual int project(const brb_params& p, const brb_batch_in& in, brb_batch_out& out) = 0;   // out valid until the next call
  virtual std::string last_error() = 0;
  // BGZF blocks of `n` bytes, compressed by the engine (callable from another thread than project()); false: not available,
  // the caller compresses on the host.  `out` stays valid until the next call.
  virtual bool bgzf_compress(const uint8_t*, size_t, const uint8_t*&, size_t&) { return false; }
  // Output records assembled AND compressed by the engine from the rows a project_slot() call left resident in `slot`
  // (brb_emit_bgzf): slots() contexts per engine, so that a bundle's rows can wait on the device for the host's flush order
  // while the next bundles are projected.  can_emit() false: the caller assembles the records itself.
  // a piece of a BGZF file (whole blocks) inflated by the engine (brb_bgzf_inflate), callable from two reader threads (worker 0 / 1);
  // false: not available or refused (malformed block) -- the caller inflates the piece itself
  virtual bool bgzf_inflate(int, const uint8_t*, size_t, std::vector<uint8_t>&) { return false; }
  virtual int slots() { return 1; }
  virtual int project_slot(int, const brb_params& p, const brb_batch_in& in, brb_batch_out& out) { return project(p, in, out); }
  virtual bool can_emit() { return false; }
  virtual void emit(int, const brb_params&, const brb_emit_in&, const uint8_t*&, size_t&) {}
};
EngineIf* make_engine();
}  // namespace brbhost
