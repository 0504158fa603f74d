// TEST INFRASTRUCTURE (oracle shim): stub for system boost/crc.hpp (gclib/gff.h:16). Unused by bramble.
#pragma once
#include <cstddef>
#include <cstdint>
namespace boost {
struct crc_32_type {
  uint32_t v = 0;
  void process_bytes(const void *p, size_t n) {
    const unsigned char *b = (const unsigned char *)p;
    for (size_t i = 0; i < n; i++) v = v * 31u + b[i];
  }
  uint32_t checksum() const { return v; }
  void reset() { v = 0; }
};
}
